// C ABI, part 1: Initialize / CreateGeometry / DestroyGeometry and the scene queries.
// Replaces PW::Tracer::Initialize, CreateGeometry, DestroyGeometry (CUTracer.cu:220-338) and the
// static scene set-up of GraphicsRTX::DoOnUpdate (GraphicsRTX.cpp:165-167).
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstring>

#include "context.h"
#include "obj_loader.h"

namespace mcpt {

Context& ctx() {
    static Context c;
    return c;
}

int fail(int code, const std::string& msg) {
    ctx().last_error = msg;
    return code;
}

int fail_cuda(cudaError_t e, const char* what) {
    return fail(MCPT_ERR_CUDA, std::string(what) + ": " + cudaGetErrorString(e));
}

Tuning& tuning() {
    static Tuning t;
    return t;
}

int bind_device() {
    Context& c = ctx();
    if (!c.has_device) return fail(MCPT_ERR_NO_DEVICE, "mcpt_initialize has not bound a CUDA device; libmcpt has no CPU path");
    const cudaError_t e = cudaSetDevice(c.device);
    return e == cudaSuccess ? MCPT_OK : fail_cuda(e, "cudaSetDevice");
}

namespace {

template <class T>
cudaError_t upload(DeviceBuffer<T>& dst, const std::vector<T>& src, cudaStream_t s) {
    cudaError_t e = dst.ensure(src.size());
    if (e != cudaSuccess) return e;
    return cudaMemcpyAsync(dst.ptr, src.data(), src.size() * sizeof(T), cudaMemcpyHostToDevice, s);
}

void release_device_scene(Context& c) {
    c.d_ref_nodes.release(); c.d_kd_nodes.release(); c.d_leaf_ids.release(); c.d_tri_isect.release(); c.d_leaf_tris.release(); c.d_leaf_blob.release(); c.d_rope_nodes.release();
    c.d_tri_shade.release(); c.d_mats.release(); c.d_emitters.release();
    c.uploaded = false;
    c.parity_uploaded = false;
}

// Builds tree and device tables into locals and commits them to the context only when all of it succeeded: a
// failed call leaves the previous geometry (host tables, tree AND device tables) exactly as it was.  If the upload
// of a committed scene fails, the context holds the new host tables with nothing on the device (uploaded = false;
// the next device call retries the upload) -- never a mix of new host tables and old device tables.
// The small tables (materials, emitters) go up, the scene descriptor is bound to the device tables, residency is chosen.
int bind_device_scene(Context& c) {
    cudaError_t e;
    if ((e = upload(c.d_mats, c.packed.mats, c.stream)) != cudaSuccess) return fail_cuda(e, "upload mats");
    if ((e = upload(c.d_emitters, c.packed.emitters, c.stream)) != cudaSuccess) return fail_cuda(e, "upload emitters");
    if ((e = cudaStreamSynchronize(c.stream)) != cudaSuccess) return fail_cuda(e, "scene upload");
    c.parity_uploaded = false;
    c.dscene = DScene{};
    DScene& d = c.dscene;
    d.rope_nodes = c.d_rope_nodes.ptr; d.leaf_blob = c.d_leaf_blob.ptr; d.tri_shade = c.d_tri_shade.ptr; d.mats = c.d_mats.ptr; d.emitters = c.d_emitters.ptr;
    for (int i = 0; i < 3; ++i) { d.bmin[i] = c.packed.bmin[i]; d.bmax[i] = c.packed.bmax[i]; }
    d.num_nodes = static_cast<uint32_t>(c.tree.num_nodes());
    d.num_tris = static_cast<uint32_t>(c.scene.num_triangles());
    d.num_mats = static_cast<uint32_t>(c.scene.num_materials());
    d.num_emitters = c.packed.emitter_area > 0 ? static_cast<uint32_t>(c.packed.emitters.size() / 3) : 0u;
    d.emitter_area = c.packed.emitter_area;
    // Scenes whose render-path tables do not fit L2 are latency-bound: they render with 8 resident blocks per SM
    // (launch.cu).  A persisting-L2 window over rope_nodes was measured there too: no effect (profiles/r2_tuning.md).
    int l2 = 0;
    cudaDeviceGetAttribute(&l2, cudaDevAttrL2CacheSize, c.device);
    c.render_path_bytes = c.tree.num_nodes() * sizeof(uint2) + (c.blob_float4s + 3 * c.scene.num_triangles()) * sizeof(float4);
    c.large_scene = l2 > 0 && c.render_path_bytes > static_cast<size_t>(l2);
    c.uploaded = true;
    return MCPT_OK;
}

// Builds tree and device tables into locals and commits them to the context only when all of it succeeded: a
// failed call leaves the previous geometry (host tables, tree) exactly as it was.  If the upload
// of a committed scene fails, the context holds the new host tables with nothing on the device (uploaded = false;
// the next device call retries the upload) -- never a mix of new host tables and old device tables.
// MCPT_BUILD_GPU with the device pack writes the new render-path tables straight into the context's device buffers: the
// previous scene's device tables are given up first (uploaded = false), and come back from its host tables on the next device
// call should the build fail.
int install_scene(HostScene&& scene, int build_mode) {
    Context& c = ctx();
    std::string err;
    if (build_mode != MCPT_BUILD_REFERENCE && build_mode != MCPT_BUILD_SAH && build_mode != MCPT_BUILD_GPU) return fail(MCPT_ERR_INVALID, "unknown build_mode");
    if (!validate_scene(scene, err)) return fail(MCPT_ERR_INVALID, err);
    HostTree tree;
    PackedScene packed;
    DeviceTables dt;
    bool devicePacked = false;
    const auto tStart = std::chrono::steady_clock::now();
    if (build_mode == MCPT_BUILD_GPU) {
        if (const int rc = bind_device()) return rc;   // the device builder has no host twin to fall back on: MCPT_BUILD_SAH is the host one
        const bool wantPack = tuning().device_pack != 0 && !tuning().leaf_bfs;
        if (wantPack) {
            dt.leaf_blob = &c.d_leaf_blob; dt.rope_nodes = &c.d_rope_nodes; dt.tri_shade = &c.d_tri_shade;
            dt.tree_words = &c.d_tree_words; dt.tree_floats = &c.d_tree_floats;   // the tree itself stays on the device until the host asks (ensure_host_tree)
            c.uploaded = false; c.parity_uploaded = false;
        }
        if (!build_gpu_tree(scene, tree, err, wantPack ? &dt : nullptr)) return fail(err.rfind("scene too large", 0) == 0 ? MCPT_ERR_INVALID : MCPT_ERR_CUDA, err);
        devicePacked = wantPack && dt.done;
    } else {
        build_kd_tree(scene, build_mode, tree);
    }
    const auto tTree = std::chrono::steady_clock::now();
    if (devicePacked) {
        pack_small_tables(scene, tree, packed);
    } else {
        if (!pack_scene(scene, tree, packed)) return fail(MCPT_ERR_INVALID, "scene too large: more than 4 GiB of leaf blocks (28-bit leaf handles)");
        if (tuning().verify_layout && !check_ropes(tree, packed, err)) return fail(MCPT_ERR_INTERNAL, "stackless layout: " + err);
    }
    const auto tPack = std::chrono::steady_clock::now();
    c.scene = std::move(scene);
    c.tree = std::move(tree);
    c.packed = std::move(packed);
    c.host_packed = !devicePacked;
    c.has_geometry = true;
    c.uploaded = false;
    if (c.tree.on_host && c.has_device && (c.d_tree_words.ptr || c.d_tree_floats.ptr)) {   // a previous device-built tree nobody asked for
        cudaSetDevice(c.device);
        c.d_tree_words.release(); c.d_tree_floats.release();
    }
    int rc = MCPT_OK;
    if (devicePacked) {
        c.blob_float4s = dt.blob_float4s;
        rc = bind_device_scene(c);
        if (rc == MCPT_OK && tuning().verify_layout) {   // the host's own pack of the same tree, checked and compared with what the device made
            uint64_t bad[3];
            if ((rc = ensure_host_packed()) == MCPT_OK && !check_ropes(c.tree, c.packed, err)) rc = fail(MCPT_ERR_INTERNAL, "stackless layout: " + err);
            if (rc == MCPT_OK && (rc = compare_device_tables(bad)) == MCPT_OK && (bad[0] | bad[1] | bad[2]))
                rc = fail(MCPT_ERR_INTERNAL, "device pack differs from the host pack");
        }
        if (rc != MCPT_OK) { cudaSetDevice(c.device); release_device_scene(c); }
    } else if (c.has_device) {
        rc = ensure_uploaded();
        if (rc != MCPT_OK) { cudaSetDevice(c.device); release_device_scene(c); }
    }
    if (tuning().build_timing) {
        auto ms = [](std::chrono::steady_clock::time_point a, std::chrono::steady_clock::time_point b) { return 1e3 * std::chrono::duration<double>(b - a).count(); };
        std::fprintf(stderr, "install: tree%s %.1f ms, %s %.1f ms, upload %.1f ms\n", devicePacked ? " + device pack" : "", ms(tStart, tTree),
                     devicePacked ? "small tables" : "pack + neighbour links", ms(tTree, tPack), ms(tPack, std::chrono::steady_clock::now()));
    }
    return rc;
}

}  // namespace

// Host tables -> HBM, one H2D copy per table on the context stream.  Two groups: what the render path reads (neighbour-link
// nodes, leaf blocks, shading and material tables: 195 of 396 MB at 1 M triangles) goes up with the geometry; what only the
// parity hooks and the A/B kernels read (reference-order boxes, stack-traversal nodes, id lists, per-triangle records) goes up
// on their first use (ensure_parity_uploaded).
int ensure_parity_uploaded() {
    Context& c = ctx();
    if (const int rc = ensure_uploaded()) return rc;
    if (c.parity_uploaded) return MCPT_OK;
    if (const int rc = ensure_host_packed()) return rc;
    pack_parity_tables(c.tree, c.packed);
    cudaError_t e;
    if ((e = upload(c.d_ref_nodes, c.packed.ref_nodes, c.stream)) != cudaSuccess) return fail_cuda(e, "upload ref_nodes");
    if ((e = upload(c.d_kd_nodes, c.packed.kd_nodes, c.stream)) != cudaSuccess) return fail_cuda(e, "upload kd_nodes");
    if ((e = upload(c.d_leaf_ids, c.packed.leaf_ids, c.stream)) != cudaSuccess) return fail_cuda(e, "upload leaf_ids");
    if ((e = upload(c.d_tri_isect, c.packed.tri_isect, c.stream)) != cudaSuccess) return fail_cuda(e, "upload tri_isect");
    if ((e = upload(c.d_leaf_tris, c.packed.leaf_tris, c.stream)) != cudaSuccess) return fail_cuda(e, "upload leaf_tris");
    if ((e = cudaStreamSynchronize(c.stream)) != cudaSuccess) return fail_cuda(e, "parity table upload");
    DScene& d = c.dscene;
    d.ref_nodes = c.d_ref_nodes.ptr; d.kd_nodes = c.d_kd_nodes.ptr; d.leaf_ids = c.d_leaf_ids.ptr; d.tri_isect = c.d_tri_isect.ptr; d.leaf_tris = c.d_leaf_tris.ptr;
    c.parity_uploaded = true;
    return MCPT_OK;
}

int ensure_host_tree() {
    Context& c = ctx();
    HostTree& t = c.tree;
    if (t.on_host) return MCPT_OK;
    if (const int rc = bind_device()) return rc;
    const size_t n = t.nodes, r = t.refs;
    if (c.d_tree_words.count < 4 * n + r || c.d_tree_floats.count < 7 * n || !c.d_tree_words.ptr || !c.d_tree_floats.ptr)
        return fail(MCPT_ERR_INTERNAL, "the device-built tree is gone from the device before the host copied it");
    t.left.resize(n); t.aabb.resize(6 * n); t.split_axis.resize(n); t.split_value.resize(n); t.leaf_offset.resize(n); t.leaf_count.resize(n); t.leaf_ids.resize(r);
    const uint32_t* w = c.d_tree_words.ptr;
    const float* f = c.d_tree_floats.ptr;
    cudaError_t e = cudaMemcpy(t.left.data(), w, n * 4, cudaMemcpyDeviceToHost);
    if (e == cudaSuccess) e = cudaMemcpy(t.split_axis.data(), w + n, n * 4, cudaMemcpyDeviceToHost);
    if (e == cudaSuccess) e = cudaMemcpy(t.leaf_offset.data(), w + 2 * n, n * 4, cudaMemcpyDeviceToHost);
    if (e == cudaSuccess) e = cudaMemcpy(t.leaf_count.data(), w + 3 * n, n * 4, cudaMemcpyDeviceToHost);
    if (e == cudaSuccess && r) e = cudaMemcpy(t.leaf_ids.data(), w + 4 * n, r * 4, cudaMemcpyDeviceToHost);
    if (e == cudaSuccess) e = cudaMemcpy(t.aabb.data(), f, n * 24, cudaMemcpyDeviceToHost);
    if (e == cudaSuccess) e = cudaMemcpy(t.split_value.data(), f + 6 * n, n * 4, cudaMemcpyDeviceToHost);
    if (e != cudaSuccess) {
        t.left.clear(); t.aabb.clear(); t.split_axis.clear(); t.split_value.clear(); t.leaf_offset.clear(); t.leaf_count.clear(); t.leaf_ids.clear();
        return fail_cuda(e, "tree download");
    }
    t.on_host = true;
    c.d_tree_words.release(); c.d_tree_floats.release();
    return MCPT_OK;
}

int ensure_host_packed() {
    Context& c = ctx();
    if (c.host_packed) return MCPT_OK;
    if (!c.has_geometry) return fail(MCPT_ERR_INVALID, "no geometry: call mcpt_create_geometry / mcpt_load_obj first");
    if (const int rc = ensure_host_tree()) return rc;
    PackedScene packed;
    if (!pack_scene(c.scene, c.tree, packed)) return fail(MCPT_ERR_INVALID, "scene too large: more than 4 GiB of leaf blocks (28-bit leaf handles)");
    c.packed = std::move(packed);
    c.host_packed = true;
    return MCPT_OK;
}

int ensure_uploaded() {
    Context& c = ctx();
    if (const int rc = bind_device()) return rc;
    if (!c.has_geometry) return fail(MCPT_ERR_INVALID, "no geometry: call mcpt_create_geometry / mcpt_load_obj first");
    if (c.uploaded) return MCPT_OK;
    if (const int rc = ensure_host_packed()) return rc;
    cudaError_t e;
    if ((e = upload(c.d_leaf_blob, c.packed.leaf_blob, c.stream)) != cudaSuccess) return fail_cuda(e, "upload leaf_blob");
    if ((e = upload(c.d_rope_nodes, c.packed.rope_nodes, c.stream)) != cudaSuccess) return fail_cuda(e, "upload rope_nodes");
    if ((e = upload(c.d_tri_shade, c.packed.tri_shade, c.stream)) != cudaSuccess) return fail_cuda(e, "upload tri_shade");
    c.blob_float4s = c.packed.leaf_blob.size();
    return bind_device_scene(c);
}

// Diagnostic: the device's render-path tables against the host's own pack of the same tree, word for word.
// bad = {leaf_blob float4s, rope_nodes, tri_shade float4s} that differ.
int compare_device_tables(uint64_t bad[3]) {
    Context& c = ctx();
    bad[0] = bad[1] = bad[2] = 0;
    if (const int rc = ensure_uploaded()) return rc;
    if (const int rc = ensure_host_packed()) return rc;
    if (c.blob_float4s != c.packed.leaf_blob.size()) return fail(MCPT_ERR_INTERNAL, "leaf_blob size differs between the device and the host pack");
    auto diff = [&](const void* dev, const void* host, size_t bytes, size_t unit, uint64_t& out) -> int {
        std::vector<unsigned char> tmp(bytes);
        const cudaError_t e = cudaMemcpy(tmp.data(), dev, bytes, cudaMemcpyDeviceToHost);
        if (e != cudaSuccess) return fail_cuda(e, "table download");
        const unsigned char* h = static_cast<const unsigned char*>(host);
        for (size_t i = 0; i < bytes; i += unit) out += std::memcmp(tmp.data() + i, h + i, unit) != 0 ? 1u : 0u;
        return MCPT_OK;
    };
    int rc;
    if ((rc = diff(c.d_leaf_blob.ptr, c.packed.leaf_blob.data(), c.packed.leaf_blob.size() * 16, 16, bad[0]))) return rc;
    if ((rc = diff(c.d_rope_nodes.ptr, c.packed.rope_nodes.data(), c.packed.rope_nodes.size() * 8, 8, bad[1]))) return rc;
    if ((rc = diff(c.d_tri_shade.ptr, c.packed.tri_shade.data(), c.packed.tri_shade.size() * 16, 16, bad[2]))) return rc;
    return MCPT_OK;
}

}  // namespace mcpt

using namespace mcpt;

extern "C" {

int mcpt_initialize(int device) {
    return guarded([&]() -> int {
    Context& c = ctx();
    std::lock_guard<std::mutex> g(c.lock);
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) {
        cudaGetLastError();
        return fail(MCPT_ERR_NO_DEVICE, std::string("no CUDA device visible (") + (e != cudaSuccess ? cudaGetErrorString(e) : "count = 0") +
                                            "); libmcpt is GPU-only and has no CPU fallback");
    }
    if (device < 0 || device >= n) return fail(MCPT_ERR_INVALID, "device index out of range");
    if (c.has_device && c.device == device) return MCPT_OK;
    if (c.has_device) return fail(MCPT_ERR_INVALID, "already initialised on another device; call mcpt_shutdown first");
    if ((e = cudaSetDevice(device)) != cudaSuccess) return fail_cuda(e, "cudaSetDevice");
    if ((e = cudaStreamCreateWithFlags(&c.stream, cudaStreamNonBlocking)) != cudaSuccess) return fail_cuda(e, "cudaStreamCreate");
    for (auto& ev : c.ev)
        if ((e = cudaEventCreate(&ev)) != cudaSuccess) return fail_cuda(e, "cudaEventCreate");
    c.has_device = true;
    c.device = device;
    cudaDeviceGetAttribute(&c.sm_count, cudaDevAttrMultiProcessorCount, device);
    if (c.has_geometry) return ensure_uploaded();
    return MCPT_OK;
    });
}

int mcpt_shutdown(void) {
    return guarded([&]() -> int {
    Context& c = ctx();
    std::lock_guard<std::mutex> g(c.lock);
    if (c.has_device) {
        cudaSetDevice(c.device);
        release_device_scene(c);
        c.d_tree_words.release(); c.d_tree_floats.release();
        release_gpu_builder_scratch();
        c.d_slices.release(); c.d_slices_sq.release(); c.d_sum.release(); c.d_sumsq.release(); c.d_rgb.release(); c.d_pool.release();
        c.d_frame_seeds.release(); c.d_counter.release(); c.d_rays.release(); c.d_t.release(); c.d_beta.release();
        c.d_gamma.release(); c.d_tri.release();
        for (auto& ev : c.ev) { if (ev) cudaEventDestroy(ev); ev = nullptr; }
        if (c.stream) cudaStreamDestroy(c.stream);
        c.stream = nullptr;
    }
    c.has_device = false;
    c.device = -1;
    c.has_geometry = false; c.host_packed = false;
    c.scene = HostScene(); c.tree = HostTree(); c.packed = PackedScene();
    return MCPT_OK;
    });
}

int mcpt_create_geometry(const mcpt_scene_desc* d) {
    return guarded([&]() -> int {
    Context& c = ctx();
    std::lock_guard<std::mutex> g(c.lock);
    if (!d || !d->vertices || !d->triangles || !d->materials || d->num_vertices <= 0 || d->num_triangles <= 0 ||
        d->num_materials <= 0 || d->num_normals < 0 || (d->num_normals > 0 && !d->normals))
        return fail(MCPT_ERR_INVALID, "mcpt_create_geometry: null or empty scene table");
    HostScene s;
    s.vertices.assign(d->vertices, d->vertices + 3 * static_cast<size_t>(d->num_vertices));
    if (d->num_normals > 0) s.normals.assign(d->normals, d->normals + 3 * static_cast<size_t>(d->num_normals));
    s.triangles.assign(d->triangles, d->triangles + 7 * static_cast<size_t>(d->num_triangles));
    s.materials.assign(d->materials, d->materials + 12 * static_cast<size_t>(d->num_materials));
    return install_scene(std::move(s), d->build_mode);
    });
}

int mcpt_load_obj(const char* obj_path, const char* mtl_base, int loader, int build_mode) {
    return guarded([&]() -> int {
    Context& c = ctx();
    std::lock_guard<std::mutex> g(c.lock);
    if (!obj_path) return fail(MCPT_ERR_INVALID, "mcpt_load_obj: null path");
    HostScene s;
    std::string err;
    if (!load_obj_scene(obj_path, mtl_base ? mtl_base : "", loader, s, err)) return fail(MCPT_ERR_IO, err);
    return install_scene(std::move(s), build_mode);
    });
}

int mcpt_load_scene_file(const char* path, int build_mode) {
    return guarded([&]() -> int {
    Context& c = ctx();
    std::lock_guard<std::mutex> g(c.lock);
    if (!path) return fail(MCPT_ERR_INVALID, "mcpt_load_scene_file: null path");
    HostScene s;
    std::string err;
    if (!load_scene_file(path, s, err)) return fail(MCPT_ERR_IO, err);
    return install_scene(std::move(s), build_mode);
    });
}

int mcpt_destroy_geometry(void) {
    return guarded([&]() -> int {
    Context& c = ctx();
    std::lock_guard<std::mutex> g(c.lock);
    if (c.has_device) { cudaSetDevice(c.device); release_device_scene(c); c.d_tree_words.release(); c.d_tree_floats.release(); }
    c.has_geometry = false; c.host_packed = false;
    c.scene = HostScene(); c.tree = HostTree(); c.packed = PackedScene();
    return MCPT_OK;
    });
}

int mcpt_check_stackless_layout(void) {
    return guarded([&]() -> int {
    Context& c = ctx();
    std::lock_guard<std::mutex> g(c.lock);
    if (!c.has_geometry) return fail(MCPT_ERR_INVALID, "no geometry");
    if (const int rc = ensure_host_packed()) return rc;
    std::string err;
    if (!check_ropes(c.tree, c.packed, err)) return fail(MCPT_ERR_INTERNAL, "stackless layout: " + err);
    return MCPT_OK;
    });
}

int mcpt_debug_compare_device_tables(uint64_t differing[3]) {
    return guarded([&]() -> int {
    Context& c = ctx();
    std::lock_guard<std::mutex> g(c.lock);
    if (!differing) return fail(MCPT_ERR_INVALID, "null result");
    return compare_device_tables(differing);
    });
}

int mcpt_debug_stackless_walk(const float* ray6, int max_hops, int* hops, uint32_t* leaf_offsets, float* exit_t, uint32_t* exit_axis,
                              uint32_t* tri_ids, int max_ids, int* num_ids) {
    return guarded([&]() -> int {
    Context& c = ctx();
    std::lock_guard<std::mutex> g(c.lock);
    if (!c.has_geometry) return fail(MCPT_ERR_INVALID, "no geometry");
    if (!ray6 || !hops || !leaf_offsets || !exit_t || !exit_axis || max_hops <= 0) return fail(MCPT_ERR_INVALID, "bad arguments");
    if (const int rc = ensure_host_packed()) return rc;
    *hops = walk_ropes_host(c.packed, ray6, max_hops, leaf_offsets, exit_t, exit_axis);
    int n = 0;
    if (tri_ids && num_ids) {   // triangle ids of the visited leaves, -1 between leaves
        for (int h = 0; h < *hops; ++h) {
            const uint32_t off = leaf_offsets[h];
            uint32_t count;
            std::memcpy(&count, &c.packed.leaf_blob[off + 1].w, 4);
            for (uint32_t k = 0; k < count && n < max_ids; ++k) {
                uint32_t id;
                std::memcpy(&id, &c.packed.leaf_blob[off + 4 + 3 * k + 1].w, 4);
                tri_ids[n++] = id;
            }
            if (n < max_ids) tri_ids[n++] = 0xFFFFFFFFu;
        }
        *num_ids = n;
    }
    return MCPT_OK;
    });
}

int mcpt_get_tree_info(mcpt_tree_info* info) {
    return guarded([&]() -> int {
    Context& c = ctx();
    std::lock_guard<std::mutex> g(c.lock);
    if (!info) return fail(MCPT_ERR_INVALID, "null info");
    if (!c.has_geometry) return fail(MCPT_ERR_INVALID, "no geometry");
    info->num_nodes = static_cast<int32_t>(c.tree.num_nodes());
    info->num_leaves = static_cast<int32_t>(c.tree.num_leaves);
    info->num_leaf_refs = static_cast<int32_t>(c.tree.num_refs());
    info->max_depth = static_cast<int32_t>(c.tree.max_depth);
    info->num_emitter_triangles = c.packed.emitter_area > 0 ? static_cast<int32_t>(c.packed.emitters.size() / 3) : 0;
    info->build_seconds = c.tree.build_seconds;
    // what is resident for rendering (+ the parity-only tables once something asked for them); before the first upload: the host pack's size
    info->device_bytes = c.uploaded ? c.render_path_bytes + (c.packed.mats.size() + c.packed.emitters.size()) * sizeof(float4) +
                                          (c.parity_uploaded ? c.packed.ref_nodes.size() * 16 + c.packed.kd_nodes.size() * 8 + c.packed.leaf_ids.size() * 4 +
                                                                   (c.packed.tri_isect.size() + c.packed.leaf_tris.size()) * 16 : 0)
                                    : c.packed.bytes();
    return MCPT_OK;
    });
}

int mcpt_get_tree(uint32_t* left, uint32_t* right, float* aabb6, uint32_t* split_axis, float* split_value,
                  uint32_t* leaf_offset, uint32_t* leaf_count, uint32_t* leaf_ids) {
    return guarded([&]() -> int {
    Context& c = ctx();
    std::lock_guard<std::mutex> g(c.lock);
    if (!c.has_geometry) return fail(MCPT_ERR_INVALID, "no geometry");
    if (const int rc = ensure_host_tree()) return rc;
    const HostTree& t = c.tree;
    const size_t n = t.num_nodes();
    if (left) std::memcpy(left, t.left.data(), n * 4);
    if (right) for (size_t k = 0; k < n; ++k) right[k] = t.split_axis[k] ? t.left[k] + 1u : 0xFFFFFFFFu;
    if (aabb6) std::memcpy(aabb6, t.aabb.data(), n * 24);
    if (split_axis) std::memcpy(split_axis, t.split_axis.data(), n * 4);
    if (split_value) std::memcpy(split_value, t.split_value.data(), n * 4);
    if (leaf_offset) std::memcpy(leaf_offset, t.leaf_offset.data(), n * 4);
    if (leaf_count) std::memcpy(leaf_count, t.leaf_count.data(), n * 4);
    if (leaf_ids) std::memcpy(leaf_ids, t.leaf_ids.data(), t.leaf_ids.size() * 4);
    return MCPT_OK;
    });
}

int mcpt_get_scene(int32_t counts[4], float* vertices, float* normals, uint32_t* triangles, float* materials) {
    return guarded([&]() -> int {
    Context& c = ctx();
    std::lock_guard<std::mutex> g(c.lock);
    if (!c.has_geometry) return fail(MCPT_ERR_INVALID, "no geometry");
    const HostScene& s = c.scene;
    if (counts) {
        counts[0] = static_cast<int32_t>(s.num_vertices()); counts[1] = static_cast<int32_t>(s.num_normals());
        counts[2] = static_cast<int32_t>(s.num_triangles()); counts[3] = static_cast<int32_t>(s.num_materials());
    }
    if (vertices) std::memcpy(vertices, s.vertices.data(), s.vertices.size() * 4);
    if (normals) std::memcpy(normals, s.normals.data(), s.normals.size() * 4);
    if (triangles) std::memcpy(triangles, s.triangles.data(), s.triangles.size() * 4);
    if (materials) std::memcpy(materials, s.materials.data(), s.materials.size() * 4);
    return MCPT_OK;
    });
}

int mcpt_set_tuning(const char* name, double value) {
    return guarded([&]() -> int {
    Context& c = ctx();
    std::lock_guard<std::mutex> g(c.lock);
    if (!name) return fail(MCPT_ERR_INVALID, "null tuning name");
    Tuning& t = tuning();
    const std::string k(name);
    auto clampi = [&](int lo, int hi) { return static_cast<int>(value < lo ? lo : (value > hi ? hi : value)); };
    auto clampf = [&](float lo, float hi) { return static_cast<float>(value < lo ? lo : (value > hi ? hi : value)); };
    if (!(value == value)) return fail(MCPT_ERR_INVALID, "tuning value is NaN");
    if (k == "shade_threshold") t.shade_threshold = clampi(1, 32);
    else if (k == "cam_threshold") t.cam_threshold = clampi(1, 32);
    else if (k == "tri_cap") t.tri_cap = clampi(1, 64);
    else if (k == "chunks") t.chunks = clampi(0, 64);
    else if (k == "min_blocks") t.min_blocks = clampi(0, 16);
    else if (k == "pool") t.pool = clampi(0, 1);
    else if (k == "variant") t.variant = clampi(0, 4095);
    else if (k == "leaf_bfs") t.leaf_bfs = clampi(0, 1);
    else if (k == "sah_resort") t.sah_resort = clampi(0, 1);
    else if (k == "build_timing") t.build_timing = clampi(0, 1);
    else if (k == "sm_stats") t.sm_stats = clampi(0, 1);
    else if (k == "verify_layout") t.verify_layout = clampi(0, 1);
    else if (k == "device_pack") t.device_pack = clampi(0, 1);
    else if (k == "sah_ci") t.sah_ci = clampf(0.05f, 100.0f);
    else if (k == "sah_eb") t.sah_eb = clampf(0.05f, 100.0f);
    else if (k == "defaults") t = Tuning();
    else return fail(MCPT_ERR_INVALID, "unknown tuning knob: " + k);
    return MCPT_OK;
    });
}

int mcpt_get_tuning(const char* name, double* value) {
    return guarded([&]() -> int {
    Context& c = ctx();
    std::lock_guard<std::mutex> g(c.lock);
    if (!name || !value) return fail(MCPT_ERR_INVALID, "null argument");
    const Tuning& t = tuning();
    const std::string k(name);
    if (k == "shade_threshold") *value = t.shade_threshold;
    else if (k == "cam_threshold") *value = t.cam_threshold;
    else if (k == "tri_cap") *value = t.tri_cap;
    else if (k == "chunks") *value = t.chunks;
    else if (k == "min_blocks") *value = t.min_blocks;
    else if (k == "pool") *value = t.pool;
    else if (k == "variant") *value = t.variant;
    else if (k == "leaf_bfs") *value = t.leaf_bfs;
    else if (k == "sah_resort") *value = t.sah_resort;
    else if (k == "build_timing") *value = t.build_timing;
    else if (k == "sm_stats") *value = t.sm_stats;
    else if (k == "verify_layout") *value = t.verify_layout;
    else if (k == "device_pack") *value = t.device_pack;
    else if (k == "sah_ci") *value = t.sah_ci;
    else if (k == "sah_eb") *value = t.sah_eb;
    else return fail(MCPT_ERR_INVALID, "unknown tuning knob: " + k);
    return MCPT_OK;
    });
}

const char* mcpt_last_error(void) { return ctx().last_error.c_str(); }
const char* mcpt_version(void) { return "mcpt-b200 0.1 (sm_100a)"; }

}  // extern "C"
