// Packs the host tables + flattened tree into the HBM layout described in device_scene.h.
// Compiled with -ffp-contract=off: e1, e2 and c3 must round exactly as the reference's
// per-ray expressions do (rtx.hlsl:167-186 with Math.hpp:169-175's determinant order).
#include "device_scene.h"

#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>

namespace mcpt {

namespace {
inline float as_float(uint32_t u) { float f; std::memcpy(&f, &u, 4); return f; }
inline uint32_t as_uint(float f) { uint32_t u; std::memcpy(&u, &f, 4); return u; }
}  // namespace

// Emitter list (triangle order, area CDF) and material table: the two small tables, shared by the host pack and by the device
// pack (kd_build_gpu.cu makes the large ones on the device; these stay a host loop over the triangles' materials).
static void collect_emitters(const HostScene& scene, PackedScene& out, std::vector<double>& areas, double& areaSum) {
    const size_t nT = scene.num_triangles();
    out.emitters.clear();
    areas.clear();
    areaSum = 0.0;
    // which materials emit (rtx.hlsl:330: any Ka component > 0), then the emitting triangles in triangle order: the scan over the
    // triangles' material ids runs in parallel for large scenes (static chunks, concatenated in chunk order)
    std::vector<unsigned char> emits(scene.num_materials(), 0);
    for (size_t m = 0; m < emits.size(); ++m) {
        const float* Ka = &scene.materials[12 * m];
        emits[m] = (Ka[0] > 0 || Ka[1] > 0 || Ka[2] > 0) ? 1 : 0;
    }
    std::vector<size_t> ids;
    if (nT > 65536) {
        const int chunks = 64;
        std::vector<std::vector<size_t>> part(chunks);
#pragma omp parallel for schedule(static, 1)
        for (int c = 0; c < chunks; ++c)
            for (size_t t = nT * c / chunks; t < nT * (c + 1) / chunks; ++t)
                if (emits[scene.triangles[7 * t + 6]]) part[c].push_back(t);
        for (const auto& p : part) ids.insert(ids.end(), p.begin(), p.end());
    } else {
        for (size_t t = 0; t < nT; ++t)
            if (emits[scene.triangles[7 * t + 6]]) ids.push_back(t);
    }
    for (const size_t t : ids) {
        const uint32_t* r = &scene.triangles[7 * t];
        const float* a = &scene.vertices[3 * static_cast<size_t>(r[0])];
        const float* b = &scene.vertices[3 * static_cast<size_t>(r[1])];
        const float* c = &scene.vertices[3 * static_cast<size_t>(r[2])];
        const double u[3] = {double(b[0]) - a[0], double(b[1]) - a[1], double(b[2]) - a[2]};
        const double v[3] = {double(c[0]) - a[0], double(c[1]) - a[1], double(c[2]) - a[2]};
        const double cx = u[1] * v[2] - u[2] * v[1], cy = u[2] * v[0] - u[0] * v[2], cz = u[0] * v[1] - u[1] * v[0];
        const double area = 0.5 * std::sqrt(cx * cx + cy * cy + cz * cz);
        areas.push_back(area);
        areaSum += area;
        out.emitters.push_back(make_float4(a[0], a[1], a[2], static_cast<float>(area)));
        out.emitters.push_back(make_float4(static_cast<float>(u[0]), static_cast<float>(u[1]), static_cast<float>(u[2]), 0.0f));
        out.emitters.push_back(make_float4(static_cast<float>(v[0]), static_cast<float>(v[1]), static_cast<float>(v[2]), as_float(static_cast<uint32_t>(t))));
    }
}
static void pack_small_tables_finish(const HostScene& scene, PackedScene& out, const std::vector<double>& areas, double areaSum) {
    const size_t nM = scene.num_materials();
    // running CDF over emitter area for area-uniform light sampling
    double run = 0.0;
    for (size_t i = 0; i < areas.size(); ++i) {
        run += areas[i];
        out.emitters[3 * i + 1].w = (i + 1 == areas.size()) ? 1.0f : static_cast<float>(run / areaSum);
    }
    out.emitter_area = static_cast<float>(areaSum);
    if (out.emitters.empty()) out.emitters.resize(3, make_float4(0, 0, 0, 0));

    out.mats.resize(3 * nM);
    for (size_t m = 0; m < nM; ++m) {
        const float* p = &scene.materials[12 * m];
        out.mats[3 * m] = make_float4(p[0], p[1], p[2], p[9]);       // Ka, Ns
        out.mats[3 * m + 1] = make_float4(p[3], p[4], p[5], p[10]);  // Kd, Tr
        out.mats[3 * m + 2] = make_float4(p[6], p[7], p[8], p[11]);  // Ks, Ni
    }
}
void pack_small_tables(const HostScene& scene, const HostTree& tree, PackedScene& out) {
    std::vector<double> areas;
    double areaSum = 0.0;
    collect_emitters(scene, out, areas, areaSum);
    pack_small_tables_finish(scene, out, areas, areaSum);
    for (int i = 0; i < 3; ++i) { out.bmin[i] = tree.root(i); out.bmax[i] = tree.root(3 + i); }
}

bool pack_scene(const HostScene& scene, const HostTree& tree, PackedScene& out) {
    const size_t nN = tree.num_nodes(), nT = scene.num_triangles();
    auto tp = std::chrono::steady_clock::now();
    auto lap = [&](const char* what) {
        if (!tuning().build_timing) return;
        const auto now = std::chrono::steady_clock::now();
        std::fprintf(stderr, "  pack: %-28s %7.1f ms\n", what, 1e3 * std::chrono::duration<double>(now - tp).count());
        tp = now;
    };
    // Device order of the leaves: depth first (left subtree first), so that the id lists, triangle records and
    // headers of spatially neighbouring leaves are neighbours in memory (the host tree keeps the reference's
    // breadth-first leaf order).  Matters once the scene outgrows L1/L2 (K4: L2 hit rate, DRAM page locality).
    // A leaf's depth-first ordinal is the number of leaves to its left: children are numbered after their parents
    // (breadth-first tree), so one backward pass counts the leaves and references under every node and one forward
    // pass hands each child its start -- two flat array passes instead of a stack walk that appends ids one by one.
    out.leaf_order.assign(nN, 0xFFFFFFFFu);
    std::vector<uint32_t> devOffset(nN, 0u);
    size_t totalRefs = 0;
    if (tuning().leaf_bfs) {  // A/B knob: keep the breadth-first leaf order
        uint32_t ordinal = 0, off = 0;
        for (uint32_t k = 0; k < nN; ++k)
            if (tree.split_axis[k] == 0) { out.leaf_order[k] = ordinal++; devOffset[k] = off; off += tree.leaf_count[k]; }
        totalRefs = off;
    } else {
        std::vector<uint32_t> leavesUnder(nN), refsUnder(nN);
        for (size_t k = nN; k-- > 0;) {
            if (tree.split_axis[k] == 0) { leavesUnder[k] = 1u; refsUnder[k] = tree.leaf_count[k]; }
            else { const uint32_t L = tree.left[k]; leavesUnder[k] = leavesUnder[L] + leavesUnder[L + 1]; refsUnder[k] = refsUnder[L] + refsUnder[L + 1]; }
        }
        std::vector<uint32_t> firstLeaf(nN, 0u);
        for (size_t k = 0; k < nN; ++k) {
            if (tree.split_axis[k] == 0) { out.leaf_order[k] = firstLeaf[k]; continue; }
            const uint32_t L = tree.left[k];
            firstLeaf[L] = firstLeaf[k]; devOffset[L] = devOffset[k];
            firstLeaf[L + 1] = firstLeaf[k] + leavesUnder[L]; devOffset[L + 1] = devOffset[k] + refsUnder[L];
        }
        totalRefs = nN ? refsUnder[0] : 0;
    }
    out.leaf_ids.assign(totalRefs ? totalRefs : 1, 0u);
#pragma omp parallel for schedule(static) if (nN > 65536)
    for (size_t k = 0; k < nN; ++k)
        if (tree.split_axis[k] == 0)
            for (uint32_t i = 0; i < tree.leaf_count[k]; ++i) out.leaf_ids[devOffset[k] + i] = tree.leaf_ids[tree.leaf_offset[k] + i];
    lap("leaf order (depth first)");
    // Parity-only tables (reference-order boxes, per-reference record copies) are built on first use: pack_parity_tables.
    out.ref_nodes.clear();
    out.leaf_tris.clear();
    out.kd_nodes.resize(nN);
#pragma omp parallel for schedule(static) if (nN > 65536)
    for (size_t k = 0; k < nN; ++k) {
        if (tree.split_axis[k] != 0) out.kd_nodes[k] = make_uint2(as_uint(tree.split_value[k]), (tree.left[k] << 2) | (tree.split_axis[k] - 1u));
        else out.kd_nodes[k] = make_uint2(devOffset[k], (tree.leaf_count[k] << 2) | 3u);
    }
    for (int i = 0; i < 3; ++i) { out.bmin[i] = tree.aabb[i]; out.bmax[i] = tree.aabb[3 + i]; }

    lap("node tables");
    out.tri_isect.resize(3 * nT);
    out.tri_shade.resize(3 * nT);
    double areaSum = 0.0;
    std::vector<double> areas;
#pragma omp parallel for schedule(static) if (nT > 65536)
    for (size_t t = 0; t < nT; ++t) {
        const uint32_t* r = &scene.triangles[7 * t];
        const float* a = &scene.vertices[3 * static_cast<size_t>(r[0])];
        const float* b = &scene.vertices[3 * static_cast<size_t>(r[1])];
        const float* c = &scene.vertices[3 * static_cast<size_t>(r[2])];
        const float e1[3] = {a[0] - b[0], a[1] - b[1], a[2] - b[2]};
        const float e2[3] = {a[0] - c[0], a[1] - c[1], a[2] - c[2]};
        const float c3 = e1[1] * e2[2] - e2[1] * e1[2];
        out.tri_isect[3 * t] = make_float4(a[0], a[1], a[2], c3);
        out.tri_isect[3 * t + 1] = make_float4(e1[0], e1[1], e1[2], as_float(static_cast<uint32_t>(t)));
        out.tri_isect[3 * t + 2] = make_float4(e2[0], e2[1], e2[2], 0.0f);
        const float* n0 = &scene.normals[3 * static_cast<size_t>(r[3])];
        const float* n1 = &scene.normals[3 * static_cast<size_t>(r[4])];
        const float* n2 = &scene.normals[3 * static_cast<size_t>(r[5])];
        out.tri_shade[3 * t] = make_float4(n0[0], n0[1], n0[2], as_float(r[6]));
        out.tri_shade[3 * t + 1] = make_float4(n1[0], n1[1], n1[2], 0.0f);
        out.tri_shade[3 * t + 2] = make_float4(n2[0], n2[1], n2[2], 0.0f);
    }
    lap("triangle tables");
    collect_emitters(scene, out, areas, areaSum);
    lap("emitters");
    const bool ropesOk = build_ropes(tree, out);
    lap("neighbour links + leaf blocks");

    pack_small_tables_finish(scene, out, areas, areaSum);
    return ropesOk;
}

// The tables only the parity hooks and the A/B kernels read: the reference-order boxes (ref_nodes) and the per-reference
// copies of the triangle records (leaf_tris).  Built on their first use (ensure_parity_uploaded, check_ropes' callers):
// at 1 M triangles they are 130 MB that a render never touches.
void pack_parity_tables(const HostTree& tree, PackedScene& out) {
    const size_t nN = tree.num_nodes();
    if (out.ref_nodes.size() == 2 * nN && out.leaf_tris.size() == 3 * out.leaf_ids.size()) return;
    out.ref_nodes.resize(2 * nN);
#pragma omp parallel for schedule(static) if (nN > 65536)
    for (size_t k = 0; k < nN; ++k) {
        const float* bb = &tree.aabb[6 * k];
        const bool inner = tree.split_axis[k] != 0;
        const uint32_t a = inner ? tree.left[k] : out.kd_nodes[k].x, b = inner ? 0u : (0x80000000u | tree.leaf_count[k]);
        out.ref_nodes[2 * k] = make_float4(bb[0], bb[1], bb[2], as_float(a));
        out.ref_nodes[2 * k + 1] = make_float4(bb[3], bb[4], bb[5], as_float(b));
    }
    out.leaf_tris.resize(3 * out.leaf_ids.size());
#pragma omp parallel for schedule(static) if (out.leaf_ids.size() > 65536)
    for (size_t k = 0; k < out.leaf_ids.size(); ++k)
        for (int j = 0; j < 3; ++j) out.leaf_tris[3 * k + j] = out.tri_isect[3 * static_cast<size_t>(out.leaf_ids[k]) + j];
}

// D3DXMatrixLookAtRH / Inverse / PerspectiveFovRH in closed form (GraphicsRTX.cpp:173-184):
// zaxis = normalize(eye - at), xaxis = normalize(up x zaxis), yaxis = zaxis x xaxis,
// inverse view rows = xaxis, yaxis, zaxis, eye; _22 = cot(fovy/2), _11 = _22 / (w/h).
DCamera make_camera(const float eye[3], const float at[3], const float up[3], float fovy, int width, int height, int twin) {
    DCamera cam;
    auto norm = [](float v[3]) {
        float l = sqrtf(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
        v[0] /= l; v[1] /= l; v[2] /= l;
    };
    float z[3] = {eye[0] - at[0], eye[1] - at[1], eye[2] - at[2]};
    norm(z);
    float x[3] = {up[1] * z[2] - up[2] * z[1], up[2] * z[0] - up[0] * z[2], up[0] * z[1] - up[1] * z[0]};
    norm(x);
    float y[3] = {z[1] * x[2] - z[2] * x[1], z[2] * x[0] - z[0] * x[2], z[0] * x[1] - z[1] * x[0]};
    for (int i = 0; i < 3; ++i) { cam.inv[0][i] = x[i]; cam.inv[1][i] = y[i]; cam.inv[2][i] = z[i]; cam.inv[3][i] = eye[i]; }
    float aspect = 1.0f * width / height;
    cam.p22 = static_cast<float>(1.0 / std::tan(static_cast<double>(fovy) * 0.5));
    cam.p11 = cam.p22 / aspect;
    cam.sx = static_cast<float>(2.0 / (static_cast<double>(width) * cam.p11));
    cam.sy = static_cast<float>(2.0 / (static_cast<double>(height) * cam.p22));
    cam.tan_half = tanf(fovy * 0.5f);  // tan(projFOV * PW_PI / 360), CUTracer.cu:202
    cam.twin = twin;
    return cam;
}

}  // namespace mcpt
