// Process-global tracer state behind the C ABI (the reference keeps module-level __device__
// globals, CUTracer.cu:30-42; here it is one struct so that DestroyGeometry really frees).
#pragma once
#include <cuda_runtime.h>

#include <exception>
#include <mutex>
#include <new>
#include <string>

#include "../../include/mcpt.h"
#include "device_scene.h"
#include "host_scene.h"

namespace mcpt {

template <class T>
struct DeviceBuffer {
    T* ptr = nullptr;
    size_t count = 0;
    cudaError_t ensure(size_t n) {
        if (n <= count && ptr) return cudaSuccess;
        release();
        cudaError_t e = cudaMalloc(reinterpret_cast<void**>(&ptr), (n ? n : 1) * sizeof(T));
        if (e == cudaSuccess) count = n;
        return e;
    }
    void release() {
        if (ptr) cudaFree(ptr);
        ptr = nullptr;
        count = 0;
    }
};

// kd_build_gpu.cu's device pack writes the render-path tables straight into these buffers of the context.
struct DeviceTables {
    DeviceBuffer<float4>* leaf_blob = nullptr;
    DeviceBuffer<uint2>* rope_nodes = nullptr;
    DeviceBuffer<float4>* tri_shade = nullptr;
    DeviceBuffer<uint32_t>* tree_words = nullptr;   // optional: the tree itself stays on the device too, [left | axis | leaf offset | leaf count] x nodes, then the leaf ids
    DeviceBuffer<float>* tree_floats = nullptr;     //           [box 6 x nodes | plane x nodes]; the HostTree then carries counts only (on_host = false)
    size_t blob_float4s = 0;   // out: float4 units of leaf_blob in use (blocks + 3 of padding)
    bool done = false;         // out: the three tables are complete on the device
};

struct Context {
    std::mutex lock;
    bool has_device = false;
    int device = -1;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};

    bool has_geometry = false, uploaded = false, parity_uploaded = false;   // render-path tables / parity-hook tables in HBM
    bool host_packed = false;     // `packed` holds the host copy of the tables (not after a device pack, until something asks for it)
    HostScene scene;
    HostTree tree;
    PackedScene packed;
    DScene dscene{};
    size_t blob_float4s = 0, render_path_bytes = 0;   // leaf_blob in use on the device; bytes of the tables a render reads
    bool large_scene = false;     // render-path tables larger than L2: latency-bound, rendered with 8 resident blocks per SM
    DeviceBuffer<float4> d_ref_nodes, d_tri_isect, d_leaf_tris, d_leaf_blob, d_tri_shade, d_mats, d_emitters;
    DeviceBuffer<uint2> d_kd_nodes, d_rope_nodes;
    DeviceBuffer<uint32_t> d_leaf_ids, d_tree_words;   // d_tree_words / d_tree_floats: a device-built tree until the host asks for it
    DeviceBuffer<float> d_tree_floats;

    // render scratch, grown on demand and kept
    DeviceBuffer<float4> d_slices, d_slices_sq, d_sum, d_sumsq;
    DeviceBuffer<float> d_rgb, d_pool;   // d_pool: path state of render_kernel_pool's resident warps
    int sm_count = 0;
    DeviceBuffer<uint32_t> d_frame_seeds;
    DeviceBuffer<unsigned long long> d_counter;  // >= 4 entries
    // intersect scratch
    DeviceBuffer<float> d_rays, d_t, d_beta, d_gamma;
    DeviceBuffer<int32_t> d_tri;

    mcpt_counters counters{};
    std::string last_error;
};

Context& ctx();
int fail(int code, const std::string& msg);
int fail_cuda(cudaError_t e, const char* what);
int compare_device_tables(uint64_t bad[3]);   // device render-path tables against the host pack of the same tree (diagnostic)
int ensure_host_tree();     // downloads a device-built tree into c.tree if it is still on the device only
int ensure_host_packed();   // the host copy of the packed tables (pack_scene on c.scene / c.tree) if a device pack skipped it
int ensure_uploaded();  // needs lock held; binds the context's device to the calling thread; render-path tables
int ensure_parity_uploaded();  // + the tables only the parity hooks and the A/B kernels read (first use)
int bind_device();      // cudaSetDevice(context device): the current device is per host thread, so every entry point
                        // that touches the device calls this first (MCPT_ERR_NO_DEVICE when none is bound)

int pick_chunks(int width, int height, int spp);   // abi_render.cpp: slices per pixel (grid.z)
int check_cfg(const mcpt_render_cfg* cfg);        // abi_render.cpp: argument checks shared by every render entry point

// No exception crosses the C boundary (include/mcpt.h): every extern "C" body runs inside guarded().
template <class F>
int guarded(F&& body) noexcept {
    try {
        return body();
    } catch (const std::bad_alloc&) {
        return fail(MCPT_ERR_INTERNAL, "out of host memory");
    } catch (const std::exception& e) {
        try { return fail(MCPT_ERR_INTERNAL, std::string("exception: ") + e.what()); } catch (...) { return MCPT_ERR_INTERNAL; }
    } catch (...) {
        try { return fail(MCPT_ERR_INTERNAL, "unknown exception"); } catch (...) { return MCPT_ERR_INTERNAL; }
    }
}

}  // namespace mcpt
