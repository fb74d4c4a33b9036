// mcpt_render_accumulate_device_noalloc: the render + slice reduce of mcpt_render_accumulate_device for callers that
// cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid allocation).
// No counterpart in the reference.
//
// The scratch the render path normally grows on demand (Context::d_slices, d_counter) lives here in ONE file-scope
// __device__ array of fixed size: 256 bytes of ray counters + MCPT_NOALLOC_MAX_SLICE_PIXELS float4 slice entries
// (16 MiB).  Every GPU that loads this module holds it until the process exits.  This entry point calls no cudaMalloc,
// cudaFree or allocating library routine; the scene tables must already be resident (they go up with the geometry).
//
// Host code may not use the array's NAME as a pointer: on the host it is a stand-in object, not a device address (a
// kernel given it faults, or silently uses host-reachable memory).  The device address is asked for once per device,
// with that device current, by cudaGetSymbolAddress, and that pointer goes to the render kernel.  The reduce kernel
// of this file is device code and names the array directly -- so a wrong address would show as an empty image.
//
// Every call shares the array and finds the previous call's contents in it: the counters are zeroed per call on the
// caller's stream, the slice entries of this call's (pixels x slices) range are all overwritten by the render kernel
// before the reduce reads them.  Two calls that overlap on the GPU (different streams, or another thread once the
// library lock is released) race on it -- the call is blocking and serialised by the context lock for that reason.
#include <algorithm>
#include <cstdint>

#include "context.h"
#include "launch.h"

namespace mcpt {
namespace {

constexpr size_t kCounterBytes = 256;   // 16 x unsigned long long ray / round counters, padded to the array's alignment
constexpr size_t kMaxSliceEntries = MCPT_NOALLOC_MAX_SLICE_PIXELS;
constexpr int kMaxDevices = 64;

static __device__ __align__(256) unsigned char noalloc_scratch[kCounterBytes + kMaxSliceEntries * sizeof(float4)];

// device address of noalloc_scratch per device ordinal (nullptr = not asked yet)
unsigned char* g_scratch_addr[kMaxDevices] = {};

// dst[i] += sum over slices of scratch_slices[c][i]; same fixed order as reduce_slices_kernel (launch.cu), the slices
// read from the file-scope array by name.
__global__ void __launch_bounds__(256) reduce_static_slices_kernel(int chunks, size_t pixels, float4* __restrict__ dst) {
    const size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= pixels) return;
    const float4* slices = reinterpret_cast<const float4*>(noalloc_scratch + kCounterBytes);
    float4 a = dst[i];
    for (int c = 0; c < chunks; ++c) {
        const float4 v = slices[static_cast<size_t>(c) * pixels + i];
        a.x += v.x; a.y += v.y; a.z += v.z; a.w += v.w;
    }
    dst[i] = a;
}

// Needs the context's device current (bind_device()).
int scratch_address(int device, unsigned char** out) {
    if (device < 0 || device >= kMaxDevices) return fail(MCPT_ERR_INVALID, "no-alloc render: device ordinal above 63");
    if (!g_scratch_addr[device]) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, noalloc_scratch);
        if (e != cudaSuccess) return fail_cuda(e, "cudaGetSymbolAddress(noalloc_scratch)");
        if (!p) return fail(MCPT_ERR_INTERNAL, "cudaGetSymbolAddress returned a null address");
        g_scratch_addr[device] = static_cast<unsigned char*>(p);
    }
    *out = g_scratch_addr[device];
    return MCPT_OK;
}

}  // namespace
}  // namespace mcpt

using namespace mcpt;

extern "C" int mcpt_render_accumulate_device_noalloc(const mcpt_render_cfg* cfg, void* dev_sum, void* stream) {
    return guarded([&]() -> int {
    Context& c = ctx();
    std::lock_guard<std::mutex> g(c.lock);
    if (!dev_sum) return fail(MCPT_ERR_INVALID, "null dev_sum");
    int rc = check_cfg(cfg);
    if (rc) return rc;
    if (cfg->rng_mode != MCPT_RNG_PHILOX) return fail(MCPT_ERR_INVALID, "no-alloc render: MCPT_RNG_PHILOX only (the reference stream uploads a frame-seed table)");
    if (cfg->flags & (MCPT_FLAG_MOMENTS | MCPT_FLAG_REF_TRAVERSAL | MCPT_FLAG_SIMPLE_KERNEL))
        return fail(MCPT_ERR_INVALID, "no-alloc render: MOMENTS, REF_TRAVERSAL and SIMPLE_KERNEL need buffers this entry point may not allocate");
    const size_t pixels = static_cast<size_t>(cfg->width) * cfg->height;
    if (pixels > kMaxSliceEntries)
        return fail(MCPT_ERR_INVALID, "no-alloc render: width*height = " + std::to_string(pixels) + " exceeds MCPT_NOALLOC_MAX_SLICE_PIXELS = " +
                                          std::to_string(kMaxSliceEntries) + " (fixed scratch); use mcpt_render_accumulate_device");
    if ((rc = bind_device())) return rc;
    if (!c.has_geometry || !c.uploaded) return fail(MCPT_ERR_INVALID, "no-alloc render: no geometry resident on the device (load it after mcpt_initialize; this call uploads nothing)");
    unsigned char* scratch = nullptr;
    if ((rc = scratch_address(c.device, &scratch))) return rc;
    unsigned long long* counters = reinterpret_cast<unsigned long long*>(scratch);
    float4* slices = reinterpret_cast<float4*>(scratch + kCounterBytes);
    // As many slices as the allocating path would take, as far as the fixed scratch holds them (fewer slices = same
    // samples, other fp32 summation order).
    const int chunks = static_cast<int>(std::max<size_t>(1, std::min<size_t>(pick_chunks(cfg->width, cfg->height, cfg->spp), kMaxSliceEntries / pixels)));
    cudaStream_t s = stream ? static_cast<cudaStream_t>(stream) : c.stream;
    cudaError_t e;
    if ((e = cudaMemsetAsync(counters, 0, kCounterBytes, s)) != cudaSuccess) return fail_cuda(e, "memset counters");

    RenderLaunch r;
    r.sc = c.dscene;
    const bool twin = cfg->tracer_mode == MCPT_TRACER_CUDA_TWIN;
    r.cam = make_camera(cfg->eye, cfg->at, cfg->up, cfg->fovy, cfg->width, cfg->height, twin ? 1 : 0);
    r.width = cfg->width; r.height = cfg->height;
    r.sample_begin = cfg->sample_offset; r.sample_count = cfg->spp; r.chunks = chunks;
    r.depth = twin ? 0xFFFFFFFFu : static_cast<uint32_t>(cfg->max_depth);          // as render_core (abi_render.cpp)
    r.kill_at = twin ? 0xFFFFFFFFu : 3u * static_cast<uint32_t>(cfg->max_depth);
    r.gather_at = twin ? static_cast<uint32_t>(cfg->max_depth) : 0xFFFFFFFFu;
    r.twin = twin ? 1u : 0u;
    r.seed = cfg->seed;
    r.jitter = (cfg->flags & MCPT_FLAG_NO_JITTER) ? 0u : 1u;
    r.emission_scale = cfg->emission_scale;
    r.slices = slices;
    r.slices_sq = nullptr;
    r.frame_seeds = nullptr;
    r.ray_counter = counters;
    r.rng_mode = MCPT_RNG_PHILOX;
    r.kd = true;
    r.nee = (cfg->flags & MCPT_FLAG_NEE) != 0;
    r.moments = false;
    r.state_machine = true;
    const Tuning& tn = tuning();
    r.shade_threshold = std::min(32, std::max(1, tn.shade_threshold));
    r.cam_threshold = std::min(32, std::max(1, tn.cam_threshold));
    r.tri_cap = std::min(64, std::max(1, tn.tri_cap));
    r.min_blocks = tn.min_blocks > 0 ? tn.min_blocks : (c.large_scene ? 8 : 4);
    r.pool = 0; r.pool_state = nullptr; r.sm_count = 0;   // the pooled kernel's scratch is cudaMalloc'ed: not on this path
    r.variant = tn.variant;

    c.counters.launches = 0;
    cudaEventRecord(c.ev[0], s);
    if ((e = launch_render(r, s, &c.counters.launches)) != cudaSuccess) return fail_cuda(e, "render kernel launch");
    cudaEventRecord(c.ev[1], s);
    ++c.counters.launches;
    reduce_static_slices_kernel<<<static_cast<unsigned>((pixels + 255) / 256), 256, 0, s>>>(chunks, pixels, static_cast<float4*>(dev_sum));
    if ((e = cudaGetLastError()) != cudaSuccess) return fail_cuda(e, "reduce launch");
    c.counters.paths = static_cast<uint64_t>(pixels) * static_cast<uint64_t>(cfg->spp);

    unsigned long long rays = 0;   // counter read-back into a stack word: a copy, not an allocation
    if ((e = cudaMemcpyAsync(&rays, counters, sizeof(rays), cudaMemcpyDeviceToHost, s)) != cudaSuccess) return fail_cuda(e, "counter readback");
    cudaEventRecord(c.ev[2], s);
    if ((e = cudaStreamSynchronize(s)) != cudaSuccess) return fail_cuda(e, "no-alloc render");
    float k = 0, t = 0;
    cudaEventElapsedTime(&k, c.ev[0], c.ev[1]);
    cudaEventElapsedTime(&t, c.ev[0], c.ev[2]);
    c.counters.rays = rays;
    c.counters.kernel_ms = k;
    c.counters.total_ms = t;
    return MCPT_OK;
    });
}
