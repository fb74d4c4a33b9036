// Render-kernel instantiations and launch wrappers (sm_100a).  The triangle test is written in explicit
// round-to-nearest intrinsics (intersect.cuh: tri_test) and is bit-exact whatever this file's flags; shading and
// the walk's steering use the hardware units (MCPT_FAST_SHADE).  The kernels whose outputs are
// compared bit for bit with the CPU oracle live in parity.cu (always -fmad=false).
#include "launch.h"

#include "../../include/mcpt.h"
#include "hooks.cuh"
#include "render_sm.cuh"
#include "render_pool.cuh"

#include <type_traits>

// Shipping configuration of the persistent-lane kernel (profiles/r1_tuning.md).
#ifndef MCPT_SM_VARIANT
#define MCPT_SM_VARIANT 503
#endif
#ifndef MCPT_SM_MIN_BLOCKS
#define MCPT_SM_MIN_BLOCKS 4
#endif

namespace mcpt {

constexpr int kPoolBlocksPerSm = 4, kPoolSets = 2;
size_t pool_state_floats(int sm_count) { return static_cast<size_t>(sm_count) * kPoolBlocksPerSm * 8 * PG_WORDS * 32 * kPoolSets; }

namespace {

__global__ void __launch_bounds__(256) reduce_slices_kernel(const float4* __restrict__ slices, int chunks, size_t pixels,
                                                            float4* __restrict__ dst, int accumulate) {
    const size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= pixels) return;
    float4 a = accumulate ? dst[i] : make_float4(0, 0, 0, 0);
    for (int c = 0; c < chunks; ++c) {
        const float4 v = slices[static_cast<size_t>(c) * pixels + i];
        a.x += v.x; a.y += v.y; a.z += v.z; a.w += v.w;
    }
    dst[i] = a;
}

__global__ void __launch_bounds__(256) finalize_kernel(const float4* __restrict__ sum, size_t pixels, float* __restrict__ rgb) {
    const size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= pixels) return;
    const float4 v = sum[i];
    const float inv = v.w > 0 ? 1.0f / v.w : 0.0f;
    rgb[3 * i] = v.x * inv; rgb[3 * i + 1] = v.y * inv; rgb[3 * i + 2] = v.z * inv;
}

// FP32 issue-rate probe for the roofline denominator (MEASURED_PEAKS.json has no FP32
// figure): 8 independent FMA chains per thread, 2 flops per FMA.
__global__ void __launch_bounds__(256) fp32_peak_kernel(float* __restrict__ out, int iters, float x, float y) {
    float a0 = threadIdx.x, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6, a7 = a0 + 7;
    for (int i = 0; i < iters; ++i) {
        a0 = __fmaf_rn(a0, x, y); a1 = __fmaf_rn(a1, x, y); a2 = __fmaf_rn(a2, x, y); a3 = __fmaf_rn(a3, x, y);
        a4 = __fmaf_rn(a4, x, y); a5 = __fmaf_rn(a5, x, y); a6 = __fmaf_rn(a6, x, y); a7 = __fmaf_rn(a7, x, y);
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = ((a0 + a1) + (a2 + a3)) + ((a4 + a5) + (a6 + a7));
}

template <class RNG, bool KD, bool NEE, bool MOMENTS>
cudaError_t launch_render_t(const RenderLaunch& r, cudaStream_t stream) {
    RenderParams p;
    p.sc = r.sc; p.cam = r.cam; p.width = r.width; p.height = r.height;
    p.sample_begin = r.sample_begin; p.sample_count = r.sample_count; p.chunks = r.chunks;
    p.depth = r.depth; p.kill_at = r.kill_at; p.gather_at = r.gather_at; p.twin = r.twin; p.seed = r.seed; p.jitter = r.jitter; p.emission_scale = r.emission_scale; p.cam_threshold = 16; p.tri_cap = 4;
    p.sum = r.slices; p.sumsq = r.slices_sq; p.frame_seeds = r.frame_seeds; p.ray_counter = r.ray_counter; p.pool_state = nullptr;
    dim3 grid((r.width + 15) / 16, (r.height + 15) / 16, r.chunks);
    render_kernel<RNG, KD, NEE, MOMENTS><<<grid, 256, 0, stream>>>(p);
    return cudaGetLastError();
}

template <class RNG, bool NEE, bool MOMENTS>
cudaError_t launch_render_sm_t(const RenderLaunch& r, cudaStream_t stream) {
    RenderParams p;
    p.sc = r.sc; p.cam = r.cam; p.width = r.width; p.height = r.height;
    p.sample_begin = r.sample_begin; p.sample_count = r.sample_count; p.chunks = r.chunks;
    p.depth = r.depth; p.kill_at = r.kill_at; p.gather_at = r.gather_at; p.twin = r.twin; p.seed = r.seed; p.jitter = r.jitter; p.emission_scale = r.emission_scale; p.cam_threshold = r.cam_threshold; p.tri_cap = r.tri_cap;
    p.sum = r.slices; p.sumsq = r.slices_sq; p.frame_seeds = r.frame_seeds; p.ray_counter = r.ray_counter; p.pool_state = nullptr;
    dim3 grid((r.width + 15) / 16, (r.height + 15) / 16, r.chunks);
    // Pooled lanes (render_pool.cuh; mcpt_set_tuning("pool", 1); off by default, profiles/r2_tuning.md): persistent warps,
    // 64 paths per warp; emission-on-hit estimator without moments.
    if constexpr (!NEE && !MOMENTS) {
        if (r.pool > 0 && r.pool_state != nullptr && r.sm_count > 0) {
            p.pool_state = r.pool_state;
            render_kernel_pool<RNG, kPoolBlocksPerSm, kPoolSets><<<static_cast<unsigned>(r.sm_count * kPoolBlocksPerSm), 256, 0, stream>>>(p);
            return cudaGetLastError();
        }
    }
#ifdef MCPT_TUNE
    // tuning build: kernel variants for the headline combination (Philox, no NEE, no moments); 511 = shipped + per-warp
    // round statistics (VAR bit 3).  The round-1 variants (5 ... 375, 128- and 64-thread blocks) were measured and
    // removed; their numbers are in profiles/r1_tuning.md.
    if constexpr (std::is_same<RNG, RngPhilox>::value && !NEE && !MOMENTS) {
        switch (r.variant) {
            case 119: render_kernel_sm<RNG, NEE, MOMENTS, 4, 119><<<grid, 256, 0, stream>>>(p, r.shade_threshold); return cudaGetLastError();
            case 511: render_kernel_sm<RNG, NEE, MOMENTS, 4, 511><<<grid, 256, 0, stream>>>(p, r.shade_threshold); return cudaGetLastError();
            case 1015: return r.min_blocks == 3 ? (render_kernel_sm<RNG, NEE, MOMENTS, 3, 1015><<<grid, 256, 0, stream>>>(p, r.shade_threshold), cudaGetLastError())
                            : r.min_blocks == 8 ? (render_kernel_sm<RNG, NEE, MOMENTS, 8, 1015><<<grid, 256, 0, stream>>>(p, r.shade_threshold), cudaGetLastError())
                                                : (render_kernel_sm<RNG, NEE, MOMENTS, 4, 1015><<<grid, 256, 0, stream>>>(p, r.shade_threshold), cudaGetLastError());
            case 1023: render_kernel_sm<RNG, NEE, MOMENTS, 4, 1023><<<grid, 256, 0, stream>>>(p, r.shade_threshold); return cudaGetLastError();
            default: break;
        }
    }
#endif
    // Residency: scenes whose render-path tables outgrow L2 (K4: 1 M triangles) are bound by the latency of dependent
    // L2 / DRAM misses (issue slots 43 % busy at 4 blocks per SM): 8 resident blocks (32 registers, the rest spilled to L1)
    // give +25 % there and -15 % on the L1-resident scenes.  abi_render.cpp picks 8 for large scenes (min_blocks = 0 = automatic);
    // 5 and 6 exist for the A/B (mcpt_set_tuning("min_blocks", n); Philox stream; profiles/r2_tuning.md).
    if constexpr (std::is_same<RNG, RngPhilox>::value) {
        switch (r.min_blocks) {
            case 5: if constexpr (!NEE && !MOMENTS) { render_kernel_sm<RNG, NEE, MOMENTS, 5, MCPT_SM_VARIANT><<<grid, 256, 0, stream>>>(p, r.shade_threshold); return cudaGetLastError(); } break;
            case 6: if constexpr (!NEE && !MOMENTS) { render_kernel_sm<RNG, NEE, MOMENTS, 6, MCPT_SM_VARIANT><<<grid, 256, 0, stream>>>(p, r.shade_threshold); return cudaGetLastError(); } break;
            case 8: render_kernel_sm<RNG, NEE, MOMENTS, 8, MCPT_SM_VARIANT><<<grid, 256, 0, stream>>>(p, r.shade_threshold); return cudaGetLastError();
            default: break;
        }
    }
    render_kernel_sm<RNG, NEE, MOMENTS, MCPT_SM_MIN_BLOCKS, MCPT_SM_VARIANT><<<grid, 256, 0, stream>>>(p, r.shade_threshold); return cudaGetLastError();
}

template <class RNG>
cudaError_t launch_render_rng(const RenderLaunch& r, cudaStream_t s) {
    if (r.state_machine && r.kd) {
        switch ((r.nee ? 2 : 0) | (r.moments ? 1 : 0)) {
            case 0: return launch_render_sm_t<RNG, false, false>(r, s);
            case 1: return launch_render_sm_t<RNG, false, true>(r, s);
            case 2: return launch_render_sm_t<RNG, true, false>(r, s);
            default: return launch_render_sm_t<RNG, true, true>(r, s);
        }
    }
    const int key = (r.kd ? 4 : 0) | (r.nee ? 2 : 0) | (r.moments ? 1 : 0);
    switch (key) {
        case 0: return launch_render_t<RNG, false, false, false>(r, s);
        case 1: return launch_render_t<RNG, false, false, true>(r, s);
        case 2: return launch_render_t<RNG, false, true, false>(r, s);
        case 3: return launch_render_t<RNG, false, true, true>(r, s);
        case 4: return launch_render_t<RNG, true, false, false>(r, s);
        case 5: return launch_render_t<RNG, true, false, true>(r, s);
        case 6: return launch_render_t<RNG, true, true, false>(r, s);
        default: return launch_render_t<RNG, true, true, true>(r, s);
    }
}

}  // namespace

cudaError_t launch_render(const RenderLaunch& r, cudaStream_t stream, uint32_t* launches) {
    if (launches) ++*launches;
    if (r.rng_mode == MCPT_RNG_REFERENCE) return launch_render_rng<RngReference>(r, stream);
    return launch_render_rng<RngPhilox>(r, stream);
}

cudaError_t launch_reduce_slices(const float4* slices, int chunks, size_t pixels, float4* dst, bool accumulate,
                                 cudaStream_t stream, uint32_t* launches) {
    if (launches) ++*launches;
    reduce_slices_kernel<<<static_cast<unsigned>((pixels + 255) / 256), 256, 0, stream>>>(slices, chunks, pixels, dst, accumulate ? 1 : 0);
    return cudaGetLastError();
}

cudaError_t launch_finalize(const float4* sum, size_t pixels, float* rgb, cudaStream_t stream, uint32_t* launches) {
    if (launches) ++*launches;
    finalize_kernel<<<static_cast<unsigned>((pixels + 255) / 256), 256, 0, stream>>>(sum, pixels, rgb);
    return cudaGetLastError();
}

cudaError_t launch_fp32_peak(float* scratch, int blocks, int iters, cudaStream_t stream) {
    fp32_peak_kernel<<<blocks, 256, 0, stream>>>(scratch, iters, 0.999f, 0.001f);
    return cudaGetLastError();
}

cudaError_t launch_primary_rays(const DCamera& cam, int width, int height, int sample, uint32_t seed, uint32_t jitter,
                                int rng_mode, uint32_t frame_seed, bool render_path, float* rays, cudaStream_t stream,
                                uint32_t* launches) {
    if (launches) ++*launches;
    if (!render_path) return launch_primary_rays_exact(cam, width, height, sample, seed, jitter, rng_mode, frame_seed, rays, stream);
    const unsigned blocks = (static_cast<unsigned>(width) * static_cast<unsigned>(height) + 255u) / 256u;
    if (rng_mode == MCPT_RNG_REFERENCE)
        primary_rays_kernel<RngReference, !kFastShade><<<blocks, 256, 0, stream>>>(cam, width, height, sample, seed, jitter, frame_seed, rays);
    else
        primary_rays_kernel<RngPhilox, !kFastShade><<<blocks, 256, 0, stream>>>(cam, width, height, sample, seed, jitter, frame_seed, rays);
    return cudaGetLastError();
}

cudaError_t launch_sample(int kind, int64_t n, const float* u, const float* normals, const float* dirs, const float* par0,
                          const float* par1, bool render_path, float* out, cudaStream_t stream, uint32_t* launches) {
    if (n <= 0) return cudaSuccess;
    if (launches) ++*launches;
    if (!render_path) return launch_sample_exact(kind, n, u, normals, dirs, par0, par1, out, stream);
    sample_kernel<kFastShade><<<static_cast<unsigned>((n + 255) / 256), 256, 0, stream>>>(kind, n, u, normals, dirs, par0, par1, out);
    return cudaGetLastError();
}

}  // namespace mcpt
