// Kernel instantiations and launch wrappers.  Compile for sm_100a with -fmad=false and WITHOUT
// -use_fast_math: the intersection-parity gate and the per-sample agreement with the CPU oracle
// rely on IEEE mul/add/div/sqrt that are not contracted.
#include "launch.h"

#include "../../include/mcpt.h"
#include "render_sm.cuh"

#include <type_traits>

// Shipping configuration of the persistent-lane kernel (profiles/r1_tuning.md).
#ifndef MCPT_SM_VARIANT
#define MCPT_SM_VARIANT 503
#endif
#ifndef MCPT_SM_MIN_BLOCKS
#define MCPT_SM_MIN_BLOCKS 4
#endif

namespace mcpt {

namespace {

__global__ void __launch_bounds__(256) intersect_kernel(const DScene sc, const float* __restrict__ rays, int64_t n,
                                                        int variant, int32_t* __restrict__ tri, float* __restrict__ t,
                                                        float* __restrict__ beta, float* __restrict__ gamma) {
    const int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float* r = rays + 6 * i;
    const float ox = __ldg(r), oy = __ldg(r + 1), oz = __ldg(r + 2), dx = __ldg(r + 3), dy = __ldg(r + 4), dz = __ldg(r + 5);
    const Hit h = variant == MCPT_ISECT_ROPES ? intersect_ropes(sc, ox, oy, oz, dx, dy, dz)
                  : variant == MCPT_ISECT_KD  ? intersect_kd(sc, ox, oy, oz, dx, dy, dz)
                                              : intersect_ref(sc, ox, oy, oz, dx, dy, dz);
    tri[i] = static_cast<int32_t>(h.tri);
    if (t) t[i] = h.t;
    if (beta) beta[i] = h.beta;
    if (gamma) gamma[i] = h.gamma;
}

// Work counters of one traversal variant over a ray batch: {nodes popped, leaves/boxes entered,
// triangle tests} -- the per-ray figures the roofline's algorithmic work is stated in.
__global__ void __launch_bounds__(256) tally_kernel(const DScene sc, const float* __restrict__ rays, int64_t n, int variant,
                                                    unsigned long long* __restrict__ out) {
    const int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    TraversalTally tally;
    if (i < n) {
        const float* r = rays + 6 * i;
        if (variant == MCPT_ISECT_ROPES) intersect_ropes(sc, r[0], r[1], r[2], r[3], r[4], r[5], &tally);
        else if (variant == MCPT_ISECT_KD) intersect_kd(sc, r[0], r[1], r[2], r[3], r[4], r[5], &tally);
        else intersect_ref(sc, r[0], r[1], r[2], r[3], r[4], r[5], &tally);
    }
    uint32_t a = tally.popped, b = tally.entered, c = tally.tris;
    for (int off = 16; off > 0; off >>= 1) {
        a += __shfl_xor_sync(0xFFFFFFFFu, a, off); b += __shfl_xor_sync(0xFFFFFFFFu, b, off); c += __shfl_xor_sync(0xFFFFFFFFu, c, off);
    }
    if ((threadIdx.x & 31) == 0) { atomicAdd(out, (unsigned long long)a); atomicAdd(out + 1, (unsigned long long)b); atomicAdd(out + 2, (unsigned long long)c); }
}

template <class RNG>
__global__ void __launch_bounds__(256) primary_rays_kernel(const DCamera cam, int width, int height, uint32_t sample,
                                                           uint32_t seed, uint32_t jitter, uint32_t frameSeed,
                                                           float* __restrict__ rays) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= static_cast<uint32_t>(width) * static_cast<uint32_t>(height)) return;
    const uint32_t px = i % static_cast<uint32_t>(width), py = i / static_cast<uint32_t>(width);
    RNG rng;
    rng.start(i, seed, sample, frameSeed);
    float u1 = 0.5f, u2 = 0.5f;
    if (jitter) { u1 = rng.next(); u2 = rng.next(); }
    F3 o, d;
    camera_ray(cam, width, height, px, py, u1, u2, o, d);
    float* r = rays + 6 * static_cast<size_t>(i);
    r[0] = o.x; r[1] = o.y; r[2] = o.z; r[3] = d.x; r[4] = d.y; r[5] = d.z;
}

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

// div3_rn against __fdiv_rn on pseudo-random operands (Philox), counters[0] = mismatching quotients,
// counters[1] = quadruples that took the shared-reciprocal branch.
__global__ void __launch_bounds__(256) selfcheck_div3_kernel(long long n, uint32_t seed, int spread, unsigned long long* counters) {
    unsigned long long bad = 0, fast = 0;
    for (long long i = blockIdx.x * 256ll + threadIdx.x; i < n; i += gridDim.x * 256ll) {
        uint32_t o[4], e[4];
        philox4x32_10(static_cast<uint32_t>(i), static_cast<uint32_t>(i >> 32), 0u, 0u, seed, 0x51ED270Bu, o);
        philox4x32_10(static_cast<uint32_t>(i), static_cast<uint32_t>(i >> 32), 1u, 0u, seed, 0x51ED270Bu, e);
        float v[4];
        for (int k = 0; k < 4; ++k) {
            // random sign and mantissa, exponent uniform in [-spread, spread]
            const int ex = static_cast<int>(e[k] % static_cast<uint32_t>(2 * spread + 1)) - spread;
            v[k] = __uint_as_float((o[k] & 0x807FFFFFu) | (static_cast<uint32_t>(127 + ex) << 23));
        }
        float a0, a1, a2;
        div3_rn(v[0], v[1], v[2], v[3], a0, a1, a2);
        const float b0 = __fdiv_rn(v[0], v[3]), b1 = __fdiv_rn(v[1], v[3]), b2 = __fdiv_rn(v[2], v[3]);
        bad += (__float_as_uint(a0) != __float_as_uint(b0)) + (__float_as_uint(a1) != __float_as_uint(b1)) +
               (__float_as_uint(a2) != __float_as_uint(b2));
        fast += div3_guard<false>(v[0], v[1], v[2], v[3]) ? 1u : 0u;
    }
    bad = __reduce_add_sync(0xFFFFFFFFu, static_cast<unsigned>(bad));
    fast = __reduce_add_sync(0xFFFFFFFFu, static_cast<unsigned>(fast));
    if ((threadIdx.x & 31u) == 0u) { atomicAdd(counters, bad); atomicAdd(counters + 1, fast); }
}

template <class RNG, bool KD, bool NEE, bool MOMENTS>
cudaError_t launch_render_t(const RenderLaunch& r, cudaStream_t stream) {
    RenderParams p;
    p.sc = r.sc; p.cam = r.cam; p.width = r.width; p.height = r.height;
    p.sample_begin = r.sample_begin; p.sample_count = r.sample_count; p.chunks = r.chunks;
    p.depth = r.depth; p.kill_at = r.kill_at; p.gather_at = r.gather_at; p.twin = r.twin; p.seed = r.seed; p.jitter = r.jitter; p.emission_scale = r.emission_scale; p.cam_threshold = 16; p.tri_cap = 4;
    p.sum = r.slices; p.sumsq = r.slices_sq; p.frame_seeds = r.frame_seeds; p.ray_counter = r.ray_counter;
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
    p.sum = r.slices; p.sumsq = r.slices_sq; p.frame_seeds = r.frame_seeds; p.ray_counter = r.ray_counter;
    dim3 grid((r.width + 15) / 16, (r.height + 15) / 16, r.chunks);
#ifdef MCPT_TUNE
    // tuning build: every kernel variant x residency for the headline combination (Philox, no NEE, no moments)
    if constexpr (std::is_same<RNG, RngPhilox>::value && !NEE && !MOMENTS) {
#define MCPT_V(V) case V: render_kernel_sm<RNG, NEE, MOMENTS, 4, V><<<grid, 256, 0, stream>>>(p, r.shade_threshold); break;
        if (r.variant == 55 && r.min_blocks == 8) {  // 128-thread blocks (16x8 tiles), 8 per SM
            grid.y = (r.height + 7) / 8;
            render_kernel_sm<RNG, NEE, MOMENTS, 8, 55, 128><<<grid, 128, 0, stream>>>(p, r.shade_threshold);
            return cudaGetLastError();
        }
        if (r.variant == 55 && r.min_blocks == 16) {  // 64-thread blocks (16x4 tiles), 16 per SM
            grid.y = (r.height + 3) / 4;
            render_kernel_sm<RNG, NEE, MOMENTS, 16, 55, 64><<<grid, 64, 0, stream>>>(p, r.shade_threshold);
            return cudaGetLastError();
        }
        switch (r.variant) {
            MCPT_V(5) MCPT_V(7) MCPT_V(21) MCPT_V(23) MCPT_V(13) MCPT_V(15) MCPT_V(4) MCPT_V(20) MCPT_V(55) MCPT_V(53) MCPT_V(119) MCPT_V(247) MCPT_V(375) MCPT_V(503)
            default: return cudaErrorInvalidValue;
        }
#undef MCPT_V
        return cudaGetLastError();
    }
#endif
    render_kernel_sm<RNG, NEE, MOMENTS, MCPT_SM_MIN_BLOCKS, MCPT_SM_VARIANT><<<grid, 256, 0, stream>>>(p, r.shade_threshold);
    return cudaGetLastError();
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

cudaError_t launch_intersect(const DScene& sc, const float* rays, int64_t n, int variant, int32_t* tri, float* t,
                             float* beta, float* gamma, cudaStream_t stream, uint32_t* launches) {
    if (n <= 0) return cudaSuccess;
    if (launches) ++*launches;
    intersect_kernel<<<static_cast<unsigned>((n + 255) / 256), 256, 0, stream>>>(sc, rays, n, variant, tri, t, beta, gamma);
    return cudaGetLastError();
}

cudaError_t launch_tally(const DScene& sc, const float* rays, int64_t n, int variant, unsigned long long* out, cudaStream_t stream) {
    if (n <= 0) return cudaSuccess;
    tally_kernel<<<static_cast<unsigned>((n + 255) / 256), 256, 0, stream>>>(sc, rays, n, variant, out);
    return cudaGetLastError();
}

cudaError_t launch_selfcheck_div3(long long n, uint32_t seed, int spread, unsigned long long* counters, cudaStream_t stream) {
    selfcheck_div3_kernel<<<148 * 8, 256, 0, stream>>>(n, seed, spread, counters);
    return cudaGetLastError();
}

cudaError_t launch_fp32_peak(float* scratch, int blocks, int iters, cudaStream_t stream) {
    fp32_peak_kernel<<<blocks, 256, 0, stream>>>(scratch, iters, 0.999f, 0.001f);
    return cudaGetLastError();
}

cudaError_t launch_primary_rays(const DCamera& cam, int width, int height, int sample, uint32_t seed, uint32_t jitter,
                                int rng_mode, uint32_t frame_seed, float* rays, cudaStream_t stream, uint32_t* launches) {
    if (launches) ++*launches;
    const unsigned blocks = (static_cast<unsigned>(width) * static_cast<unsigned>(height) + 255u) / 256u;
    if (rng_mode == MCPT_RNG_REFERENCE)
        primary_rays_kernel<RngReference><<<blocks, 256, 0, stream>>>(cam, width, height, sample, seed, jitter, frame_seed, rays);
    else
        primary_rays_kernel<RngPhilox><<<blocks, 256, 0, stream>>>(cam, width, height, sample, seed, jitter, frame_seed, rays);
    return cudaGetLastError();
}

}  // namespace mcpt
