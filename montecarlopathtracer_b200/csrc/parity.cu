// Parity hooks: the kernels whose results are compared bit for bit (or to an ulp) with the CPU oracle.
// ALWAYS compiled with -fmad=false and without -use_fast_math, whatever the render kernels' translation unit
// (launch.cu) uses: intersection in both traversal orders, the IEEE camera rays, the IEEE samplers, the
// shared-reciprocal division self-check.
#include "launch.h"

#include "../../include/mcpt.h"
#include "hooks.cuh"

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

}  // namespace

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

cudaError_t launch_primary_rays_exact(const DCamera& cam, int width, int height, int sample, uint32_t seed, uint32_t jitter,
                                      int rng_mode, uint32_t frame_seed, float* rays, cudaStream_t stream) {
    const unsigned blocks = (static_cast<unsigned>(width) * static_cast<unsigned>(height) + 255u) / 256u;
    if (rng_mode == MCPT_RNG_REFERENCE)
        primary_rays_kernel<RngReference, true><<<blocks, 256, 0, stream>>>(cam, width, height, sample, seed, jitter, frame_seed, rays);
    else
        primary_rays_kernel<RngPhilox, true><<<blocks, 256, 0, stream>>>(cam, width, height, sample, seed, jitter, frame_seed, rays);
    return cudaGetLastError();
}

cudaError_t launch_sample_exact(int kind, int64_t n, const float* u, const float* normals, const float* dirs,
                                const float* par0, const float* par1, float* out, cudaStream_t stream) {
    if (n <= 0) return cudaSuccess;
    sample_kernel<false><<<static_cast<unsigned>((n + 255) / 256), 256, 0, stream>>>(kind, n, u, normals, dirs, par0, par1, out);
    return cudaGetLastError();
}

}  // namespace mcpt
