// Kernels of the parity hooks that exist in two arithmetic forms, instantiated once per translation unit:
//   parity.cu (always -fmad=false)  the IEEE forms compared bit for bit / to an ulp with the CPU oracle;
//   launch.cu (the render kernels' flags)  the forms the render kernels really run (MCPT_FAST_SHADE, hardware units).
#pragma once
#include "render.cuh"

namespace mcpt {

// rtx.hlsl:376-398 for one sample index.  EXACT: the shader's expressions, rays bit-equal to the oracle's;
// !EXACT: camera_ray as render_kernel / render_kernel_sm call it.
template <class RNG, bool EXACT>
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
    camera_ray<EXACT>(cam, width, height, px, py, u1, u2, o, d);
    float* r = rays + 6 * static_cast<size_t>(i);
    r[0] = o.x; r[1] = o.y; r[2] = o.z; r[3] = d.x; r[4] = d.y; r[5] = d.z;
}

// The BSDF samplers of shading.cuh on n (uniforms, normal, dir, parameters) tuples (mcpt_sample_batch).
// kind: 0 sample_hemi, 1 sample_phong (par0 = Ns), 2 sample_fresnel (par0 = Tr, par1 = Ni), 3 the CUDA twin's
// sample_fresnel (Utils.hpp:97-137).  u: 2 floats per tuple (fresnel uses the first).
template <bool FAST>
__global__ void __launch_bounds__(256) sample_kernel(int kind, int64_t n, const float* __restrict__ u,
                                                     const float* __restrict__ normals, const float* __restrict__ dirs,
                                                     const float* __restrict__ par0, const float* __restrict__ par1,
                                                     float* __restrict__ out) {
    const int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) return;
    RngFixed rng{u[2 * i], u[2 * i + 1], 0};
    const F3 nrm = f3(normals[3 * i], normals[3 * i + 1], normals[3 * i + 2]);
    F3 d = f3(0, 0, 1), o;
    if (dirs) d = f3(dirs[3 * i], dirs[3 * i + 1], dirs[3 * i + 2]);
    if (kind == 0) o = sample_hemi<RngFixed, FAST>(rng, nrm);
    else if (kind == 1) o = sample_phong<RngFixed, FAST>(rng, nrm, d, par0[i]);
    else o = sample_fresnel<RngFixed, FAST>(rng, nrm, d, par0[i], par1[i], kind == 3);
    out[3 * i] = o.x; out[3 * i + 1] = o.y; out[3 * i + 2] = o.z;
}

}  // namespace mcpt
