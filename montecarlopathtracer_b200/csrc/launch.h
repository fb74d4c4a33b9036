// Host-callable launchers of the sm_100a kernels (launch.cu).
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

#include "device_scene.h"

namespace mcpt {

struct RenderLaunch {
    DScene sc;
    DCamera cam;
    int width, height;
    int sample_begin, sample_count, chunks;
    uint32_t depth, kill_at, gather_at, twin, seed, jitter;
    float emission_scale;
    float4* slices;        // [chunks][h][w] scratch (or the destination itself when chunks == 1 and !accumulate)
    float4* slices_sq;     // nullptr unless moments
    const uint32_t* frame_seeds;
    unsigned long long* ray_counter;
    int rng_mode;          // MCPT_RNG_*
    bool kd, nee, moments;
    bool state_machine;    // render_kernel_sm (warp-voted state machine) instead of render_kernel
    int shade_threshold;   // lanes that must be waiting before the warp leaves the traverse phase
    int min_blocks;        // tuning builds only: 2, 3 or 4 resident 256-thread blocks per SM (register cap 128 / 85 / 64)
    int pool;              // 1: render_kernel_pool (64 pooled paths per warp; emission-on-hit, no moments); experimental, off by default
    float* pool_state;     // its scratch: pool_state_floats(sm_count) floats
    int sm_count;
    int variant;           // tuning builds only: VAR bits of render_kernel_sm
    int cam_threshold;     // camera-ray queue refill threshold (lanes)
    int tri_cap;           // triangle tests per lane and round (VAR bit 6)
};

size_t pool_state_floats(int sm_count);   // scratch of render_kernel_pool for a device of that many SMs

// Number of kernels launched is returned through *launches (adds to it).
cudaError_t launch_render(const RenderLaunch& r, cudaStream_t stream, uint32_t* launches);

// dst[i] (+)= sum over chunks of slices[c][i]   (fixed order -> deterministic)
cudaError_t launch_reduce_slices(const float4* slices, int chunks, size_t pixels, float4* dst, bool accumulate,
                                 cudaStream_t stream, uint32_t* launches);

// rgb[i*3..] = sum[i].xyz / sum[i].w
cudaError_t launch_finalize(const float4* sum, size_t pixels, float* rgb, cudaStream_t stream, uint32_t* launches);

cudaError_t launch_intersect(const DScene& sc, const float* rays, int64_t n, int variant, int32_t* tri, float* t,
                             float* beta, float* gamma, cudaStream_t stream, uint32_t* launches);

// render_path: the rays as the render kernels generate them (hardware rsqrt, folded factors) instead of the IEEE form.
cudaError_t launch_primary_rays(const DCamera& cam, int width, int height, int sample, uint32_t seed, uint32_t jitter,
                                int rng_mode, uint32_t frame_seed, bool render_path, float* rays, cudaStream_t stream,
                                uint32_t* launches);
cudaError_t launch_primary_rays_exact(const DCamera& cam, int width, int height, int sample, uint32_t seed, uint32_t jitter,
                                      int rng_mode, uint32_t frame_seed, float* rays, cudaStream_t stream);  // parity.cu

// BSDF samplers on n tuples (kind: 0 hemi, 1 phong, 2 fresnel, 3 twin fresnel); render_path as above.
cudaError_t launch_sample(int kind, int64_t n, const float* u, const float* normals, const float* dirs, const float* par0,
                          const float* par1, bool render_path, float* out, cudaStream_t stream, uint32_t* launches);
cudaError_t launch_sample_exact(int kind, int64_t n, const float* u, const float* normals, const float* dirs,
                                const float* par0, const float* par1, float* out, cudaStream_t stream);  // parity.cu

cudaError_t launch_tally(const DScene& sc, const float* rays, int64_t n, int variant, unsigned long long* out3, cudaStream_t stream);

// scratch: blocks*256 floats. flops = blocks*256*iters*16.
cudaError_t launch_fp32_peak(float* scratch, int blocks, int iters, cudaStream_t stream);
cudaError_t launch_selfcheck_div3(long long n, uint32_t seed, int spread, unsigned long long* counters, cudaStream_t stream);

}  // namespace mcpt
