// C ABI, part 2: RenderScene and the intersection parity hook.
// Replaces PW::Tracer::RenderScene (CUTracer.cu:340-404) and the per-frame part of
// GraphicsRTX::DoOnUpdate (GraphicsRTX.cpp:173-209): camera, frame seeds, launch, read-back.
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>

#include "context.h"
#include "launch.h"

using namespace mcpt;

namespace mcpt {

// How many slices the spp of a pixel are cut into (grid.z; each slice accumulates into its own buffer, one
// reduce pass adds them).  Two reasons to cut: (1) small images: enough threads to fill 148 SMs several times
// over; (2) the tail: a block carries all the samples of its 16x16 tile, the 4 096 blocks of a 1024^2 image
// are 6.9 waves of 148 x 4 resident blocks and the last wave ran the SMs 7 % below their average active time
// (profiles/r1d_*: sm__cycles_active.avg against .max) -- aim at >= 24 waves while a slice keeps >= 64 spp
// (block start-up and drain are per slice).  K2: 4 slices of 256 spp, +5 % (profiles/r1_tuning.md).
int pick_chunks(int width, int height, int spp) {
    const long long pixels = static_cast<long long>(width) * height;
    const long long target = 148LL * 2048 * 2;
    long long chunks = (target + pixels - 1) / pixels;
    const long long tiles = ((width + 15LL) / 16) * ((height + 15LL) / 16);
    const long long forTail = std::min<long long>((24LL * 148 * 4 + tiles - 1) / tiles, spp / 64);
    chunks = std::max(chunks, forTail);
    chunks = std::max(1LL, std::min<long long>(chunks, spp));
    chunks = std::min<long long>(chunks, 64);
    const int forced = tuning().chunks;  // mcpt_set_tuning("chunks", n): tests (slice independence) and tuning
    if (forced > 0) chunks = std::min(forced, spp);
    return static_cast<int>(chunks);
}


int check_cfg(const mcpt_render_cfg* cfg) {
    if (!cfg) return fail(MCPT_ERR_INVALID, "null render cfg");
    if (cfg->width <= 0 || cfg->height <= 0 || cfg->width > 32768 || cfg->height > 32768) return fail(MCPT_ERR_INVALID, "bad image size");
    if (cfg->spp <= 0) return fail(MCPT_ERR_INVALID, "spp must be positive");
    if (cfg->max_depth < 0 || cfg->max_depth > 1000000) return fail(MCPT_ERR_INVALID, "bad max_depth");
    if (cfg->sample_offset < 0) return fail(MCPT_ERR_INVALID, "negative sample_offset");
    if (cfg->rng_mode != MCPT_RNG_PHILOX && cfg->rng_mode != MCPT_RNG_REFERENCE) return fail(MCPT_ERR_INVALID, "unknown rng_mode");
    if (!(cfg->fovy > 0.0f && cfg->fovy < 3.14159f)) return fail(MCPT_ERR_INVALID, "fovy out of range");
    if (cfg->tracer_mode != MCPT_TRACER_MCRT && cfg->tracer_mode != MCPT_TRACER_CUDA_TWIN) return fail(MCPT_ERR_INVALID, "unknown tracer_mode");
    if (cfg->tracer_mode == MCPT_TRACER_CUDA_TWIN && cfg->rng_mode == MCPT_RNG_REFERENCE) return fail(MCPT_ERR_INVALID, "the CUDA twin has no reproducible reference stream (cuRAND seeded from random_device, CUTracer.cu:375-376): use MCPT_RNG_PHILOX");
    return MCPT_OK;
}

// Launches the path-sampling kernel for cfg and leaves per-pixel sums in *sum (and *sumsq).
// When `dst` is given the samples are ADDED into it (multi-GPU / progressive accumulation),
// otherwise the result pointer refers to library scratch.
int render_core(const mcpt_render_cfg* cfg, cudaStream_t stream, float4* dst, float4* dstSq, const float4** sum,
                const float4** sumsq) {
    Context& c = ctx();
    int rc = check_cfg(cfg);
    if (rc) return rc;
    const bool needsParityTables = (cfg->flags & (MCPT_FLAG_REF_TRAVERSAL | MCPT_FLAG_SIMPLE_KERNEL)) != 0;   // the A/B kernels walk kd_nodes / ref_nodes
    if ((rc = needsParityTables ? ensure_parity_uploaded() : ensure_uploaded())) return rc;
    const bool moments = (cfg->flags & MCPT_FLAG_MOMENTS) != 0;
    const size_t pixels = static_cast<size_t>(cfg->width) * cfg->height;
    const int chunks = pick_chunks(cfg->width, cfg->height, cfg->spp);
    cudaError_t e;
    if ((e = c.d_slices.ensure(pixels * chunks)) != cudaSuccess) return fail_cuda(e, "alloc slices");
    if (moments && (e = c.d_slices_sq.ensure(pixels * chunks)) != cudaSuccess) return fail_cuda(e, "alloc slices_sq");
    if ((e = c.d_counter.ensure(16)) != cudaSuccess) return fail_cuda(e, "alloc counter");
    if ((e = cudaMemsetAsync(c.d_counter.ptr, 0, 16 * sizeof(unsigned long long), stream)) != cudaSuccess) return fail_cuda(e, "memset");

    RenderLaunch r;
    r.sc = c.dscene;
    const bool twin = cfg->tracer_mode == MCPT_TRACER_CUDA_TWIN;
    r.cam = make_camera(cfg->eye, cfg->at, cfg->up, cfg->fovy, cfg->width, cfg->height, twin ? 1 : 0);
    r.width = cfg->width; r.height = cfg->height;
    r.sample_begin = cfg->sample_offset; r.sample_count = cfg->spp; r.chunks = chunks;
    // rtx.hlsl:313-328: roulette from `depth`, kill at 3*depth.  CUTracer.cu:103-176: `depth` scattering events,
    // then one gather, no roulette.
    r.depth = twin ? 0xFFFFFFFFu : static_cast<uint32_t>(cfg->max_depth);
    r.kill_at = twin ? 0xFFFFFFFFu : 3u * static_cast<uint32_t>(cfg->max_depth);
    r.gather_at = twin ? static_cast<uint32_t>(cfg->max_depth) : 0xFFFFFFFFu;
    r.twin = twin ? 1u : 0u;
    r.seed = cfg->seed;
    r.jitter = (cfg->flags & MCPT_FLAG_NO_JITTER) ? 0u : 1u;
    r.emission_scale = cfg->emission_scale;
    r.slices = c.d_slices.ptr;
    r.slices_sq = moments ? c.d_slices_sq.ptr : nullptr;
    r.frame_seeds = nullptr;
    r.ray_counter = c.d_counter.ptr;
    r.rng_mode = cfg->rng_mode;
    r.kd = (cfg->flags & MCPT_FLAG_REF_TRAVERSAL) == 0;
    r.nee = (cfg->flags & MCPT_FLAG_NEE) != 0;
    r.moments = moments;
    r.state_machine = (cfg->flags & MCPT_FLAG_SIMPLE_KERNEL) == 0;
    // Knobs come from mcpt_set_tuning only (never the environment) and are clamped there AND here: a vote threshold
    // above the warp size could never be met and render_kernel_sm would spin for ever.
    const Tuning& tn = tuning();
    r.shade_threshold = std::min(32, std::max(1, tn.shade_threshold));
    r.cam_threshold = std::min(32, std::max(1, tn.cam_threshold));
    r.tri_cap = std::min(64, std::max(1, tn.tri_cap));
    r.min_blocks = tn.min_blocks > 0 ? tn.min_blocks : (c.large_scene ? 8 : 4);   // launch.cu: residency
    r.pool = 0; r.pool_state = nullptr; r.sm_count = c.sm_count;
    if (tn.pool > 0 && !moments && !(cfg->flags & (MCPT_FLAG_NEE | MCPT_FLAG_REF_TRAVERSAL | MCPT_FLAG_SIMPLE_KERNEL)) && c.sm_count > 0) {
        if ((e = c.d_pool.ensure(pool_state_floats(c.sm_count))) != cudaSuccess) return fail_cuda(e, "alloc pool state");
        r.pool = tn.pool; r.pool_state = c.d_pool.ptr;
    }
    r.variant = tn.variant;         // MCPT_TUNE builds only
    if (cfg->rng_mode == MCPT_RNG_REFERENCE) {
        // GraphicsRTX.cpp:170-171,192: one draw of uniform_int_distribution<unsigned>(mt19937(1234))
        // per frame, i.e. the raw generator output.
        std::vector<uint32_t> seeds(static_cast<size_t>(cfg->spp));
        std::mt19937 gen(1234);
        gen.discard(static_cast<unsigned long long>(cfg->sample_offset));
        for (auto& s : seeds) s = static_cast<uint32_t>(gen());
        if ((e = c.d_frame_seeds.ensure(seeds.size())) != cudaSuccess) return fail_cuda(e, "alloc seeds");
        if ((e = cudaMemcpyAsync(c.d_frame_seeds.ptr, seeds.data(), seeds.size() * 4, cudaMemcpyHostToDevice, stream)) != cudaSuccess)
            return fail_cuda(e, "upload seeds");
        if ((e = cudaStreamSynchronize(stream)) != cudaSuccess) return fail_cuda(e, "upload seeds");
        r.frame_seeds = c.d_frame_seeds.ptr;
    }

    c.counters.launches = 0;
    cudaEventRecord(c.ev[0], stream);
    if ((e = launch_render(r, stream, &c.counters.launches)) != cudaSuccess) return fail_cuda(e, "render kernel launch");
    cudaEventRecord(c.ev[1], stream);

    if (dst) {
        if ((e = launch_reduce_slices(c.d_slices.ptr, chunks, pixels, dst, true, stream, &c.counters.launches)) != cudaSuccess) return fail_cuda(e, "reduce");
        if (moments && dstSq &&
            (e = launch_reduce_slices(c.d_slices_sq.ptr, chunks, pixels, dstSq, true, stream, &c.counters.launches)) != cudaSuccess) return fail_cuda(e, "reduce sq");
        if (sum) *sum = dst;
        if (sumsq) *sumsq = dstSq;
    } else if (chunks > 1) {
        if ((e = c.d_sum.ensure(pixels)) != cudaSuccess) return fail_cuda(e, "alloc sum");
        if ((e = launch_reduce_slices(c.d_slices.ptr, chunks, pixels, c.d_sum.ptr, false, stream, &c.counters.launches)) != cudaSuccess) return fail_cuda(e, "reduce");
        if (moments) {
            if ((e = c.d_sumsq.ensure(pixels)) != cudaSuccess) return fail_cuda(e, "alloc sumsq");
            if ((e = launch_reduce_slices(c.d_slices_sq.ptr, chunks, pixels, c.d_sumsq.ptr, false, stream, &c.counters.launches)) != cudaSuccess) return fail_cuda(e, "reduce sq");
        }
        if (sum) *sum = c.d_sum.ptr;
        if (sumsq) *sumsq = moments ? c.d_sumsq.ptr : nullptr;
    } else {
        if (sum) *sum = c.d_slices.ptr;
        if (sumsq) *sumsq = moments ? c.d_slices_sq.ptr : nullptr;
    }
    c.counters.paths = static_cast<uint64_t>(pixels) * static_cast<uint64_t>(cfg->spp);
    return MCPT_OK;
}

int finish_counters(cudaStream_t stream) {
    Context& c = ctx();
    cudaError_t e;
    unsigned long long cnt[16] = {0};
    if ((e = cudaMemcpyAsync(cnt, c.d_counter.ptr, sizeof(cnt), cudaMemcpyDeviceToHost, stream)) != cudaSuccess) return fail_cuda(e, "counter readback");
    cudaEventRecord(c.ev[2], stream);
    if ((e = cudaStreamSynchronize(stream)) != cudaSuccess) return fail_cuda(e, "render");
    float k = 0, t = 0;
    cudaEventElapsedTime(&k, c.ev[0], c.ev[1]);
    cudaEventElapsedTime(&t, c.ev[0], c.ev[2]);
    const unsigned long long rays = cnt[0];
    c.counters.rays = rays;
    if (cnt[1] && tuning().sm_stats) {  // tuning builds, VAR bit 3
        std::fprintf(stderr, "rounds %llu lanes/round %.2f | node steps: mean/lane %.2f max %.2f | tri tests: mean/lane %.2f max %.2f | shade phases %llu lanes/phase %.2f done/phase %.2f | rounds/phase %.2f | raw3 %llu raw4 %llu\n",
                     cnt[1], double(cnt[2]) / cnt[1], double(cnt[3]) / cnt[2], double(cnt[4]) / cnt[1], double(cnt[5]) / cnt[2], double(cnt[6]) / cnt[1],
                     cnt[7], double(cnt[8]) / cnt[7], double(cnt[9]) / cnt[7], double(cnt[1]) / cnt[7], cnt[3], cnt[4]);
        std::fprintf(stderr, "rays %llu | rounds/ray %.3f (warp rounds x 32 / rays) | (lane, triangle) pairs per round %.2f, round's longest triangle loop %.2f -> packed iterations ceil(P/32) ~ %.2f | point locations per ray %.3f, steps per location %.2f, rounds with a location %.1f %%, longest walk in such a round %.2f\n",
                     rays, 32.0 * cnt[1] / rays, double(cnt[5]) / cnt[1], double(cnt[6]) / cnt[1], double(cnt[5]) / cnt[1] / 32.0,
                     double(cnt[10]) / rays, cnt[10] ? double(cnt[11]) / cnt[10] : 0.0, 100.0 * cnt[12] / cnt[1], cnt[12] ? double(cnt[13]) / cnt[12] : 0.0);
    }
    c.counters.kernel_ms = k;
    c.counters.total_ms = t;
    return MCPT_OK;
}

}  // namespace mcpt

extern "C" {

void mcpt_default_render_cfg(mcpt_render_cfg* cfg, int width, int height, int spp, int max_depth) {
    if (!cfg) return;
    std::memset(cfg, 0, sizeof(*cfg));
    cfg->width = width; cfg->height = height; cfg->spp = spp; cfg->max_depth = max_depth;
    cfg->sample_offset = 0; cfg->seed = 0; cfg->rng_mode = MCPT_RNG_PHILOX; cfg->flags = 0;
    cfg->eye[0] = 0; cfg->eye[1] = 5; cfg->eye[2] = 17;  // GraphicsRTX.cpp:175-177
    cfg->at[0] = 0; cfg->at[1] = 5; cfg->at[2] = 16;
    cfg->up[0] = 0; cfg->up[1] = 1; cfg->up[2] = 0;
    cfg->fovy = 3.141592654f / 4.0f;  // float(D3DX_PI) / 4, GraphicsRTX.cpp:183
    cfg->emission_scale = 1.0f;
}

void mcpt_cuda_twin_render_cfg(mcpt_render_cfg* cfg, int width, int height, int spp, int scene_id) {
    if (!cfg) return;
    mcpt_default_render_cfg(cfg, width, height, spp, 7);                 // depth literal, CUTracer.cu:212
    cfg->tracer_mode = MCPT_TRACER_CUDA_TWIN;
    cfg->eye[2] = scene_id == 1 ? 17.0f : 23.0f;                        // CUTracer.cu:349,363
    cfg->at[2] = cfg->eye[2] - 1.0f;
    cfg->fovy = 60.0f * 3.14159265359f / 180.0f;                         // projFOV = 60, PW_PI (stdafx.h:46)
    cfg->emission_scale = 10.0f;                                         // ILLUM (stdafx.h:45)
}

int mcpt_render_scene(const mcpt_render_cfg* cfg, float* host_rgb, float* host_m2) {
    return guarded([&]() -> int {
    Context& c = ctx();
    std::lock_guard<std::mutex> g(c.lock);
    if (!host_rgb) return fail(MCPT_ERR_INVALID, "null host_rgb");
    if (host_m2 && !(cfg && (cfg->flags & MCPT_FLAG_MOMENTS))) return fail(MCPT_ERR_INVALID, "host_m2 needs MCPT_FLAG_MOMENTS");
    const float4 *sum = nullptr, *sumsq = nullptr;
    int rc = render_core(cfg, c.stream, nullptr, nullptr, &sum, &sumsq);
    if (rc) return rc;
    const size_t pixels = static_cast<size_t>(cfg->width) * cfg->height;
    cudaError_t e;
    if ((e = c.d_rgb.ensure(3 * pixels)) != cudaSuccess) return fail_cuda(e, "alloc rgb");
    if ((e = launch_finalize(sum, pixels, c.d_rgb.ptr, c.stream, &c.counters.launches)) != cudaSuccess) return fail_cuda(e, "finalize");
    if ((e = cudaMemcpyAsync(host_rgb, c.d_rgb.ptr, 3 * pixels * sizeof(float), cudaMemcpyDeviceToHost, c.stream)) != cudaSuccess) return fail_cuda(e, "D2H");
    if (host_m2) {
        if ((e = cudaStreamSynchronize(c.stream)) != cudaSuccess) return fail_cuda(e, "render");
        if ((e = launch_finalize(sumsq, pixels, c.d_rgb.ptr, c.stream, &c.counters.launches)) != cudaSuccess) return fail_cuda(e, "finalize m2");
        if ((e = cudaMemcpyAsync(host_m2, c.d_rgb.ptr, 3 * pixels * sizeof(float), cudaMemcpyDeviceToHost, c.stream)) != cudaSuccess) return fail_cuda(e, "D2H m2");
    }
    return finish_counters(c.stream);
    });
}

int mcpt_render_accumulate_device(const mcpt_render_cfg* cfg, void* dev_sum, void* dev_sumsq, void* stream) {
    return guarded([&]() -> int {
    Context& c = ctx();
    std::lock_guard<std::mutex> g(c.lock);
    if (!dev_sum) return fail(MCPT_ERR_INVALID, "null dev_sum");
    cudaStream_t s = stream ? static_cast<cudaStream_t>(stream) : c.stream;
    int rc = render_core(cfg, s, static_cast<float4*>(dev_sum), static_cast<float4*>(dev_sumsq), nullptr, nullptr);
    if (rc) return rc;
    return finish_counters(s);
    });
}

int mcpt_intersect_batch_device(const void* dev_rays, int64_t n, int variant, void* dev_tri, void* dev_t,
                                void* dev_beta, void* dev_gamma, void* stream) {
    return guarded([&]() -> int {
    Context& c = ctx();
    std::lock_guard<std::mutex> g(c.lock);
    if (n < 0 || (n > 0 && (!dev_rays || !dev_tri))) return fail(MCPT_ERR_INVALID, "bad ray batch");
    if (variant != MCPT_ISECT_REFERENCE && variant != MCPT_ISECT_KD && variant != MCPT_ISECT_ROPES) return fail(MCPT_ERR_INVALID, "unknown variant");
    int rc = variant == MCPT_ISECT_ROPES ? ensure_uploaded() : ensure_parity_uploaded();
    if (rc) return rc;
    cudaStream_t s = stream ? static_cast<cudaStream_t>(stream) : c.stream;
    c.counters.launches = 0;
    cudaEventRecord(c.ev[0], s);
    cudaError_t e = launch_intersect(c.dscene, static_cast<const float*>(dev_rays), n, variant, static_cast<int32_t*>(dev_tri),
                                     static_cast<float*>(dev_t), static_cast<float*>(dev_beta), static_cast<float*>(dev_gamma),
                                     s, &c.counters.launches);
    if (e != cudaSuccess) return fail_cuda(e, "intersect launch");
    cudaEventRecord(c.ev[1], s);
    if ((e = cudaStreamSynchronize(s)) != cudaSuccess) return fail_cuda(e, "intersect");
    float k = 0;
    cudaEventElapsedTime(&k, c.ev[0], c.ev[1]);
    c.counters.paths = 0; c.counters.rays = static_cast<uint64_t>(n); c.counters.kernel_ms = k; c.counters.total_ms = k;
    return MCPT_OK;
    });
}

int mcpt_intersect_batch(const float* rays, int64_t n, int variant, int32_t* tri, float* t, float* beta, float* gamma) {
    return guarded([&]() -> int {
    Context& c = ctx();
    {
        std::lock_guard<std::mutex> g(c.lock);
        if (n < 0 || (n > 0 && (!rays || !tri))) return fail(MCPT_ERR_INVALID, "bad ray batch");
        int rc = ensure_uploaded();
        if (rc) return rc;
        if (n == 0) return MCPT_OK;
        const size_t N = static_cast<size_t>(n);
        cudaError_t e;
        if ((e = c.d_rays.ensure(6 * N)) != cudaSuccess || (e = c.d_tri.ensure(N)) != cudaSuccess || (e = c.d_t.ensure(N)) != cudaSuccess ||
            (e = c.d_beta.ensure(N)) != cudaSuccess || (e = c.d_gamma.ensure(N)) != cudaSuccess)
            return fail_cuda(e, "alloc ray batch");
        if ((e = cudaMemcpyAsync(c.d_rays.ptr, rays, 6 * N * sizeof(float), cudaMemcpyHostToDevice, c.stream)) != cudaSuccess) return fail_cuda(e, "H2D rays");
    }
    int rc = mcpt_intersect_batch_device(c.d_rays.ptr, n, variant, c.d_tri.ptr, c.d_t.ptr, c.d_beta.ptr, c.d_gamma.ptr, nullptr);
    if (rc) return rc;
    std::lock_guard<std::mutex> g(c.lock);
    const size_t N = static_cast<size_t>(n);
    cudaError_t e = cudaMemcpy(tri, c.d_tri.ptr, N * 4, cudaMemcpyDeviceToHost);
    if (e == cudaSuccess && t) e = cudaMemcpy(t, c.d_t.ptr, N * 4, cudaMemcpyDeviceToHost);
    if (e == cudaSuccess && beta) e = cudaMemcpy(beta, c.d_beta.ptr, N * 4, cudaMemcpyDeviceToHost);
    if (e == cudaSuccess && gamma) e = cudaMemcpy(gamma, c.d_gamma.ptr, N * 4, cudaMemcpyDeviceToHost);
    if (e != cudaSuccess) return fail_cuda(e, "D2H hits");
    return MCPT_OK;
    });
}

int mcpt_primary_rays(const mcpt_render_cfg* cfg, int sample, float* host_rays) {
    return guarded([&]() -> int {
    Context& c = ctx();
    std::lock_guard<std::mutex> g(c.lock);
    int rc = check_cfg(cfg);
    if (rc) return rc;
    if (!host_rays || sample < 0) return fail(MCPT_ERR_INVALID, "bad arguments");
    if (const int brc = bind_device()) return brc;
    const size_t pixels = static_cast<size_t>(cfg->width) * cfg->height;
    cudaError_t e;
    if ((e = c.d_rays.ensure(6 * pixels)) != cudaSuccess) return fail_cuda(e, "alloc rays");
    uint32_t frameSeed = 0;
    if (cfg->rng_mode == MCPT_RNG_REFERENCE) {
        std::mt19937 gen(1234);
        gen.discard(static_cast<unsigned long long>(sample));
        frameSeed = static_cast<uint32_t>(gen());
    }
    DCamera cam = make_camera(cfg->eye, cfg->at, cfg->up, cfg->fovy, cfg->width, cfg->height, cfg->tracer_mode == MCPT_TRACER_CUDA_TWIN ? 1 : 0);
    c.counters.launches = 0;
    if ((e = launch_primary_rays(cam, cfg->width, cfg->height, sample, cfg->seed, (cfg->flags & MCPT_FLAG_NO_JITTER) ? 0u : 1u,
                                 cfg->rng_mode, frameSeed, (cfg->flags & MCPT_FLAG_RENDER_ARITHMETIC) != 0, c.d_rays.ptr, c.stream,
                                 &c.counters.launches)) != cudaSuccess)
        return fail_cuda(e, "primary rays launch");
    if ((e = cudaMemcpyAsync(host_rays, c.d_rays.ptr, 6 * pixels * sizeof(float), cudaMemcpyDeviceToHost, c.stream)) != cudaSuccess) return fail_cuda(e, "D2H rays");
    if ((e = cudaStreamSynchronize(c.stream)) != cudaSuccess) return fail_cuda(e, "primary rays");
    return MCPT_OK;
    });
}

int mcpt_sample_batch(int kind, int64_t n, const float* u, const float* normals, const float* dirs, const float* par0,
                      const float* par1, int render_arithmetic, float* out) {
    return guarded([&]() -> int {
    Context& c = ctx();
    std::lock_guard<std::mutex> g(c.lock);
    if (kind < MCPT_SAMPLE_HEMI || kind > MCPT_SAMPLE_FRESNEL_TWIN) return fail(MCPT_ERR_INVALID, "unknown sampler kind");
    if (n < 0 || (n > 0 && (!u || !normals || !out))) return fail(MCPT_ERR_INVALID, "bad sampler batch");
    if (n > 0 && kind != MCPT_SAMPLE_HEMI && !dirs) return fail(MCPT_ERR_INVALID, "this sampler needs incoming directions");
    if (n > 0 && kind != MCPT_SAMPLE_HEMI && !par0) return fail(MCPT_ERR_INVALID, "missing sampler parameter (Ns / Tr)");
    if (n > 0 && kind >= MCPT_SAMPLE_FRESNEL && !par1) return fail(MCPT_ERR_INVALID, "missing sampler parameter (Ni)");
    if (const int brc = bind_device()) return brc;
    if (n == 0) return MCPT_OK;
    const size_t N = static_cast<size_t>(n);
    // one scratch buffer: u (2N) | normals (3N) | dirs (3N) | par0 (N) | par1 (N) | out (3N)
    cudaError_t e;
    if ((e = c.d_rays.ensure(13 * N)) != cudaSuccess) return fail_cuda(e, "alloc sampler batch");
    float* d = c.d_rays.ptr;
    float *du = d, *dn = d + 2 * N, *dd = d + 5 * N, *dp0 = d + 8 * N, *dp1 = d + 9 * N, *dout = d + 10 * N;
    auto up = [&](float* dst, const float* src, size_t cnt) {
        return src ? cudaMemcpyAsync(dst, src, cnt * sizeof(float), cudaMemcpyHostToDevice, c.stream) : cudaSuccess;
    };
    if ((e = up(du, u, 2 * N)) != cudaSuccess || (e = up(dn, normals, 3 * N)) != cudaSuccess || (e = up(dd, dirs, 3 * N)) != cudaSuccess ||
        (e = up(dp0, par0, N)) != cudaSuccess || (e = up(dp1, par1, N)) != cudaSuccess)
        return fail_cuda(e, "H2D sampler batch");
    c.counters.launches = 0;
    if ((e = launch_sample(kind, n, du, dn, dirs ? dd : nullptr, dp0, dp1, render_arithmetic != 0, dout, c.stream, &c.counters.launches)) != cudaSuccess)
        return fail_cuda(e, "sampler launch");
    if ((e = cudaMemcpyAsync(out, dout, 3 * N * sizeof(float), cudaMemcpyDeviceToHost, c.stream)) != cudaSuccess) return fail_cuda(e, "D2H");
    if ((e = cudaStreamSynchronize(c.stream)) != cudaSuccess) return fail_cuda(e, "sampler batch");
    return MCPT_OK;
    });
}

int mcpt_traversal_stats(const float* rays, int64_t n, int variant, uint64_t out[3]) {
    return guarded([&]() -> int {
    Context& c = ctx();
    std::lock_guard<std::mutex> g(c.lock);
    if (n <= 0 || !rays || !out) return fail(MCPT_ERR_INVALID, "bad ray batch");
    int rc = variant == MCPT_ISECT_ROPES ? ensure_uploaded() : ensure_parity_uploaded();
    if (rc) return rc;
    const size_t N = static_cast<size_t>(n);
    cudaError_t e;
    if ((e = c.d_rays.ensure(6 * N)) != cudaSuccess || (e = c.d_counter.ensure(4)) != cudaSuccess) return fail_cuda(e, "alloc");
    if ((e = cudaMemcpyAsync(c.d_rays.ptr, rays, 6 * N * sizeof(float), cudaMemcpyHostToDevice, c.stream)) != cudaSuccess) return fail_cuda(e, "H2D rays");
    if ((e = cudaMemsetAsync(c.d_counter.ptr, 0, 4 * sizeof(unsigned long long), c.stream)) != cudaSuccess) return fail_cuda(e, "memset");
    if ((e = launch_tally(c.dscene, c.d_rays.ptr, n, variant, c.d_counter.ptr, c.stream)) != cudaSuccess) return fail_cuda(e, "tally launch");
    unsigned long long h[3];
    if ((e = cudaMemcpyAsync(h, c.d_counter.ptr, sizeof(h), cudaMemcpyDeviceToHost, c.stream)) != cudaSuccess) return fail_cuda(e, "D2H");
    if ((e = cudaStreamSynchronize(c.stream)) != cudaSuccess) return fail_cuda(e, "tally");
    out[0] = h[0]; out[1] = h[1]; out[2] = h[2];
    return MCPT_OK;
    });
}

int mcpt_measure_fp32_peak(double* gflops) {
    return guarded([&]() -> int {
    Context& c = ctx();
    std::lock_guard<std::mutex> g(c.lock);
    if (!gflops) return fail(MCPT_ERR_INVALID, "null gflops");
    if (const int brc = bind_device()) return brc;
    const int blocks = 148 * 16, iters = 1 << 14;
    cudaError_t e;
    if ((e = c.d_rgb.ensure(static_cast<size_t>(blocks) * 256)) != cudaSuccess) return fail_cuda(e, "alloc");
    double best = 0.0;
    for (int rep = 0; rep < 5; ++rep) {
        cudaEventRecord(c.ev[0], c.stream);
        if ((e = launch_fp32_peak(c.d_rgb.ptr, blocks, iters, c.stream)) != cudaSuccess) return fail_cuda(e, "fp32 probe");
        cudaEventRecord(c.ev[1], c.stream);
        if ((e = cudaStreamSynchronize(c.stream)) != cudaSuccess) return fail_cuda(e, "fp32 probe");
        float ms = 0;
        cudaEventElapsedTime(&ms, c.ev[0], c.ev[1]);
        if (rep > 0) best = std::max(best, static_cast<double>(blocks) * 256 * iters * 16 / (ms * 1e6));
    }
    *gflops = best;
    return MCPT_OK;
    });
}

int mcpt_selfcheck_div3(int64_t n, uint32_t seed, int spread, uint64_t* mismatches, uint64_t* fast_path) {
    return guarded([&]() -> int {
    Context& c = ctx();
    std::lock_guard<std::mutex> g(c.lock);
    if (!mismatches || !fast_path || n < 0 || spread < 0 || spread > 126) return fail(MCPT_ERR_INVALID, "bad self-check arguments");
    if (const int brc = bind_device()) return brc;
    cudaError_t e;
    if ((e = c.d_counter.ensure(16)) != cudaSuccess) return fail_cuda(e, "alloc counters");
    if ((e = cudaMemsetAsync(c.d_counter.ptr, 0, 16 * sizeof(unsigned long long), c.stream)) != cudaSuccess) return fail_cuda(e, "memset");
    if ((e = launch_selfcheck_div3(n, seed, spread, c.d_counter.ptr, c.stream)) != cudaSuccess) return fail_cuda(e, "self-check");
    unsigned long long h[2];
    if ((e = cudaMemcpyAsync(h, c.d_counter.ptr, sizeof(h), cudaMemcpyDeviceToHost, c.stream)) != cudaSuccess) return fail_cuda(e, "D2H");
    if ((e = cudaStreamSynchronize(c.stream)) != cudaSuccess) return fail_cuda(e, "self-check");
    *mismatches = h[0]; *fast_path = h[1];
    return MCPT_OK;
    });
}

int mcpt_get_counters(mcpt_counters* out) {
    return guarded([&]() -> int {
    if (!out) return fail(MCPT_ERR_INVALID, "null counters");
    *out = ctx().counters;
    return MCPT_OK;
    });
}

}  // extern "C"
