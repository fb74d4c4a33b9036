// The path-sampling kernel (rtx.hlsl:304-405 re-designed for sm_100a) and its small companions.
//
// One thread owns one pixel for a contiguous range of samples and keeps the running sums in
// registers (one float4 store per pixel per launch instead of the reference's read-modify-write
// of a gamma-encoded texture every sample, rtx.hlsl:401-404).  The loop is flat: every iteration
// traces exactly ONE ray per lane -- a path segment or a pending shadow ray -- and a lane whose
// path ended regenerates the next camera ray in place, so the warp re-converges at the single
// intersect call site however different the lanes' path lengths are.
#pragma once
#include "intersect.cuh"
#include "shading.cuh"

namespace mcpt {

struct RenderParams {
    DScene sc;
    DCamera cam;
    int width, height;
    int sample_begin;      // absolute index of the first sample of this launch
    int sample_count;      // samples per pixel in this launch
    int chunks;            // sample range is split over gridDim.z chunks, one output slice each
    uint32_t depth;        // roulette from this bounce on (the shader's `depth`, rtx.hlsl:317); 0xFFFFFFFF = never
    uint32_t kill_at;      // a hit at this bounce index returns 0 (3 * depth, rtx.hlsl:313)
    uint32_t gather_at;    // CUDA twin: at this bounce index only emission is gathered, no scattering (CUTracer.cu:162-175)
    uint32_t twin;         // CUDA twin shading rules: glass tinted by Kd, Ns truncated to unsigned
    uint32_t seed;
    uint32_t jitter;
    int cam_threshold;     // render_kernel_sm, camera-ray queue: lanes that must need a refill before the warp refills
    int tri_cap;           // render_kernel_sm, VAR bit 6: triangle tests per lane and round
    float emission_scale;
    float4* sum;           // [chunks][height][width]  r,g,b sums + sample count
    float4* sumsq;         // same layout, or nullptr
    const uint32_t* frame_seeds;  // reference stream: one seed per sample of this launch
    unsigned long long* ray_counter;
    float* pool_state;     // render_kernel_pool: per resident warp, PG_WORDS x (32 * SETS) words of path state (stays in L2)
};

// rtx.hlsl:384-398.  EXACT: the shader's expressions with IEEE divides and square root (primary_rays_kernel: rays
// bit-equal to the oracle's).  The render kernels (MCPT_FAST_SHADE) use the same map with the two quotients per
// coordinate folded into one host-computed factor and the hardware rsqrt: a few ulps in the ray, ~45 instructions less.
template <bool EXACT = true>
__device__ __forceinline__ void camera_ray(const DCamera& cam, int width, int height, uint32_t px, uint32_t py,
                                           float u1, float u2, F3& org, F3& dir) {
    if (cam.twin) {
        // CUTracer.cu:195-212: +-1 px, horizontal fov, the 2.0 / 1.0 literals make this double arithmetic
        const float bx = static_cast<float>(px) + (u1 * 2.0f - 1.0f);
        const float by = static_cast<float>(py) + (u2 * 2.0f - 1.0f);
        const float t0 = static_cast<float>((2.0 * bx / width - 1) * cam.tan_half);
        const float t1 = static_cast<float>((1.0 * height / width - 2.0 * by / width) * cam.tan_half);
        F3 w;
        w.x = cam.inv[0][0] * t0 + cam.inv[1][0] * t1 - cam.inv[2][0];
        w.y = cam.inv[0][1] * t0 + cam.inv[1][1] * t1 - cam.inv[2][1];
        w.z = cam.inv[0][2] * t0 + cam.inv[1][2] * t1 - cam.inv[2][2];
        const float len = sqrtf(w.x * w.x + w.y * w.y + w.z * w.z);
        if (fabsf(len) > MCPT_EPS) { w.x /= len; w.y /= len; w.z /= len; }  // Utils.hpp:27-34
        dir = w;
        org = f3(cam.inv[3][0], cam.inv[3][1], cam.inv[3][2]);
        return;
    }
    const float biasx = static_cast<float>(px) + (u1 - 0.5f);
    const float biasy = static_cast<float>(py) + (u2 - 0.5f);
    float d0, d1;
    if (EXACT) {
        d0 = (2.0f * biasx / static_cast<float>(width) - 1) / cam.p11;
        d1 = (1 - 2.0f * biasy / static_cast<float>(height)) / cam.p22;
    } else {
        d0 = (biasx - 0.5f * static_cast<float>(width)) * cam.sx;
        d1 = (0.5f * static_cast<float>(height) - biasy) * cam.sy;
    }
    const float d2 = -1.0f;
    float w[3], f[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        w[i] = d0 * cam.inv[0][i] + d1 * cam.inv[1][i] + d2 * cam.inv[2][i];
        f[i] = w[i] + cam.inv[3][i];
    }
    org = f3(f[0], f[1], f[2]);
    const float len2 = w[0] * w[0] + w[1] * w[1] + w[2] * w[2];
    const float inv = EXACT ? 1.0f / sqrtf(len2) : shade_rsqrt(len2);
    dir = f3(w[0] * inv, w[1] * inv, w[2] * inv);
}

template <bool KD>
__device__ __forceinline__ Hit trace(const DScene& sc, F3 o, F3 d) {
    if (KD) return intersect_kd(sc, o.x, o.y, o.z, d.x, d.y, d.z);
    return intersect_ref(sc, o.x, o.y, o.z, d.x, d.y, d.z);
}

// Pixel of a thread: each warp covers an 8x4 tile, a block of BLOCK threads a 16 x (BLOCK/16) tile.
template <int BLOCK = 256>
__device__ __forceinline__ void thread_pixel(uint32_t& px, uint32_t& py) {
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    px = blockIdx.x * 16u + (warp & 1u) * 8u + (lane & 7u);
    py = blockIdx.y * static_cast<uint32_t>(BLOCK / 16) + (warp >> 1) * 4u + (lane >> 3);
}

template <class RNG, bool KD, bool NEE, bool MOMENTS>
__global__ void __launch_bounds__(256) render_kernel(const RenderParams p) {
    uint32_t px, py;
    thread_pixel(px, py);
    const bool inside = px < static_cast<uint32_t>(p.width) && py < static_cast<uint32_t>(p.height);
    const uint32_t pixel = py * static_cast<uint32_t>(p.width) + px;

    // this chunk's share of the sample range
    const int chunk = blockIdx.z;
    const int base = p.sample_count / p.chunks, rem = p.sample_count % p.chunks;
    const int local0 = chunk * base + min(chunk, rem);
    const int localN = base + (chunk < rem ? 1 : 0);
    int s = local0;
    const int sEnd = local0 + localN;

    RNG rng;
    F3 org = f3(0, 0, 0), dir = f3(0, 0, 1), thr = f3(1, 1, 1), cur = f3(0, 0, 0);
    F3 shOrg = f3(0, 0, 0), shDir = f3(0, 0, 1), shVal = f3(0, 0, 0);
    uint32_t shTri = 0, bounce = 0, nrays = 0;
    bool shadow = false, lastWasNee = false;
    F3 acc = f3(0, 0, 0), acc2 = f3(0, 0, 0);

    auto start_sample = [&]() {
        rng.start(pixel, p.seed, static_cast<uint32_t>(p.sample_begin + s), p.frame_seeds ? p.frame_seeds[s] : 0u);
        float u1 = 0.5f, u2 = 0.5f;
        if (p.jitter) { u1 = rng.next(); u2 = rng.next(); }
        camera_ray<!MCPT_FAST_SHADE>(p.cam, p.width, p.height, px, py, u1, u2, org, dir);
        thr = f3(1, 1, 1); cur = f3(0, 0, 0);
        bounce = 0; lastWasNee = false;
    };

    bool alive = inside && s < sEnd;
    if (alive) start_sample();
    while (alive) {
        const Hit h = trace<KD>(p.sc, shadow ? shOrg : org, shadow ? shDir : dir);
        ++nrays;
        if (NEE && shadow) {
            if (h.tri == shTri) cur = cur + shVal;
            shadow = false;
            continue;
        }
        bool done = false;
        if (h.tri == 0xFFFFFFFFu) {
            done = true;  // rtx.hlsl:370
        } else {
            rng.set_slot(bounce + 1u);
            if (bounce >= p.kill_at) {
                done = true;  // :313-316
            } else if (bounce >= p.depth) {
                const float illum = fmaxf(fmaxf(thr.x, thr.y), thr.z);  // :317-328
                if (illum > rng.next()) thr = MCPT_FAST_SHADE ? thr * shade_rcp(illum) : thr / illum; else done = true;
            }
            if (!done) {
                const float4 s0 = ldg4(p.sc.tri_shade + 3 * h.tri);
                const uint32_t mat = __float_as_uint(s0.w);
                const float4 mKa = ldg4(p.sc.mats + 3 * mat);
                if (mKa.x > 0 || mKa.y > 0 || mKa.z > 0) {
                    // :329-334; with NEE the light was already gathered at the previous Lambert vertex
                    if (!(NEE && lastWasNee)) cur = cur + thr * f3(mKa.x, mKa.y, mKa.z) * p.emission_scale;
                    done = true;
                } else if (bounce >= p.gather_at) {
                    done = true;  // CUDA twin: the last intersect only gathers emission (CUTracer.cu:162-175)
                } else {
                    const float4 s1 = ldg4(p.sc.tri_shade + 3 * h.tri + 1), s2 = ldg4(p.sc.tri_shade + 3 * h.tri + 2);
                    const float4 mKd = ldg4(p.sc.mats + 3 * mat + 1);
                    F3 normal = f3(s0.x, s0.y, s0.z) * (1.0f - h.beta - h.gamma) + f3(s1.x, s1.y, s1.z) * h.beta +
                                f3(s2.x, s2.y, s2.z) * h.gamma;  // :336-340
                    normal = normalize(normal);
                    const F3 hp = f3(org.x + h.t * dir.x, org.y + h.t * dir.y, org.z + h.t * dir.z);  // :197-199
                    lastWasNee = false;
                    if (mKd.w > 0) {  // Tr > 0, :342-346
                        const float4 mKs = ldg4(p.sc.mats + 3 * mat + 2);
                        dir = sample_fresnel(rng, normal, dir, mKd.w, mKs.w, p.twin != 0u);  // twin: Utils.hpp:97-137
                        if (p.twin) thr = thr * f3(mKd.x, mKd.y, mKd.z);  // CUTracer.cu:131-133
                    } else if (mKa.w > 1) {  // Ns > 1, :348-352
                        const float4 mKs = ldg4(p.sc.mats + 3 * mat + 2);
                        dir = sample_phong(rng, normal, dir, p.twin ? static_cast<float>(static_cast<uint32_t>(mKa.w)) : mKa.w);  // Utils.hpp:72
                        thr = thr * f3(mKs.x, mKs.y, mKs.z);
                    } else {  // :354-365
                        const bool flip = dot(dir, normal) > 0;
                        dir = flip ? -sample_hemi(rng, normal) : sample_hemi(rng, normal);
                        thr = thr * f3(mKd.x, mKd.y, mKd.z);
                        if (NEE && p.sc.num_emitters > 0 && bounce + 1u < p.kill_at) {
                            // Next-event estimation of what the reference would collect at bounce+1:
                            // thr*Kd * Ka * (cos_x / pi) * |cos_y| * A / r^2, visible iff the first hit from
                            // hp + 0.01 w is the sampled emitter triangle.
                            rng.set_slot(0x80000000u | (bounce + 1u));
                            const float u0 = rng.next(), ub = rng.next(), uc = rng.next();
                            uint32_t e = 0;
                            while (e + 1 < p.sc.num_emitters && u0 > ldg4(p.sc.emitters + 3 * e + 1).w) ++e;
                            const float4 ea = ldg4(p.sc.emitters + 3 * e), eu = ldg4(p.sc.emitters + 3 * e + 1),
                                         ev = ldg4(p.sc.emitters + 3 * e + 2);
                            const float su = shade_sqrt(ub);
                            const float b1 = 1.0f - su, b2 = uc * su;  // uniform on the triangle
                            const F3 y = f3(ea.x + b1 * eu.x + b2 * ev.x, ea.y + b1 * eu.y + b2 * ev.y, ea.z + b1 * eu.z + b2 * ev.z);
                            F3 w = y - hp;
                            const float r2 = dot(w, w);
                            w = w * shade_rsqrt(r2);
                            const F3 ng = f3(eu.y * ev.z - eu.z * ev.y, eu.z * ev.x - eu.x * ev.z, eu.x * ev.y - eu.y * ev.x);
                            const float cosy = fabsf(dot(ng, w)) * shade_rcp(2.0f * ea.w);  // |ng| = 2 * area
                            const F3 nf = flip ? -normal : normal;
                            const float cosx = dot(nf, w);
                            if (cosx > 0 && cosy > 0 && r2 > 0) {
                                shTri = __float_as_uint(ev.w);  // id of the sampled emitting triangle
                                const uint32_t lmat = __float_as_uint(ldg4(p.sc.tri_shade + 3 * shTri).w);
                                const float4 lKa = ldg4(p.sc.mats + 3 * lmat);
                                const float g = cosx * cosy * p.sc.emitter_area * shade_rcp(MCPT_PI * r2);
                                shVal = thr * f3(lKa.x, lKa.y, lKa.z) * (g * p.emission_scale);
                                shOrg = hp + w * 0.01f;
                                shDir = w;
                                shadow = true;
                            }
                            lastWasNee = true;
                        }
                    }
                    org = hp + dir * 0.01f;  // :366
                    ++bounce;
                }
            }
        }
        if (done) {
            acc = acc + cur;
            if (MOMENTS) acc2 = acc2 + cur * cur;
            ++s;
            shadow = false;
            if (s < sEnd) start_sample(); else alive = false;
        }
    }

    if (inside) {
        const size_t o = (static_cast<size_t>(chunk) * p.height + py) * p.width + px;
        p.sum[o] = make_float4(acc.x, acc.y, acc.z, static_cast<float>(localN));
        if (MOMENTS) p.sumsq[o] = make_float4(acc2.x, acc2.y, acc2.z, static_cast<float>(localN));
    }
    // one 64-bit atomic per warp for the ray count
    for (int off = 16; off > 0; off >>= 1) nrays += __shfl_xor_sync(0xFFFFFFFFu, nrays, off);
    if ((threadIdx.x & 31u) == 0u && nrays) atomicAdd(p.ray_counter, static_cast<unsigned long long>(nrays));
}

}  // namespace mcpt
