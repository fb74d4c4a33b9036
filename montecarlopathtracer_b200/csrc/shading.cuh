// Random streams and BSDF direction samplers (device).
//
// Samplers follow rtx.hlsl:213-302 statement by statement (same special cases, same operand
// order); this translation unit is compiled with -fmad=false and without fast-math.  The differences
// from the CPU oracle are a few ulps per direction: libm-vs-CUDA pow/sin/cos, and (MCPT_FAST_SHADE,
// default) the hardware's approximate reciprocal square root / square root (MUFU.RSQ / MUFU.SQRT,
// <= 2 ulp) and sincospif in place of the IEEE sequences -- HLSL's own rsqrt/sqrt/sincos carry no
// tighter guarantee than that.  The intersection arithmetic (intersect.cuh) is not touched by this.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

namespace mcpt {

struct F3 {
    float x, y, z;
};
__device__ __forceinline__ F3 f3(float x, float y, float z) { return {x, y, z}; }
__device__ __forceinline__ F3 operator+(F3 a, F3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
__device__ __forceinline__ F3 operator-(F3 a, F3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
__device__ __forceinline__ F3 operator*(F3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
__device__ __forceinline__ F3 operator*(F3 a, F3 b) { return {a.x * b.x, a.y * b.y, a.z * b.z}; }
__device__ __forceinline__ F3 operator/(F3 a, float s) { return {a.x / s, a.y / s, a.z / s}; }
__device__ __forceinline__ F3 operator-(F3 a) { return {-a.x, -a.y, -a.z}; }
__device__ __forceinline__ float dot(F3 a, F3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
#ifndef MCPT_FAST_SHADE
#define MCPT_FAST_SHADE 1
#endif
// 1/sqrt(x) and sqrt(x) for shading directions: one MUFU instruction each instead of ~10-20 (IEEE sqrt + IEEE divide).
// FAST is a template parameter (default: the build's MCPT_FAST_SHADE) so that the sampler parity hook
// (mcpt_sample_batch) can run both forms from one library.
constexpr bool kFastShade = MCPT_FAST_SHADE != 0;
template <bool FAST = kFastShade>
__device__ __forceinline__ float shade_rsqrt(float x) {
    if (FAST) {
        float y;
        asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
        return y;
    }
    return __fdiv_rn(1.0f, __fsqrt_rn(x));
}
template <bool FAST = kFastShade>
__device__ __forceinline__ float shade_sqrt(float x) {
    if (FAST) {
        float y;
        asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
        return y;
    }
    return __fsqrt_rn(x);
}
// sin and cos of 2 pi y
template <bool FAST = kFastShade>
__device__ __forceinline__ void shade_sincos_2pi(float y, float* s, float* c) {
    if (FAST) sincospif(2.0f * y, s, c);
    else sincosf(__fmul_rn(__fmul_rn(2.0f, 3.14159265358979323846f), y), s, c);
}
template <bool FAST = kFastShade>
__device__ __forceinline__ float shade_rcp(float x) {
    if (FAST) {
        float y;
        asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
        return y;
    }
    return __fdiv_rn(1.0f, x);
}
template <bool FAST = kFastShade>
__device__ __forceinline__ F3 normalize(F3 v) { return v * shade_rsqrt<FAST>(dot(v, v)); }

// ---- Philox4x32-10 (Salmon et al. 2011), counter-based: key = (pixel, seed),
// counter = (sample, slot, block, 0); slot 0 = camera jitter, slot b+1 = scattering event b,
// slot 0x80000000|(b+1) = next-event estimation at event b.  u = ((x >> 8) + 1) * 2^-24 in (0,1].
__device__ __forceinline__ void philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0,
                                              uint32_t k1, uint32_t out[4]) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
        const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
        c0 = hi1 ^ c1 ^ k0; c1 = lo1; c2 = hi0 ^ c3 ^ k1; c3 = lo0;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

struct RngPhilox {
    uint32_t k0, k1, sample, slot, draw;
    uint32_t b0, b1, b2, b3;
    bool need;  // the block that holds draw `draw` has not been generated yet
    __device__ __forceinline__ void start(uint32_t pixel, uint32_t seed, uint32_t sampleIndex, uint32_t /*frameSeed*/) {
        k0 = pixel; k1 = seed; sample = sampleIndex; slot = 0; draw = 0; need = true;
    }
    __device__ __forceinline__ void set_slot(uint32_t s) { slot = s; draw = 0; need = true; }
    // Generates the current block now.  Same numbers as leaving it to the first next(); a caller uses it to run the
    // ten rounds once, with the whole warp, in front of code whose branches would each run their own copy.
    __device__ __forceinline__ void fill() {
        uint32_t o[4];
        philox4x32_10(sample, slot, draw >> 2, 0u, k0, k1, o);
        b0 = o[0]; b1 = o[1]; b2 = o[2]; b3 = o[3];
        need = false;
    }
    // A stream parked after the camera draws and picked up later (counter-based: nothing to carry).
    __device__ __forceinline__ uint32_t save() const { return 0u; }
    __device__ __forceinline__ void resume(uint32_t pixel, uint32_t seed, uint32_t sampleIndex, uint32_t) { start(pixel, seed, sampleIndex, 0u); }
    __device__ __forceinline__ float next() {
        if (need) fill();
        const uint32_t i = draw & 3u;
        const uint32_t x = i == 0 ? b0 : (i == 1 ? b1 : (i == 2 ? b2 : b3));
        ++draw;
        need = (draw & 3u) == 0u;
        return static_cast<float>((x >> 8) + 1u) * (1.0f / 16777216.0f);
    }
};

// ---- the reference stream: rtx.hlsl:61-72 (TEA, 16 rounds) and :74-82 (Park-Miller via
// Schrage in uint arithmetic), seeded per (pixel, frame) and warmed by two draws (:380-382).
struct RngReference {
    uint32_t sd;
    __device__ __forceinline__ void start(uint32_t pixel, uint32_t /*seed*/, uint32_t /*sampleIndex*/, uint32_t frameSeed) {
        uint32_t v0 = pixel, v1 = frameSeed, sum = 0;
#pragma unroll
        for (int n = 0; n < 16; ++n) {
            sum += 0x9e3779b9u;
            v0 += ((v1 << 4) + 0xa341316cu) ^ (v1 + sum) ^ ((v1 >> 5) + 0xc8013ea4u);
            v1 += ((v0 << 4) + 0xad90777du) ^ (v0 + sum) ^ ((v0 >> 5) + 0x7e95761eu);
        }
        sd = v0;
        next();
        next();
    }
    __device__ __forceinline__ void set_slot(uint32_t) {}
    __device__ __forceinline__ void fill() {}
    __device__ __forceinline__ uint32_t save() const { return sd; }
    __device__ __forceinline__ void resume(uint32_t, uint32_t, uint32_t, uint32_t saved) { sd = saved; }
    __device__ __forceinline__ float next() {
        sd = 16807u * (sd % 127773u) - 2836u * (sd / 127773u);
        if (sd > 0x7FFFFFFFu) sd += 0x7FFFFFFFu;
        return __fdiv_rn(__uint2float_rn(sd), 2147483648.0f);
    }
};

#define MCPT_PI 3.14159265358979323846f
#define MCPT_EPS 1.192092896e-07f

// rtx.hlsl:213-251.  TWIN: CUDA/Utils.hpp:97-137 -- the same branches, but only the two refract branches normalise
// (by Utils.hpp:27-34's divide by the length unless it is below FLT_EPSILON); the reflect branches return
// dir - 2 (dir.n) n as it is.
template <class RNG, bool FAST = kFastShade>
__device__ __forceinline__ F3 sample_fresnel(RNG& rng, F3 normal, F3 dir, float Tr, float Ni, bool twin = false) {
    const float x = rng.next();
    F3 outdir;
    bool refracted = false;
    const float ndoti = dot(dir, normal);
    if (FAST) {
        // pow(c, 5) by three products; the quotients by Ni as products with one reciprocal (each within an ulp or two of
        // the IEEE sequence below, which costs ~50 + 5 x 10 instructions)
        const float c1 = 1 - fabsf(ndoti), c2 = c1 * c1;
        Tr = Tr * (1 - c2 * c2 * c1);
        if (x < Tr) {
            const float rNi = shade_rcp<true>(Ni);
            if (ndoti <= 0) {
                const float alpha = -ndoti * rNi - shade_sqrt<true>(1 - (1 - ndoti * ndoti) * rNi * rNi);
                outdir = normal * alpha + dir * rNi;
                refracted = true;
            } else {
                const float test = 1 - (1 - ndoti * ndoti) * Ni * Ni;
                if (test < 0) {
                    outdir = dir - normal * ndoti * 2;
                } else {
                    const float alpha = -ndoti * Ni + shade_sqrt<true>(test);
                    outdir = normal * alpha + dir * Ni;
                    refracted = true;
                }
            }
        } else {
            outdir = dir - normal * ndoti * 2;
        }
        if (twin && !refracted) return outdir;
        return normalize<true>(outdir);
    }
    Tr = Tr * (1 - powf(1 - fabsf(ndoti), 5.0f));
    if (x < Tr) {
        if (ndoti <= 0) {
            const float alpha = -ndoti / Ni - sqrtf(1 - (1 - ndoti * ndoti) / Ni / Ni);
            outdir = normal * alpha + dir / Ni;
            refracted = true;
        } else {
            const float test = 1 - (1 - ndoti * ndoti) * Ni * Ni;
            if (test < 0) {
                outdir = dir - normal * dot(dir, normal) * 2;
            } else {
                const float alpha = -ndoti * Ni + sqrtf(test);
                outdir = normal * alpha + dir * Ni;
                refracted = true;
            }
        }
    } else {
        outdir = dir - normal * dot(dir, normal) * 2;
    }
    if (twin) {
        if (refracted) {
            const float len = sqrtf(dot(outdir, outdir));
            if (fabsf(len) > MCPT_EPS) outdir = outdir / len;
        }
        return outdir;
    }
    return normalize<false>(outdir);
}

// Rotation of a +Y-centred lobe direction onto `normal` (rtx.hlsl:262-274 / :287-300).
// divideByInvlen selects samplePhong's `d.z / invlen` against sampleHemi's `d.z * (1/invlen)`.
template <bool FAST = kFastShade>
__device__ __forceinline__ F3 rotate_to(F3 d, F3 normal, bool divideByInvlen) {
    if (fabsf(normal.y + 1) < MCPT_EPS) return -d;
    if (fabsf(normal.y - 1) >= MCPT_EPS) {
        const float invlen = shade_rsqrt<FAST>(1.0f - normal.y * normal.y);
        F3 o;
        o.x = (normal.z * d.x + normal.x * normal.y * d.z) * invlen + normal.x * d.y;
        if (FAST) o.y = normal.y * d.y - d.z * ((1.0f - normal.y * normal.y) * invlen);  // 1 / invlen
        else o.y = divideByInvlen ? (normal.y * d.y - d.z / invlen) : (normal.y * d.y - d.z * (1.0f / invlen));
        o.z = (-normal.x * d.x + normal.z * normal.y * d.z) * invlen + normal.z * d.y;
        return o;
    }
    return d;
}

// rtx.hlsl:253-276
template <class RNG, bool FAST = kFastShade>
__device__ __forceinline__ F3 sample_phong(RNG& rng, F3 normal, F3 dir, float Ns) {
    const float x = rng.next();
    const float y = rng.next();
    float cosT;
    if (FAST) {
        // x^(1/(Ns+1)) = 2^(log2(x) / (Ns+1)): x in (0,1], the exponent in (0, 1/2): MUFU.LG2's absolute error (2^-22) is
        // scaled down by it, the result is within ~1 ulp of powf's
        cosT = exp2f(__log2f(x) * shade_rcp<true>(Ns + 1));
    } else {
        cosT = powf(x, 1.0f / (Ns + 1));
    }
    // FAST clamps the radicand: MUFU arithmetic can land cosT an ulp above 1 (the IEEE form cannot: powf(x <= 1, e > 0) <= 1)
    const float sinT = FAST ? shade_sqrt<true>(fmaxf(1 - cosT * cosT, 0.0f)) : sqrtf(1 - cosT * cosT);
    float s, c;
    shade_sincos_2pi<FAST>(y, &s, &c);
    const F3 halfdir = rotate_to<FAST>(f3(sinT * c, cosT, sinT * s), normal, true);
    return dir - halfdir * dot(dir, halfdir) * 2;
}

// rtx.hlsl:278-302
template <class RNG, bool FAST = kFastShade>
__device__ __forceinline__ F3 sample_hemi(RNG& rng, F3 normal) {
    const float x = rng.next();
    const float y = rng.next();
    const float sinT = shade_sqrt<FAST>(x);
    const float cosT = shade_sqrt<FAST>(1 - x);
    float s, c;
    shade_sincos_2pi<FAST>(y, &s, &c);
    return rotate_to<FAST>(f3(sinT * c, cosT, sinT * s), normal, false);
}

// Fixed uniforms handed in by the sampler parity hook (mcpt_sample_batch).
struct RngFixed {
    float u0, u1;
    int i;
    __device__ __forceinline__ float next() { return i++ == 0 ? u0 : u1; }
};

}  // namespace mcpt
