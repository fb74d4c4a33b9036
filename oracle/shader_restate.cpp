// TEST INFRASTRUCTURE -- CPU oracle, device half.  NOT product code.
//
// Line-by-line restatement of the reference's compute shader
//   /root/reference/MCRT/QuinEngine/Shader/rtx.hlsl
// as plain C++ (binary32, statement order kept, compile with -ffp-contract=off and
// without -ffast-math).  Conventions fixed where HLSL leaves them open:
//   * `determinant()` expands in the order of Matrix33f::det
//     (/root/reference/CVMCTracer/CVMCTracer/Framework/Math.hpp:169-175), the only
//     explicit order the reference gives;
//   * 1/x and a/b are IEEE divides, sqrt is IEEE, pow/sin/cos are libm's;
//   * HLSL min/max ignore NaN -> fminf/fmaxf;
//   * normalize(v) = v * (1 / sqrt(dot(v,v))), dot sums left to right.
// Parity status: rtx.hlsl itself cannot be compiled here (no HLSL compiler, no graphics stack) and the
// reference ships no golden outputs.  What IS pinned to the reference's own code: the three samplers are bit-equal
// to CVMCTracer/CVMCTracer/CUDA/Utils.hpp:46-137 -- the same math as rtx.hlsl:213-302, compiled unmodified on the
// host into oracle/_ref/libmcpt_refcv.so -- on the uniforms that code itself drew (tests/test_oracle_pins.py);
// the determinant order to Matrix33f::det; the tree to KDTree.hpp.  The traversal/path loop (:144-211, :304-371)
// and ray generation (:373-398) correspond by citation, statement by statement.
#include <cmath>

#include "oracle_internal.h"

namespace oracle {
namespace {

const float CS_PI = 3.14159265358979323846f;   // rtx.hlsl:1
const float CS_FLT_EPSILON = 1.192092896e-07f;  // rtx.hlsl:2

inline V3 operator+(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline V3 operator-(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline V3 operator*(V3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
inline V3 operator*(V3 a, V3 b) { return {a.x * b.x, a.y * b.y, a.z * b.z}; }
inline V3 operator/(V3 a, float s) { return {a.x / s, a.y / s, a.z / s}; }
inline V3 operator-(V3 a) { return {-a.x, -a.y, -a.z}; }
inline float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline V3 normalize(V3 v) { return v * (1.0f / sqrtf(dot(v, v))); }
inline V3 load3(const std::vector<float>& a, uint32_t i) {
    return {a[3 * static_cast<size_t>(i)], a[3 * static_cast<size_t>(i) + 1], a[3 * static_cast<size_t>(i) + 2]};
}

// Math.hpp:169-175, rows (m0 m1 m2 / m3 m4 m5 / m6 m7 m8).
inline float det3(float m0, float m1, float m2, float m3, float m4, float m5, float m6, float m7, float m8) {
    float res1 = m0 * (m4 * m8 - m5 * m7);
    float res2 = -m1 * (m3 * m8 - m5 * m6);
    float res3 = m2 * (m3 * m7 - m4 * m6);
    return res1 + res2 + res3;
}

// rtx.hlsl:84-142
bool intersect_aabb(const Tree& tree, V3 from, V3 dir, uint32_t nodeId, float curt) {
    const float* bb = &tree.aabb[6 * static_cast<size_t>(nodeId)];
    const float minx = bb[0], miny = bb[1], minz = bb[2], maxx = bb[3], maxy = bb[4], maxz = bb[5];
    float t_min, t_max, t_xmin, t_xmax, t_ymin, t_ymax, t_zmin, t_zmax;
    float x_a = 1.0f / dir.x, y_a = 1.0f / dir.y, z_a = 1.0f / dir.z;
    float x_e = from.x, y_e = from.y, z_e = from.z;
    if (dir.x == 0 && (from.x < minx || from.x > maxx)) return false;
    if (dir.y == 0 && (from.y < miny || from.y > maxy)) return false;
    if (dir.z == 0 && (from.z < minz || from.z > maxz)) return false;
    if (x_a >= 0) { t_xmin = (minx - x_e) * x_a; t_xmax = (maxx - x_e) * x_a; }
    else          { t_xmin = (maxx - x_e) * x_a; t_xmax = (minx - x_e) * x_a; }
    if (y_a >= 0) { t_ymin = (miny - y_e) * y_a; t_ymax = (maxy - y_e) * y_a; }
    else          { t_ymin = (maxy - y_e) * y_a; t_ymax = (miny - y_e) * y_a; }
    if (z_a >= 0) { t_zmin = (minz - z_e) * z_a; t_zmax = (maxz - z_e) * z_a; }
    else          { t_zmin = (maxz - z_e) * z_a; t_zmax = (minz - z_e) * z_a; }
    t_min = fmaxf(t_xmin, fmaxf(t_ymin, t_zmin));
    t_max = fminf(t_xmax, fminf(t_ymax, t_zmax));
    return (t_min <= t_max && t_min <= curt);
}

}  // namespace

// rtx.hlsl:61-72
uint32_t rand_seed(uint32_t v0, uint32_t v1) {
    uint32_t sum = 0;
    for (uint32_t n = 0; n < 16; n++) {
        sum += 0x9e3779b9u;
        v0 += ((v1 << 4) + 0xa341316cu) ^ (v1 + sum) ^ ((v1 >> 5) + 0xc8013ea4u);
        v1 += ((v0 << 4) + 0xad90777du) ^ (v0 + sum) ^ ((v0 >> 5) + 0x7e95761eu);
    }
    return v0;
}

// rtx.hlsl:74-82
float rand_gen(uint32_t& sd) {
    sd = 16807u * (sd % 127773u) - 2836u * (sd / 127773u);
    if (sd > 0x7FFFFFFFu) sd += 0x7FFFFFFFu;
    return 1.0f * sd / 0x7FFFFFFF;
}

// Philox4x32-10 (Salmon et al., SC'11), the stream north_star asks the new build to use.
void philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) {
    uint32_t c0 = ctr[0], c1 = ctr[1], c2 = ctr[2], c3 = ctr[3], k0 = key[0], k1 = key[1];
    for (int r = 0; r < 10; ++r) {
        uint64_t p0 = static_cast<uint64_t>(0xD2511F53u) * c0;
        uint64_t p1 = static_cast<uint64_t>(0xCD9E8D57u) * c2;
        uint32_t n0 = static_cast<uint32_t>(p1 >> 32) ^ c1 ^ k0;
        uint32_t n1 = static_cast<uint32_t>(p1);
        uint32_t n2 = static_cast<uint32_t>(p0 >> 32) ^ c3 ^ k1;
        uint32_t n3 = static_cast<uint32_t>(p0);
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

float Rng::next() {
    if (mode == 0) return rand_gen(sd);
    if (mode == 2) return fixed[draw++];
    if ((draw & 3u) == 0u) {
        uint32_t ctr[4] = {sample, slot, draw >> 2, 0u};
        philox4x32_10(ctr, key, buf);
    }
    uint32_t x = buf[draw & 3u];
    ++draw;
    return static_cast<float>((x >> 8) + 1u) * (1.0f / 16777216.0f);
}

// rtx.hlsl:144-211
Hit intersect(const Scene& scene, V3 from, V3 dir, Tally& tally) {
    const Tree& tree = scene.tree;
    Hit hitInfo;
    hitInfo.triID = 0xFFFFFFFFu;
    hitInfo.beta = hitInfo.gamma = hitInfo.t = 0.0f;
    hitInfo.p = {0.0f, 0.0f, 0.0f};
    ++tally.rays;

    float tmin = 10000;
    uint32_t localstack[64];
    uint32_t stacktop = 0;
    localstack[stacktop++] = 0;
    while (stacktop != 0) {
        uint32_t node = localstack[--stacktop];
        ++tally.popped;
        if (!intersect_aabb(tree, from, dir, node, tmin)) continue;
        ++tally.entered;
        if (tree.splitAxis[node] == 0) {
            const uint32_t off = tree.leafOffset[node], cnt = tree.leafCount[node];
            for (uint32_t i = 0; i < cnt; i++) {
                const uint32_t id = tree.leafIds[off + i];
                const Tri& tri = scene.tris[id];
                V3 a = load3(scene.verts, tri.v[0]);
                V3 b = load3(scene.verts, tri.v[1]);
                V3 c = load3(scene.verts, tri.v[2]);
                ++tally.tris;
                // :167-186, the 4x4 matrices have a unit last row/column -> 3x3 minors
                float detA = det3(a.x - b.x, a.x - c.x, dir.x,
                                  a.y - b.y, a.y - c.y, dir.y,
                                  a.z - b.z, a.z - c.z, dir.z);
                float beta = det3(a.x - from.x, a.x - c.x, dir.x,
                                  a.y - from.y, a.y - c.y, dir.y,
                                  a.z - from.z, a.z - c.z, dir.z) / detA;
                float gamma = det3(a.x - b.x, a.x - from.x, dir.x,
                                   a.y - b.y, a.y - from.y, dir.y,
                                   a.z - b.z, a.z - from.z, dir.z) / detA;
                float t = det3(a.x - b.x, a.x - c.x, a.x - from.x,
                               a.y - b.y, a.y - c.y, a.y - from.y,
                               a.z - b.z, a.z - c.z, a.z - from.z) / detA;
                if (beta + gamma < 1 && beta > 0 && gamma > 0 && t > 0 && t < tmin) {
                    tmin = t;
                    hitInfo.triID = id;
                    hitInfo.beta = beta;
                    hitInfo.gamma = gamma;
                    hitInfo.p.x = from.x + t * dir.x;
                    hitInfo.p.y = from.y + t * dir.y;
                    hitInfo.p.z = from.z + t * dir.z;
                    hitInfo.t = t;
                }
            }
        } else {
            localstack[stacktop++] = tree.left[node];
            localstack[stacktop++] = tree.right[node];
        }
    }
    return hitInfo;
}

// rtx.hlsl:213-251
V3 sample_fresnel(Rng& rng, V3 normal, V3 dir, float Tr, float Ni) {
    float x = rng.next();
    V3 outdir;
    float ndoti = dot(dir, normal);
    Tr = Tr * (1 - powf(1 - fabsf(ndoti), 5));
    if (x < Tr) {
        if (ndoti <= 0) {
            float alpha = -ndoti / Ni - sqrtf(1 - (1 - ndoti * ndoti) / Ni / Ni);
            outdir = normal * alpha + dir / Ni;
        } else {
            float test = 1 - (1 - ndoti * ndoti) * Ni * Ni;
            if (test < 0) {
                outdir = dir - normal * dot(dir, normal) * 2;
            } else {
                float alpha = -ndoti * Ni + sqrtf(test);
                outdir = normal * alpha + dir * Ni;
            }
        }
    } else {
        outdir = dir - normal * dot(dir, normal) * 2;
    }
    return normalize(outdir);
}

// rtx.hlsl:253-276
V3 sample_phong(Rng& rng, V3 normal, V3 dir, float Ns) {
    float x = rng.next();
    float y = rng.next();
    float cosT = powf(x, 1.0f / (Ns + 1));
    float sinT = sqrtf(1 - cosT * cosT);
    float phi = 2 * CS_PI * y;
    V3 halfdir = {sinT * cosf(phi), cosT, sinT * sinf(phi)};
    if (fabsf(normal.y + 1) < CS_FLT_EPSILON) {
        halfdir = -halfdir;
    } else if (fabsf(normal.y - 1) >= CS_FLT_EPSILON) {
        V3 d = halfdir;  // the shader shadows `dir` here (:269)
        float invlen = 1.0f / sqrtf(1.0f - normal.y * normal.y);
        halfdir.x = (normal.z * d.x + normal.x * normal.y * d.z) * invlen + normal.x * d.y;
        halfdir.y = normal.y * d.y - d.z / invlen;
        halfdir.z = (-normal.x * d.x + normal.z * normal.y * d.z) * invlen + normal.z * d.y;
    }
    return dir - halfdir * dot(dir, halfdir) * 2;
}

// rtx.hlsl:278-302
V3 sample_hemi(Rng& rng, V3 normal) {
    float x = rng.next();
    float y = rng.next();
    float sinT = sqrtf(x);
    float cosT = sqrtf(1 - x);
    float phi = 2 * CS_PI * y;
    V3 outdir = {sinT * cosf(phi), cosT, sinT * sinf(phi)};
    if (fabsf(normal.y + 1) < CS_FLT_EPSILON) {
        outdir = -outdir;
    } else if (fabsf(normal.y - 1) >= CS_FLT_EPSILON) {
        V3 d = outdir;
        float invlen = 1.0f / sqrtf(1.0f - normal.y * normal.y);
        float len = 1.0f / invlen;
        outdir.x = (normal.z * d.x + normal.x * normal.y * d.z) * invlen + normal.x * d.y;
        outdir.y = normal.y * d.y - d.z * len;
        outdir.z = (-normal.x * d.x + normal.z * normal.y * d.z) * invlen + normal.z * d.y;
    }
    return outdir;
}

// rtx.hlsl:304-371
V3 sample_mc(const Scene& scene, Rng& rng, V3 from, V3 dir, uint32_t depth, Tally& tally) {
    Hit hit;
    V3 color = {1, 1, 1};
    uint32_t bounce = 0;
    hit = intersect(scene, from, dir, tally);
    while (hit.triID != 0xFFFFFFFFu) {
        rng.set_slot(bounce + 1);  // no-op for the reference stream
        if (bounce >= depth * 3) {
            return {0, 0, 0};
        } else if (bounce >= depth) {
            float illum = fmaxf(fmaxf(color.x, color.y), color.z);
            if (illum > rng.next()) {
                color = color / illum;
            } else {
                return {0, 0, 0};
            }
        }
        const Tri& tri = scene.tris[hit.triID];
        const Material& m = scene.mats[tri.mat];
        if (m.Ka[0] > 0 || m.Ka[1] > 0 || m.Ka[2] > 0) {
            color = color * V3{m.Ka[0], m.Ka[1], m.Ka[2]};
            return color;
        }
        V3 normal1 = load3(scene.normals, tri.n[0]);
        V3 normal2 = load3(scene.normals, tri.n[1]);
        V3 normal3 = load3(scene.normals, tri.n[2]);
        V3 normal = normal1 * (1.0f - hit.beta - hit.gamma) + normal2 * hit.beta + normal3 * hit.gamma;
        normal = normalize(normal);
        if (m.Tr > 0) {
            dir = sample_fresnel(rng, normal, dir, m.Tr, m.Ni);
        } else if (m.Ns > 1) {
            dir = sample_phong(rng, normal, dir, m.Ns);
            color = color * V3{m.Ks[0], m.Ks[1], m.Ks[2]};
        } else {
            if (dot(dir, normal) > 0) {
                dir = -sample_hemi(rng, normal);
            } else {
                dir = sample_hemi(rng, normal);
            }
            color = color * V3{m.Kd[0], m.Kd[1], m.Kd[2]};
        }
        from = hit.p + dir * 0.01f;
        hit = intersect(scene, from, dir, tally);
        ++bounce;
    }
    return {0, 0, 0};
}

// ---- the CUDA twin (CVMCTracer/CVMCTracer/CUDA) ---------------------------------------------------------
// CUDA/Utils.hpp:27-34: divide by the length unless it is (almost) zero.
V3 normalize_twin(V3 v) {
    float len = sqrtf(v.x * v.x + v.y * v.y + v.z * v.z);
    if (fabsf(len) > CS_FLT_EPSILON) {
        v.x /= len; v.y /= len; v.z /= len;
    }
    return v;
}

// CUDA/Utils.hpp:97-137: the twin's sampleFresnel.  Same branches as rtx.hlsl:213-251, but the result is normalised
// only in the two refract branches, by Utils.hpp:27-34's divide-by-length (the reflect branches :120,:134 return
// indir - 2 (indir.n) n as it is), where the shader normalises every branch with rsqrt (:250).
// Pinned bit for bit against the reference's own Utils.hpp compiled on the host (tests/test_oracle_pins.py).
V3 sample_fresnel_twin(Rng& rng, V3 normal, V3 dir, float Tr, float Ni) {
    float x = rng.next();
    V3 outdir;
    float ndoti = dot(dir, normal);
    Tr = Tr * (1 - powf(1 - fabsf(ndoti), 5));
    if (x < Tr) {
        if (ndoti <= 0) {
            float alpha = -ndoti / Ni - sqrtf(1 - (1 - ndoti * ndoti) / Ni / Ni);
            outdir = normalize_twin(normal * alpha + dir / Ni);
        } else {
            float test = 1 - (1 - ndoti * ndoti) * Ni * Ni;
            if (test < 0) {
                outdir = dir - normal * dot(dir, normal) * 2;
            } else {
                float alpha = -ndoti * Ni + sqrtf(test);
                outdir = normalize_twin(normal * alpha + dir * Ni);
            }
        }
    } else {
        outdir = dir - normal * dot(dir, normal) * 2;
    }
    return outdir;
}

// CUTracer.cu:98-177 over the same scene tables (the twin binds one material per OBJ group; the loader
// already gives every triangle its group's material, CUTracer.cu:300-308).  Same samplers as rtx.hlsl
// (Utils.hpp:46-137 is the same math) except that samplePhong takes Ns as an unsigned (Utils.hpp:72).
V3 sample_mc_twin(const Scene& scene, Rng& rng, V3 from, V3 dir, uint32_t maxDepth, float illum, Tally& tally) {
    V3 color = {1, 1, 1};
    for (uint32_t depth = 0; depth < maxDepth; ++depth) {
        Hit hit = intersect(scene, from, dir, tally);
        if (hit.triID == 0xFFFFFFFFu) return {0, 0, 0};
        rng.set_slot(depth + 1);
        const Tri& tri = scene.tris[hit.triID];
        const Material& m = scene.mats[tri.mat];
        if (m.Ka[0] > 0 || m.Ka[1] > 0 || m.Ka[2] > 0) {
            color.x *= m.Ka[0] * illum; color.y *= m.Ka[1] * illum; color.z *= m.Ka[2] * illum;
            return color;
        }
        V3 normal = load3(scene.normals, tri.n[0]) * (1.0f - hit.beta - hit.gamma) + load3(scene.normals, tri.n[1]) * hit.beta +
                    load3(scene.normals, tri.n[2]) * hit.gamma;
        normal = normalize_twin(normal);
        if (m.Tr > 0) {
            dir = sample_fresnel_twin(rng, normal, dir, m.Tr, m.Ni);
            color = color * V3{m.Kd[0], m.Kd[1], m.Kd[2]};  // :131-133, the tint rtx.hlsl:345 comments out
        } else if (m.Ns > 1) {
            dir = sample_phong(rng, normal, dir, static_cast<float>(static_cast<uint32_t>(m.Ns)));
            color = color * V3{m.Ks[0], m.Ks[1], m.Ks[2]};
        } else {
            color = color * V3{m.Kd[0], m.Kd[1], m.Kd[2]};
            dir = dot(dir, normal) > 0 ? -sample_hemi(rng, normal) : sample_hemi(rng, normal);
        }
        from = hit.p + dir * 0.01f;
    }
    // :162-175 the final gather: whatever is hit contributes throughput * Ka * ILLUM (zero for non-emitters)
    Hit hit = intersect(scene, from, dir, tally);
    if (hit.triID == 0xFFFFFFFFu) return {0, 0, 0};
    const Material& m = scene.mats[scene.tris[hit.triID].mat];
    color.x *= m.Ka[0] * illum; color.y *= m.Ka[1] * illum; color.z *= m.Ka[2] * illum;
    return color;
}

}  // namespace oracle
