// TEST INFRASTRUCTURE -- CPU oracle, C entry points and the render driver.  NOT product code.
//
// Driver part restated from /root/reference/MCRT/QuinEngine/RTX/GraphicsRTX.cpp:163-209
// (camera, frame-seed stream, prevCount) and rtx.hlsl:373-405 (ray generation and the
// gamma-space running mean).  See oracle.h for the parity status.
#include <omp.h>

#include <chrono>
#include <cmath>
#include <cstring>
#include <random>

#include "oracle_internal.h"

using namespace oracle;

namespace {

struct Camera {
    float inv[4][4];  // the matrix the shader calls viewMatrix: rows xaxis, yaxis, zaxis, eye
    float p11, p22;
};

// D3DXMatrixLookAtRH + D3DXMatrixInverse (GraphicsRTX.cpp:173-181) in closed form:
// zaxis = normalize(eye - at), xaxis = normalize(up x zaxis), yaxis = zaxis x xaxis; the inverse
// of the view matrix has rows (xaxis,0) (yaxis,0) (zaxis,0) (eye,1).  The two transposes at
// :180/:184 cancel against HLSL's column-major cbuffer packing.
// D3DXMatrixPerspectiveFovRH (:183): _22 = cot(fovy/2), _11 = _22 / aspect.
Camera make_camera(const oracle_cfg& cfg) {
    Camera cam;
    auto norm = [](float v[3]) {
        float l = sqrtf(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
        v[0] /= l; v[1] /= l; v[2] /= l;
    };
    float z[3] = {cfg.eye[0] - cfg.at[0], cfg.eye[1] - cfg.at[1], cfg.eye[2] - cfg.at[2]};
    norm(z);
    float x[3] = {cfg.up[1] * z[2] - cfg.up[2] * z[1], cfg.up[2] * z[0] - cfg.up[0] * z[2],
                  cfg.up[0] * z[1] - cfg.up[1] * z[0]};
    norm(x);
    float y[3] = {z[1] * x[2] - z[2] * x[1], z[2] * x[0] - z[0] * x[2], z[0] * x[1] - z[1] * x[0]};
    for (int i = 0; i < 3; ++i) {
        cam.inv[0][i] = x[i]; cam.inv[1][i] = y[i]; cam.inv[2][i] = z[i]; cam.inv[3][i] = cfg.eye[i];
    }
    cam.inv[0][3] = cam.inv[1][3] = cam.inv[2][3] = 0.0f;
    cam.inv[3][3] = 1.0f;
    float aspect = 1.0f * cfg.width / cfg.height;
    cam.p22 = static_cast<float>(1.0 / std::tan(static_cast<double>(cfg.fovy) * 0.5));
    cam.p11 = cam.p22 / aspect;
    return cam;
}

// rtx.hlsl:384-398.  u1,u2 are the two jitter draws (0.5 each when jitter is off).
void camera_ray(const oracle_cfg& cfg, const Camera& cam, uint32_t px, uint32_t py, float u1, float u2,
                V3& from, V3& dir) {
    if (cfg.tracer_mode == 1) {
        // CUTracer.cu:195-212: +-1 px jitter, horizontal fov, square pixels, the 2.0 / 1.0 literals make the
        // arithmetic double; origin = camPos.  cam.inv rows are right, up, -forward, eye.
        float biasx = px + (u1 * 2.0f - 1.0f);
        float biasy = py + (u2 * 2.0f - 1.0f);
        const float tanH = tanf(cfg.fovy * 0.5f);
        float d[3] = {static_cast<float>((2.0 * biasx / cfg.width - 1) * tanH),
                      static_cast<float>((1.0 * cfg.height / cfg.width - 2.0 * biasy / cfg.width) * tanH), -1.0f};
        V3 w;
        w.x = cam.inv[0][0] * d[0] + cam.inv[1][0] * d[1] + cam.inv[2][0] * d[2];
        w.y = cam.inv[0][1] * d[0] + cam.inv[1][1] * d[1] + cam.inv[2][1] * d[2];
        w.z = cam.inv[0][2] * d[0] + cam.inv[1][2] * d[1] + cam.inv[2][2] * d[2];
        dir = normalize_twin(w);
        from = {cam.inv[3][0], cam.inv[3][1], cam.inv[3][2]};
        return;
    }
    float biasx = px + (u1 - 0.5f);
    float biasy = py + (u2 - 0.5f);
    float d[3] = {(2.0f * biasx / cfg.width - 1) / cam.p11, (1 - 2.0f * biasy / cfg.height) / cam.p22, -1.0f};
    float f[3], w[3];
    for (int i = 0; i < 3; ++i) {
        // mul(float4(d, 1), M) and mul(float4(d, 0), M): row vector times matrix
        w[i] = d[0] * cam.inv[0][i] + d[1] * cam.inv[1][i] + d[2] * cam.inv[2][i];
        f[i] = w[i] + cam.inv[3][i];
    }
    from = {f[0], f[1], f[2]};
    float inv = 1.0f / sqrtf(w[0] * w[0] + w[1] * w[1] + w[2] * w[2]);
    dir = {w[0] * inv, w[1] * inv, w[2] * inv};
}

// GraphicsRTX.cpp:170-171,192: mt19937(1234) through uniform_int_distribution<unsigned>
// over its full range, i.e. the raw 32-bit outputs, one per frame.
void frame_seeds(int first, int count, uint32_t* out) {
    std::mt19937 rng(1234);
    rng.discard(static_cast<unsigned long long>(first));
    for (int i = 0; i < count; ++i) out[i] = static_cast<uint32_t>(rng());
}

void window(const oracle_cfg& cfg, int& x0, int& y0, int& x1, int& y1) {
    x0 = cfg.x0; y0 = cfg.y0; x1 = cfg.x1; y1 = cfg.y1;
    if (x1 <= x0 || y1 <= y0) { x0 = 0; y0 = 0; x1 = cfg.width; y1 = cfg.height; }
}

// Sets up the per-(pixel, frame) random source and draws the two jitter numbers the way
// rtx.hlsl:380-386 does.
void start_sample(const oracle_cfg& cfg, uint32_t pixel, uint32_t frame, uint32_t frameSeed, Rng& rng,
                  float& u1, float& u2) {
    rng.mode = cfg.rng_mode;
    if (cfg.rng_mode == 0) {
        rng.sd = rand_seed(pixel, frameSeed);
        rng.next();
        rng.next();
    } else {
        rng.key[0] = pixel;
        rng.key[1] = cfg.seed;
        rng.sample = frame;
        rng.set_slot(0);
    }
    if (cfg.jitter) {
        u1 = rng.next();
        u2 = rng.next();
    } else {
        u1 = u2 = 0.5f;
    }
}

}  // namespace

extern "C" {

void oracle_default_cfg(oracle_cfg* cfg, int width, int height, int spp, int depth) {
    std::memset(cfg, 0, sizeof(*cfg));
    cfg->width = width; cfg->height = height; cfg->spp = spp; cfg->depth = depth;
    cfg->rng_mode = 0; cfg->seed = 0; cfg->sample_offset = 0; cfg->jitter = 1;
    cfg->eye[0] = 0; cfg->eye[1] = 5; cfg->eye[2] = 17;  // GraphicsRTX.cpp:175-177
    cfg->at[0] = 0; cfg->at[1] = 5; cfg->at[2] = 16;
    cfg->up[0] = 0; cfg->up[1] = 1; cfg->up[2] = 0;
    cfg->fovy = 3.141592654f / 4.0f;  // static_cast<FLOAT>(D3DX_PI) / 4.0f
    cfg->x0 = cfg->y0 = cfg->x1 = cfg->y1 = 0;
    cfg->threads = 0;
    cfg->tracer_mode = 0;
    cfg->emission_scale = 1.0f;
}

void* oracle_scene_create(const float* verts, int nV, const float* normals, int nN, const uint32_t* tris,
                          int nT, const float* mats, int nM) {
    auto* s = new Scene;
    s->verts.assign(verts, verts + 3 * static_cast<size_t>(nV));
    if (normals && nN > 0) s->normals.assign(normals, normals + 3 * static_cast<size_t>(nN));
    s->tris.resize(static_cast<size_t>(nT));
    for (int i = 0; i < nT; ++i) {
        Tri& t = s->tris[static_cast<size_t>(i)];
        for (int k = 0; k < 3; ++k) { t.v[k] = tris[7 * i + k]; t.n[k] = tris[7 * i + 3 + k]; }
        t.mat = tris[7 * i + 6];
    }
    s->mats.resize(static_cast<size_t>(nM));
    for (int i = 0; i < nM; ++i) {
        Material& m = s->mats[static_cast<size_t>(i)];
        for (int k = 0; k < 3; ++k) { m.Ka[k] = mats[12 * i + k]; m.Kd[k] = mats[12 * i + 3 + k]; m.Ks[k] = mats[12 * i + 6 + k]; }
        m.Ns = mats[12 * i + 9]; m.Tr = mats[12 * i + 10]; m.Ni = mats[12 * i + 11];
    }
    return s;
}

void oracle_scene_free(void* scene) { delete static_cast<Scene*>(scene); }

int oracle_tree_build(void* scene) {
    auto* s = static_cast<Scene*>(scene);
    build_tree(*s, s->tree);
    return static_cast<int>(s->tree.left.size());
}

void oracle_tree_set(void* scene, int nNodes, const uint32_t* left, const uint32_t* right, const float* aabb6,
                     const uint32_t* splitAxis, const float* splitValue, const uint32_t* leafOffset,
                     const uint32_t* leafCount, const uint32_t* leafIds, int nLeafIds) {
    Tree& t = static_cast<Scene*>(scene)->tree;
    t.clear();
    size_t n = static_cast<size_t>(nNodes);
    t.left.assign(left, left + n); t.right.assign(right, right + n);
    t.aabb.assign(aabb6, aabb6 + 6 * n);
    t.splitAxis.assign(splitAxis, splitAxis + n); t.splitValue.assign(splitValue, splitValue + n);
    t.leafOffset.assign(leafOffset, leafOffset + n); t.leafCount.assign(leafCount, leafCount + n);
    t.leafIds.assign(leafIds, leafIds + nLeafIds);
    for (size_t k = 0; k < n; ++k) if (t.splitAxis[k] == 0) ++t.numLeaves;
    // depth by one pass: children always come after their parent in BFS order
    std::vector<uint32_t> depth(n, 0);
    for (size_t k = 0; k < n; ++k) {
        if (t.splitAxis[k] != 0) { depth[t.left[k]] = depth[k] + 1; depth[t.right[k]] = depth[k] + 1; }
        if (depth[k] > t.maxDepth) t.maxDepth = depth[k];
    }
}

void oracle_tree_counts(void* scene, int* out) {
    const Tree& t = static_cast<Scene*>(scene)->tree;
    out[0] = static_cast<int>(t.left.size());
    out[1] = static_cast<int>(t.numLeaves);
    out[2] = static_cast<int>(t.leafIds.size());
    out[3] = static_cast<int>(t.maxDepth);
}

void oracle_tree_copy(void* scene, uint32_t* left, uint32_t* right, float* aabb6, uint32_t* splitAxis,
                      float* splitValue, uint32_t* leafOffset, uint32_t* leafCount, uint32_t* leafIds) {
    const Tree& t = static_cast<Scene*>(scene)->tree;
    size_t n = t.left.size();
    std::memcpy(left, t.left.data(), n * 4); std::memcpy(right, t.right.data(), n * 4);
    std::memcpy(aabb6, t.aabb.data(), n * 24);
    std::memcpy(splitAxis, t.splitAxis.data(), n * 4); std::memcpy(splitValue, t.splitValue.data(), n * 4);
    std::memcpy(leafOffset, t.leafOffset.data(), n * 4); std::memcpy(leafCount, t.leafCount.data(), n * 4);
    std::memcpy(leafIds, t.leafIds.data(), t.leafIds.size() * 4);
}

void oracle_primary_rays(const oracle_cfg* cfg, int frame, float* rays) {
    Camera cam = make_camera(*cfg);
    int x0, y0, x1, y1;
    window(*cfg, x0, y0, x1, y1);
    uint32_t seed = 0;
    if (cfg->rng_mode == 0) frame_seeds(frame, 1, &seed);
    size_t k = 0;
    for (int y = y0; y < y1; ++y) {
        for (int x = x0; x < x1; ++x, ++k) {
            Rng rng;
            float u1, u2;
            start_sample(*cfg, static_cast<uint32_t>(y) * cfg->width + x, static_cast<uint32_t>(frame), seed, rng, u1, u2);
            V3 from, dir;
            camera_ray(*cfg, cam, x, y, u1, u2, from, dir);
            float* r = rays + 6 * k;
            r[0] = from.x; r[1] = from.y; r[2] = from.z; r[3] = dir.x; r[4] = dir.y; r[5] = dir.z;
        }
    }
}

void oracle_intersect(void* scene, const float* rays, int64_t n, int32_t* tri, float* t, float* beta,
                      float* gamma, oracle_counters* counters, int threads) {
    const Scene& s = *static_cast<Scene*>(scene);
    if (threads <= 0) threads = omp_get_max_threads();
    uint64_t cr = 0, cp = 0, ce = 0, ct = 0;
    auto t0 = std::chrono::steady_clock::now();
#pragma omp parallel for schedule(dynamic, 1024) num_threads(threads) reduction(+ : cr, cp, ce, ct)
    for (int64_t i = 0; i < n; ++i) {
        Tally tally;
        const float* r = rays + 6 * i;
        Hit h = intersect(s, {r[0], r[1], r[2]}, {r[3], r[4], r[5]}, tally);
        tri[i] = static_cast<int32_t>(h.triID);
        if (t) t[i] = h.t;
        if (beta) beta[i] = h.beta;
        if (gamma) gamma[i] = h.gamma;
        cr += tally.rays; cp += tally.popped; ce += tally.entered; ct += tally.tris;
    }
    auto t1 = std::chrono::steady_clock::now();
    if (counters) {
        counters->paths = 0; counters->rays = cr; counters->nodes_popped = cp;
        counters->nodes_entered = ce; counters->tris_tested = ct;
        counters->seconds = std::chrono::duration<double>(t1 - t0).count();
    }
}

void oracle_render(void* scene, const oracle_cfg* cfgp, double* sum, double* sumsq, float* display,
                   oracle_counters* counters) {
    const Scene& s = *static_cast<Scene*>(scene);
    const oracle_cfg cfg = *cfgp;
    Camera cam = make_camera(cfg);
    int x0, y0, x1, y1;
    window(cfg, x0, y0, x1, y1);
    const int ww = x1 - x0;
    int threads = cfg.threads > 0 ? cfg.threads : omp_get_max_threads();
    std::vector<uint32_t> seeds(static_cast<size_t>(cfg.spp > 0 ? cfg.spp : 1), 0u);
    if (cfg.rng_mode == 0) frame_seeds(cfg.sample_offset, cfg.spp, seeds.data());
    uint64_t cr = 0, cp = 0, ce = 0, ct = 0;
    auto t0 = std::chrono::steady_clock::now();
#pragma omp parallel for schedule(dynamic, 1) num_threads(threads) reduction(+ : cr, cp, ce, ct)
    for (int y = y0; y < y1; ++y) {
        Tally tally;
        for (int x = x0; x < x1; ++x) {
            const size_t k = static_cast<size_t>(y - y0) * ww + (x - x0);
            double acc[3] = {0, 0, 0}, acc2[3] = {0, 0, 0};
            float shown[4] = {0, 0, 0, 0};
            for (int f = 0; f < cfg.spp; ++f) {
                const uint32_t frame = static_cast<uint32_t>(cfg.sample_offset + f);
                Rng rng;
                float u1, u2;
                start_sample(cfg, static_cast<uint32_t>(y) * cfg.width + x, frame, seeds[f], rng, u1, u2);
                V3 from, dir;
                camera_ray(cfg, cam, x, y, u1, u2, from, dir);
                V3 c = cfg.tracer_mode == 1 ? sample_mc_twin(s, rng, from, dir, static_cast<uint32_t>(cfg.depth), cfg.emission_scale, tally)
                                            : sample_mc(s, rng, from, dir, static_cast<uint32_t>(cfg.depth), tally);
                if (cfg.tracer_mode == 0 && cfg.emission_scale != 1.0f) { c.x *= cfg.emission_scale; c.y *= cfg.emission_scale; c.z *= cfg.emission_scale; }
                acc[0] += c.x; acc[1] += c.y; acc[2] += c.z;
                acc2[0] += static_cast<double>(c.x) * c.x; acc2[1] += static_cast<double>(c.y) * c.y;
                acc2[2] += static_cast<double>(c.z) * c.z;
                if (display) {
                    // rtx.hlsl:401-402, prevCount = frame index (GraphicsRTX.cpp:193)
                    const float col[4] = {c.x, c.y, c.z, 1.0f};
                    const float prev = static_cast<float>(frame);
                    for (int ch = 0; ch < 4; ++ch)
                        shown[ch] = powf((powf(shown[ch], 2.2f) * prev + col[ch]) / (prev + 1), 1 / 2.2f);
                }
            }
            if (sum) { sum[3 * k] = acc[0]; sum[3 * k + 1] = acc[1]; sum[3 * k + 2] = acc[2]; }
            if (sumsq) { sumsq[3 * k] = acc2[0]; sumsq[3 * k + 1] = acc2[1]; sumsq[3 * k + 2] = acc2[2]; }
            if (display) std::memcpy(display + 4 * k, shown, sizeof(shown));
        }
        cr += tally.rays; cp += tally.popped; ce += tally.entered; ct += tally.tris;
    }
    auto t1 = std::chrono::steady_clock::now();
    if (counters) {
        counters->paths = static_cast<uint64_t>(ww) * (y1 - y0) * cfg.spp;
        counters->rays = cr; counters->nodes_popped = cp; counters->nodes_entered = ce;
        counters->tris_tested = ct;
        counters->seconds = std::chrono::duration<double>(t1 - t0).count();
    }
}

uint32_t oracle_rand_seed(uint32_t v0, uint32_t v1) { return rand_seed(v0, v1); }
float oracle_rand_gen(uint32_t* sd) { return rand_gen(*sd); }
void oracle_frame_seeds(int first, int count, uint32_t* out) { frame_seeds(first, count, out); }
void oracle_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) { philox4x32_10(ctr, key, out); }

namespace {
// Feeds fixed uniforms into the samplers exactly as given (no quantisation: the reference's own streams --
// curand_uniform, Park-Miller / 0x7FFFFFFF -- do not live on Philox's 2^-24 grid).
struct FixedRng {
    float u[2];
    Rng r;
    FixedRng(float a, float b) {
        u[0] = a; u[1] = b;
        r.mode = 2; r.fixed = u; r.draw = 0;
    }
};
}  // namespace

void oracle_sample_hemi(float u1, float u2, const float n[3], float out[3]) {
    FixedRng f(u1, u2);
    V3 o = sample_hemi(f.r, {n[0], n[1], n[2]});
    out[0] = o.x; out[1] = o.y; out[2] = o.z;
}
void oracle_sample_phong(float u1, float u2, const float n[3], const float dir[3], float Ns, float out[3]) {
    FixedRng f(u1, u2);
    V3 o = sample_phong(f.r, {n[0], n[1], n[2]}, {dir[0], dir[1], dir[2]}, Ns);
    out[0] = o.x; out[1] = o.y; out[2] = o.z;
}
void oracle_sample_fresnel(float u, const float n[3], const float dir[3], float Tr, float Ni, float out[3]) {
    FixedRng f(u, u);
    V3 o = sample_fresnel(f.r, {n[0], n[1], n[2]}, {dir[0], dir[1], dir[2]}, Tr, Ni);
    out[0] = o.x; out[1] = o.y; out[2] = o.z;
}
void oracle_sample_fresnel_twin(float u, const float n[3], const float dir[3], float Tr, float Ni, float out[3]) {
    FixedRng f(u, u);
    V3 o = sample_fresnel_twin(f.r, {n[0], n[1], n[2]}, {dir[0], dir[1], dir[2]}, Tr, Ni);
    out[0] = o.x; out[1] = o.y; out[2] = o.z;
}
void oracle_normalize_twin(float v[3]) {
    V3 o = normalize_twin({v[0], v[1], v[2]});
    v[0] = o.x; v[1] = o.y; v[2] = o.z;
}

// Batched forms (n tuples) for the parity tests: u is 2 floats per call (hemi, phong) or 1 (fresnel).
// kind: 0 hemi, 1 phong (par0 = Ns), 2 fresnel (par0 = Tr, par1 = Ni), 3 the twin's fresnel (Utils.hpp:97-137).
void oracle_sample_batch(int kind, int64_t n, const float* u, const float* normals, const float* dirs,
                         const float* par0, const float* par1, float* out) {
    for (int64_t i = 0; i < n; ++i) {
        const float* nn = normals + 3 * i;
        const float* dd = dirs ? dirs + 3 * i : nullptr;
        float* o = out + 3 * i;
        switch (kind) {
            case 0: oracle_sample_hemi(u[2 * i], u[2 * i + 1], nn, o); break;
            case 1: oracle_sample_phong(u[2 * i], u[2 * i + 1], nn, dd, par0[i], o); break;
            case 2: oracle_sample_fresnel(u[i], nn, dd, par0[i], par1[i], o); break;
            default: oracle_sample_fresnel_twin(u[i], nn, dd, par0[i], par1[i], o); break;
        }
    }
}

}  // extern "C"
