/* TEST INFRASTRUCTURE -- the CPU oracle.  NOT product code.
 *
 * A CPU restatement of the reference's KD-tree path tracer
 * (/root/reference/MCRT/QuinEngine: Utils/KDTree.hpp, RTX/ShaderResource.hpp,
 * RTX/GraphicsRTX.cpp, Shader/rtx.hlsl).  Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs may load this library, and only as
 * the checker or the timed CPU baseline -- never as part of the product path.
 *
 * Parity status: the host half (OBJ ingest, tree build, flatten) is PINNED against the
 * reference's own code compiled here (oracle/_ref/libmcpt_ref.so, see ref_bridge.cpp).
 * Device half: rtx.hlsl itself cannot be compiled or run on this box (no HLSL compiler, no
 * graphics stack) and the reference ships no golden vectors; the three BSDF samplers
 * (rtx.hlsl:213-302) are PINNED bit for bit to the reference's own CUDA/Utils.hpp:46-137 --
 * the same math, compiled unmodified on the host into oracle/_ref/libmcpt_refcv.so -- and the
 * determinant order to Matrix33f::det; traversal, path loop and ray generation correspond
 * statement by statement (each function cites its lines) and are frozen by tests/golden/.
 */
#ifndef MCPT_ORACLE_H
#define MCPT_ORACLE_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct oracle_cfg {
    int32_t width, height;
    int32_t spp;            /* number of frames; the shader takes ONE sample per pixel per frame */
    int32_t depth;          /* rtx.hlsl:400 literal 5: RR from `depth`, hard kill at 3*depth    */
    int32_t rng_mode;       /* 0: reference TEA+LCG, frame seeds from mt19937(1234); 1: Philox  */
    uint32_t seed;          /* Philox seed (rng_mode 1)                                         */
    int32_t sample_offset;  /* index of the first frame/sample rendered                          */
    int32_t jitter;         /* 1: +-0.5 px jitter (rtx.hlsl:385-386); 0: bias = 0                */
    float eye[3], at[3], up[3];
    float fovy;             /* radians; GraphicsRTX.cpp:183 uses float(D3DX_PI)/4                */
    int32_t x0, y0, x1, y1; /* pixel window [x0,x1) x [y0,y1) actually rendered (bounded CPU samples) */
    int32_t threads;        /* OpenMP threads, 0 = all                                           */
    int32_t tracer_mode;    /* 0: MCRT rtx.hlsl (KD path, canonical); 1: the CUDA twin's estimator and camera
                               (CUTracer.cu:98-218): +-1 px jitter, fovy = HORIZONTAL fov, origin at the eye,
                               `depth` scattering events + one final gather, no roulette, Kd-tinted glass,
                               Ns truncated to unsigned, emission x emission_scale (ILLUM)                 */
    float emission_scale;   /* 1 for rtx.hlsl; ILLUM = 10 for the twin (stdafx.h:45)                     */
} oracle_cfg;

typedef struct oracle_counters {
    uint64_t paths;         /* primary rays started                       */
    uint64_t rays;          /* intersect() calls                          */
    uint64_t nodes_popped;  /* stack pops (each does one slab test)       */
    uint64_t nodes_entered; /* slab tests that passed                     */
    uint64_t tris_tested;   /* Cramer evaluations                         */
    double seconds;         /* wall time of the call                      */
} oracle_counters;

void oracle_default_cfg(oracle_cfg* cfg, int width, int height, int spp, int depth);

/* tris: 7 x uint32 {v0,v1,v2,n0,n1,n2,matId}; mats: 12 x float {Ka,Kd,Ks,Ns,Tr,Ni}. */
void* oracle_scene_create(const float* verts, int nV, const float* normals, int nN,
                          const uint32_t* tris, int nT, const float* mats, int nM);
void oracle_scene_free(void* scene);

/* Restated KDTree.hpp build + ShaderResource.hpp flatten. Returns the node count. */
int oracle_tree_build(void* scene);
/* Install an externally built tree (e.g. the one libmcpt_ref.so produced). */
void oracle_tree_set(void* scene, int nNodes, const uint32_t* left, const uint32_t* right,
                     const float* aabb6, const uint32_t* splitAxis, const float* splitValue,
                     const uint32_t* leafOffset, const uint32_t* leafCount,
                     const uint32_t* leafIds, int nLeafIds);
/* out = {numNodes, numLeaves, numLeafRefs, maxDepth} */
void oracle_tree_counts(void* scene, int* out);
void oracle_tree_copy(void* scene, uint32_t* left, uint32_t* right, float* aabb6,
                      uint32_t* splitAxis, float* splitValue, uint32_t* leafOffset,
                      uint32_t* leafCount, uint32_t* leafIds);

/* rtx.hlsl:376-398 for one frame; rays = 6 floats {from, dir} per pixel of the window,
 * row-major, top row first. */
void oracle_primary_rays(const oracle_cfg* cfg, int frame, float* rays);

/* rtx.hlsl:144-211. tri = -1 on miss; t/beta/gamma untouched (0) on miss. */
void oracle_intersect(void* scene, const float* rays, int64_t n, int32_t* tri, float* t,
                      float* beta, float* gamma, oracle_counters* counters, int threads);

/* rtx.hlsl:304-405 over cfg.spp frames. sum/sumsq: 3 doubles per window pixel (linear
 * radiance sum and sum of squares over the samples); either may be NULL.
 * display (optional, 4 floats per window pixel) replays the shader's own gamma-space
 * running mean (rtx.hlsl:401-403). */
void oracle_render(void* scene, const oracle_cfg* cfg, double* sum, double* sumsq,
                   float* display, oracle_counters* counters);

/* Small hooks so tests can pin the building blocks separately. */
uint32_t oracle_rand_seed(uint32_t v0, uint32_t v1);
float oracle_rand_gen(uint32_t* sd);
void oracle_frame_seeds(int first, int count, uint32_t* out);
void oracle_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]);
void oracle_sample_hemi(float u1, float u2, const float n[3], float out[3]);
void oracle_sample_phong(float u1, float u2, const float n[3], const float dir[3], float Ns, float out[3]);
void oracle_sample_fresnel(float u, const float n[3], const float dir[3], float Tr, float Ni, float out[3]);
/* The CUDA twin's sampleFresnel and normalize (CVMCTracer/CVMCTracer/CUDA/Utils.hpp:97-137, :27-34). */
void oracle_sample_fresnel_twin(float u, const float n[3], const float dir[3], float Tr, float Ni, float out[3]);
void oracle_normalize_twin(float v[3]);
/* n tuples at once; the uniforms are used exactly as given (raw floats, any grid).
 * kind 0 hemi (u: 2/call), 1 phong (u: 2/call, par0 = Ns), 2 fresnel (u: 1/call, par0 = Tr, par1 = Ni),
 * 3 the twin's fresnel. */
void oracle_sample_batch(int kind, int64_t n, const float* u, const float* normals, const float* dirs,
                         const float* par0, const float* par1, float* out);

#ifdef __cplusplus
}
#endif
#endif
