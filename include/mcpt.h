/* mcpt.h -- C ABI of libmcpt.so, the B200-native drop-in for the reference's path-sampling
 * hot path (per-pixel Monte Carlo path sampling over a KD-tree-accelerated triangle scene).
 *
 * The entry points are what a binding of the reference's tracer interface would call.
 * The reference exposes the path as four free functions in namespace PW::Tracer
 * (CVMCTracer/CVMCTracer/CUDA/CUTracer.h:9-12):
 *     cudaError_t Initialize();
 *     cudaError_t CreateGeometry(const PW::FileReader::ObjModel* model);
 *     cudaError_t DestroyGeometry();
 *     cudaError_t RenderScene(const PWint sceneID, PWVector3f* hostcolor);
 * and, on the KD-tree side, as Core::Graphics::OnUpdate() ("one more sample per pixel",
 * MCRT/QuinEngine/Core/Graphics.hpp:17-20) over the buffer contract CB0 / CSMaterial /
 * CSTriangle / CSKDTree (MCRT/QuinEngine/Utils/Structure.hpp:186-223).
 * Compile-time knobs of the reference (CVMCTracer/CVMCTracer/stdafx.h:41-45; depth literal
 * rtx.hlsl:400; camera GraphicsRTX.cpp:173-184) become fields of mcpt_render_cfg.
 *
 * Conventions: plain pointers and sizes only; every call is blocking; 0 = success, non-zero =
 * error class with the text in mcpt_last_error(); no exception crosses the boundary; the
 * caller owns every host pointer; the library owns all device memory behind one process-global
 * context (the reference keeps module-level __device__ globals, CUTracer.cu:30-42).
 */
#ifndef MCPT_H
#define MCPT_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum {
    MCPT_OK = 0,
    MCPT_ERR_INVALID = 1,   /* bad argument / call order                */
    MCPT_ERR_IO = 2,        /* file missing or malformed                */
    MCPT_ERR_CUDA = 3,      /* CUDA runtime error (text in last_error)  */
    MCPT_ERR_NO_DEVICE = 4, /* no usable GPU: the product has no CPU path */
    MCPT_ERR_NCCL = 5,      /* libnccl.so.2 missing, or an NCCL call failed (multi-GPU entry points only) */
    MCPT_ERR_INTERNAL = 6   /* a self-check of the library's own tables failed */
};

/* KD-tree construction rule. */
enum {
    MCPT_BUILD_REFERENCE = 0, /* node-for-node the tree of KDTree.hpp:58-287 (parity anchor) */
    MCPT_BUILD_SAH = 1,       /* this library's own O(n log n) SAH builder on the host (same hits, other tree) */
    MCPT_BUILD_GPU = 2        /* level-synchronous SAH build on the bound device (binned above 32 references per node, exact
                                 below) AND the render-path tables made there too (leaf blocks, neighbour links, shading records:
                                 nothing packed on the host, nothing uploaded): the fast ingest for large scenes; needs mcpt_initialize */
};

/* Which OBJ/MTL reader semantics mcpt_load_obj follows. */
enum {
    MCPT_LOADER_TINYOBJ = 0,  /* tinyobjloader 1.1.1 as used by QuinModel (Structure.hpp:9-12)  */
    MCPT_LOADER_OBJREADER = 1 /* PW::FileReader::ObjModel (CVMCTracer/.../ObjReader.cpp:8-259)   */
};

enum {
    MCPT_RNG_PHILOX = 0,    /* Philox4x32-10 keyed by (pixel, sample, bounce) -- the product stream */
    MCPT_RNG_REFERENCE = 1  /* rtx.hlsl:61-82 TEA+LCG with the mt19937(1234) frame seeds (GraphicsRTX.cpp:170-171) */
};

/* Which of the reference's two estimators / cameras a render follows (never mixed). */
enum {
    MCPT_TRACER_MCRT = 0,      /* rtx.hlsl: +-0.5 px jitter, fovy, origin on the image plane, roulette from max_depth,
                                  kill at 3*max_depth, untinted glass -- the KD-tree path, canonical            */
    MCPT_TRACER_CUDA_TWIN = 1  /* CUTracer.cu:98-218: +-1 px jitter, `fovy` is the HORIZONTAL fov (60 deg) with square
                                  pixels, origin at the eye, max_depth (7) scattering events + one final gather, no
                                  roulette, glass tinted by Kd, Ns truncated to unsigned                        */
};

enum {
    MCPT_FLAG_NEE = 1u,        /* next-event estimation to the emitters at Lambert vertices   */
    MCPT_FLAG_MOMENTS = 2u,    /* also accumulate per-pixel sum of squares                    */
    MCPT_FLAG_NO_JITTER = 4u,  /* pixel-corner rays ("bias = 0"), for the intersection gate   */
    MCPT_FLAG_REF_TRAVERSAL = 8u, /* render with the reference-order AABB traversal instead of the KD traversal */
    MCPT_FLAG_SIMPLE_KERNEL = 16u, /* one-ray-per-iteration kernel instead of the warp-voted state machine (A/B) */
    MCPT_FLAG_RENDER_ARITHMETIC = 32u /* mcpt_primary_rays only: the rays as the render kernels compute them (hardware
                                        rsqrt, folded camera factors) instead of the IEEE form that is bit-equal to the oracle */
};

/* Sampler kinds of mcpt_sample_batch. */
enum {
    MCPT_SAMPLE_HEMI = 0,         /* rtx.hlsl:278-302 = CUDA/Utils.hpp:46-70                      */
    MCPT_SAMPLE_PHONG = 1,        /* rtx.hlsl:253-276 = CUDA/Utils.hpp:72-95   (par0 = Ns)        */
    MCPT_SAMPLE_FRESNEL = 2,      /* rtx.hlsl:213-251                          (par0 = Tr, par1 = Ni) */
    MCPT_SAMPLE_FRESNEL_TWIN = 3  /* CUDA/Utils.hpp:97-137: normalised in the refract branches only */
};

/* Intersection kernels selectable through mcpt_intersect_batch. */
enum {
    MCPT_ISECT_REFERENCE = 0, /* rtx.hlsl:144-211 visit order and slab test, strict IEEE arithmetic */
    MCPT_ISECT_KD = 1,        /* front-to-back plane traversal of the compact KD nodes with a stack   */
    MCPT_ISECT_ROPES = 2      /* stackless leaf-to-leaf traversal over neighbour links (render path)  */
};

/* Flat scene tables: exactly the fields of CSTriangle / CSMaterial (Structure.hpp:196-211) and
 * Geometry::Material/Triangle (CVMCTracer/CVMCTracer/Framework/Geometry.h:14-35). */
typedef struct mcpt_scene_desc {
    const float* vertices;     /* 3 floats per vertex                              */
    int32_t num_vertices;
    const float* normals;      /* 3 floats per normal                              */
    int32_t num_normals;
    const uint32_t* triangles; /* 7 uint32 per triangle: v0 v1 v2 n0 n1 n2 matId   */
    int32_t num_triangles;
    const float* materials;    /* 12 floats per material: Ka[3] Kd[3] Ks[3] Ns Tr Ni */
    int32_t num_materials;
    int32_t build_mode;        /* MCPT_BUILD_*                                     */
} mcpt_scene_desc;

typedef struct mcpt_render_cfg {
    int32_t width, height;
    int32_t spp;             /* samples per pixel rendered by this call                         */
    int32_t max_depth;       /* the shader's `depth` (rtx.hlsl:400): RR from max_depth, kill at 3*max_depth */
    int32_t sample_offset;   /* absolute index of the first sample (spp-split across ranks)     */
    uint32_t seed;           /* Philox key word                                                 */
    int32_t rng_mode;        /* MCPT_RNG_*                                                      */
    uint32_t flags;          /* MCPT_FLAG_*                                                     */
    float eye[3], at[3], up[3]; /* D3DXMatrixLookAtRH arguments (GraphicsRTX.cpp:175-178)       */
    float fovy;              /* radians (GraphicsRTX.cpp:183)                                   */
    float emission_scale;    /* 1 for the KD-tree tracer; the CUDA twin multiplies Ka by ILLUM  */
    int32_t tracer_mode;     /* MCPT_TRACER_*                                                   */
} mcpt_render_cfg;

typedef struct mcpt_counters {
    uint64_t paths;          /* camera paths started by the last render call            */
    uint64_t rays;           /* intersect invocations (camera + bounce + shadow rays)   */
    double kernel_ms;        /* device time of the render kernels (CUDA events)         */
    double total_ms;         /* device time of the whole call incl. reduce and copies   */
    uint32_t launches;       /* kernels launched by the last call                       */
    uint32_t reserved;
} mcpt_counters;

typedef struct mcpt_tree_info {
    int32_t num_nodes, num_leaves, num_leaf_refs, max_depth;
    int32_t num_emitter_triangles;
    double build_seconds;
    uint64_t device_bytes;   /* tables resident in HBM: what a render reads, plus the parity-only tables once a hook asked for them */
} mcpt_tree_info;

/* ---- PW::Tracer::Initialize (CUTracer.cu:220-223) ---------------------------------------- */
int mcpt_initialize(int device);
int mcpt_shutdown(void);

/* ---- PW::Tracer::CreateGeometry (CUTracer.cu:225-314) ------------------------------------- */
int mcpt_create_geometry(const mcpt_scene_desc* scene);
/* readObj + CreateGeometry (CVMCTracer/main.cpp:16-17) or QuinModel + KDTree + upload
 * (GraphicsRTX.cpp:165-167) in one call. */
int mcpt_load_obj(const char* obj_path, const char* mtl_base, int loader, int build_mode);
/* Same from a baked scene container (scenes/ *.mcpt, magic "MCPTSCN1"). */
int mcpt_load_scene_file(const char* path, int build_mode);
int mcpt_get_tree_info(mcpt_tree_info* info);
/* Host-side structural check of the stackless-traversal tables built for the current geometry (leaf blocks, leaf
 * handles, neighbour links: csrc/kd_ropes.cpp check_ropes); needs no device.  MCPT_OK, or MCPT_ERR_INTERNAL with the
 * first inconsistency in mcpt_last_error(). */
int mcpt_check_stackless_layout(void);
/* Diagnostic: the render-path tables resident on the device (leaf blocks, location nodes, shading records) against the host's own
 * pack of the same tree, word for word; differing = {leaf-block float4s, nodes, shading float4s}.  After MCPT_BUILD_GPU this
 * compares what the device pack made with what pack_scene + build_ropes make on the host (must be {0, 0, 0}). */
int mcpt_debug_compare_device_tables(uint64_t differing[3]);
/* Diagnostic, host only: the leaves (block offsets in float4 units) the stackless walk visits for one ray {origin, direction},
 * with each leaf's exit distance and axis, and the triangle ids those leaves hold (0xFFFFFFFF between leaves).  Mirrors the
 * WALK of the device code operation for operation; it tests no triangle and is not a render path. */
int mcpt_debug_stackless_walk(const float* ray6, int max_hops, int* hops, uint32_t* leaf_offsets, float* exit_t, uint32_t* exit_axis,
                              uint32_t* tri_ids, int max_ids, int* num_ids);
/* Copies the built tree out in the reference's BFS layout (ShaderResource.hpp:128-179);
 * any pointer may be NULL. aabb6: 6 floats per node. */
int mcpt_get_tree(uint32_t* left, uint32_t* right, float* aabb6, uint32_t* split_axis,
                  float* split_value, uint32_t* leaf_offset, uint32_t* leaf_count, uint32_t* leaf_ids);
/* Copies the flat scene tables out (after any load); pointers may be NULL; counts = {V,N,T,M}. */
int mcpt_get_scene(int32_t counts[4], float* vertices, float* normals, uint32_t* triangles, float* materials);

/* ---- PW::Tracer::DestroyGeometry (CUTracer.cu:316-338) ------------------------------------ */
int mcpt_destroy_geometry(void);

/* ---- PW::Tracer::RenderScene (CUTracer.cu:340-404) / Graphics::OnUpdate x spp -------------- */
void mcpt_default_render_cfg(mcpt_render_cfg* cfg, int width, int height, int spp, int max_depth);
/* RenderScene(sceneID, ...) of the CUDA tracer as a cfg: CUDA_TWIN mode, depth 7, hfov 60 deg, ILLUM = 10,
 * eye (0,5,17) for scene 1 and (0,5,23) otherwise (CUTracer.cu:347-374, stdafx.h:41-45). */
void mcpt_cuda_twin_render_cfg(mcpt_render_cfg* cfg, int width, int height, int spp, int scene_id);
/* host_rgb: width*height*3 floats, row-major, top row first, LINEAR MEAN radiance (same layout as
 * `hostcolor`). host_m2 (optional): mean of squared samples per channel (needs MCPT_FLAG_MOMENTS). */
int mcpt_render_scene(const mcpt_render_cfg* cfg, float* host_rgb, float* host_m2);
/* Device-resident variant: ADDS this call's samples into caller-owned device buffers
 * dev_sum / dev_sumsq (width*height float4: r,g,b sums and the sample count in .w) on `stream`
 * (a cudaStream_t, may be NULL).  Used for spp-split multi-GPU: each rank accumulates its share,
 * then one reduce of dev_sum combines them. */
int mcpt_render_accumulate_device(const mcpt_render_cfg* cfg, void* dev_sum, void* dev_sumsq, void* stream);
/* No counterpart in the reference.  mcpt_render_accumulate_device(cfg,
 * dev_sum, NULL, stream) for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses
 * that forbid it): this call allocates NOTHING on the device (no cudaMalloc, cudaFree or allocating library routine).
 * The slice scratch of the render kernel and its counters live in one fixed `static __device__` array of
 * 256 B + MCPT_NOALLOC_MAX_SLICE_PIXELS * 16 B (16 MiB) that every GPU loading the library holds until exit
 * (csrc/noalloc.cu).  Limits: width*height <= MCPT_NOALLOC_MAX_SLICE_PIXELS (MCPT_ERR_INVALID beyond it);
 * MCPT_RNG_PHILOX only; no MCPT_FLAG_MOMENTS / REF_TRAVERSAL / SIMPLE_KERNEL; the geometry must already be resident
 * (loaded after mcpt_initialize: this call uploads nothing).  Same samples as the allocating entry point; the same
 * bits too whenever width*height*slices fits the scratch with the slice count that one picks, otherwise fewer slices
 * (another fp32 summation order).  All calls share the one array, which keeps the last call's contents (the call
 * clears what must start at zero): calls that overlap on the GPU would race, so the call is blocking and serialised. */
#define MCPT_NOALLOC_MAX_SLICE_PIXELS (1u << 20)
int mcpt_render_accumulate_device_noalloc(const mcpt_render_cfg* cfg, void* dev_sum, void* stream);

/* ---- multi-GPU: spp split + ONE NCCL reduce, behind the boundary ------------------
 * The reference is single-GPU (CUTracer.cu:222 `cudaSetDevice(0)` is its only device call); this is the net-new
 * scaling path.  One process per GPU: each calls mcpt_initialize(its device), loads the same scene, then
 *     rank 0:  mcpt_comm_get_unique_id(id)  -> hand the 128 bytes to the other ranks (file, socket, MPI, ...)
 *     all:     mcpt_comm_init(rank, nranks, id)
 *     all:     mcpt_render_scene_distributed(cfg_of_the_WHOLE_job, rank == 0 ? hostcolor : NULL)
 * Rank r renders the r-th contiguous share of cfg's sample range (mcpt_rank_share; Philox counters carry the absolute
 * sample index, so the union over ranks is exactly the 1-GPU sample set), one ncclReduce(float4 sums, root 0) on the
 * render stream combines the accumulation buffers over NVLink, and rank 0 normalises on the device and copies ONE
 * float3 image to the host.  NCCL is bound at run time (dlopen libnccl.so.2): single-GPU callers need none. */
#define MCPT_UNIQUE_ID_BYTES 128
void mcpt_rank_share(int32_t spp, int32_t sample_offset, int rank, int nranks, int32_t* share_offset, int32_t* share_spp);
int mcpt_comm_get_unique_id(void* id128);
int mcpt_comm_init(int rank, int nranks, const void* id128);   /* after mcpt_initialize; nranks == 1 needs no id / no NCCL */
int mcpt_comm_destroy(void);
int mcpt_comm_info(int* rank, int* nranks, int* nccl_version);
/* host_rgb (rank 0 only; others may pass NULL): width*height*3 floats, the `hostcolor` layout of RenderScene. */
int mcpt_render_scene_distributed(const mcpt_render_cfg* whole_job, float* host_rgb);
/* Device-resident form: ADDS this rank's share into dev_sum (width*height float4, caller-zeroed) on `stream`
 * and reduces dev_sum in place onto rank 0. */
int mcpt_render_accumulate_distributed_device(const mcpt_render_cfg* whole_job, void* dev_sum, void* stream);

/* ---- intersection parity hook (rtx.hlsl:144-211 `intersect`) ------------------------------ */
/* rays: 6 floats {origin, direction} each. tri = -1 on miss (t, beta, gamma = 0 then). */
int mcpt_intersect_batch(const float* rays, int64_t n, int variant, int32_t* tri, float* t,
                         float* beta, float* gamma);
int mcpt_intersect_batch_device(const void* dev_rays, int64_t n, int variant, void* dev_tri,
                                void* dev_t, void* dev_beta, void* dev_gamma, void* stream);
/* Primary rays of one sample index (6 floats/pixel): rtx.hlsl:376-398 in IEEE arithmetic, bit-equal to the oracle's;
 * with MCPT_FLAG_RENDER_ARITHMETIC in cfg->flags the rays exactly as the render kernels generate them (same map, one
 * host-computed factor per axis and the hardware rsqrt: a few ulp from the IEEE form). */
int mcpt_primary_rays(const mcpt_render_cfg* cfg, int sample, float* host_rays);
/* ---- sampler parity hook (rtx.hlsl:213-302 / CUDA/Utils.hpp:46-137) ------------------------ */
/* Runs the device BSDF samplers on n tuples.  u: 2 uniforms per tuple, used as given (fresnel takes the first);
 * normals, dirs: 3 floats per tuple (dirs may be NULL for MCPT_SAMPLE_HEMI); par0/par1 per tuple as the kind needs;
 * out: 3 floats per tuple.  render_arithmetic = 0: the IEEE forms (compared with the oracle to the ulp);
 * 1: the forms the render kernels run (MUFU rsqrt/sqrt/rcp/lg2/ex2, sincospif). */
int mcpt_sample_batch(int kind, int64_t n, const float* u, const float* normals, const float* dirs,
                      const float* par0, const float* par1, int render_arithmetic, float* out);

/* Work counters of one traversal variant over a host ray batch: out = {nodes popped, leaves (KD) or
 * boxes (reference order) entered, triangle tests}.  Feeds the roofline's algorithmic-work figures. */
int mcpt_traversal_stats(const float* rays, int64_t n, int variant, uint64_t out[3]);
int mcpt_get_counters(mcpt_counters* out);
/* Measured FP32 FMA issue rate of the bound device in GFLOP/s (roofline denominator for the
 * intersection / shading arithmetic; the reference publishes nothing comparable). */
int mcpt_measure_fp32_peak(double* gflops);
/* Self-check of the kernels' shared-reciprocal division (intersect.cuh: div3_rn) against the IEEE
 * divide it stands for: n pseudo-random (n0, n1, n2, d) quadruples with exponents spread over
 * 2^-spread .. 2^+spread, every quotient compared bit for bit.  *mismatches must come back 0;
 * *fast_path = how many quadruples took the shared-reciprocal branch (so the check is not vacuous). */
int mcpt_selfcheck_div3(int64_t n, uint32_t seed, int spread, uint64_t* mismatches, uint64_t* fast_path);
/* ---- image output (CUTracer.cu:383-396 step PNGs, GraphicsRTX.cpp:223-226 temp.png) ---------- */
/* rgb: width*height*3 linear floats, top row first.  PNG stores clamp(pow(v, 1/gamma)) in 8 bits
 * (gamma 2.2 = the display value of rtx.hlsl:402; 1.0 = the CUDA tracer's linear dump). */
int mcpt_write_png(const char* path, const float* rgb, int width, int height, float gamma);
int mcpt_write_pfm(const char* path, const float* rgb, int width, int height);
/* ---- tuning / A-B knobs ---------------------------------------------------------------------- */
/* The library never reads the caller's environment.  Knobs (clamped to their valid range): "shade_threshold",
 * "cam_threshold" [1,32], "tri_cap" [1,64], "chunks" [0,64; 0 = automatic], "sah_ci", "sah_eb" [0.05,100] (next
 * tree build), "leaf_bfs", "sah_resort", "build_timing", "sm_stats", "verify_layout", "device_pack" {0,1}, "min_blocks" {0 = automatic, 4, 5, 6, 8}, "pool" {0, 1: the experimental pooled-lane kernel, csrc/render_pool.cuh}, "variant" (MCPT_TUNE builds);
 * "defaults" restores the shipped configuration.  None of them changes a result beyond fp32 summation order. */
int mcpt_set_tuning(const char* name, double value);
int mcpt_get_tuning(const char* name, double* value);
const char* mcpt_last_error(void);
const char* mcpt_version(void);

#ifdef __cplusplus
}
#endif
#endif /* MCPT_H */
