// Device-resident scene layout (HBM) and the host routine that packs it.
//
// The reference uploads 304-byte CSKDTree nodes with 64 inline triangle ids, 28-byte index
// triangles and separate float3 vertex/normal buffers (Structure.hpp:196-223), so one triangle
// test costs three dependent gathers (rtx.hlsl:164-166).  Here everything a traversal step needs
// is one or two aligned 128-bit loads:
//
//   ref_nodes  2 x float4 per node   {lo.xyz, a} {hi.xyz, b}      reference-order AABB traversal
//                 inner: a = left child index (right = left+1), b = 0
//                 leaf : a = offset into leaf_ids,             b = 0x80000000 | count
//   kd_nodes   1 x uint2 per node (8 B)                            plane traversal (render path)
//                 inner: x = float bits of the split plane, y = (left << 2) | axis(0..2)
//                 leaf : x = offset into leaf_ids,           y = (count << 2) | 3
//   leaf_ids   uint32, ascending inside a leaf (the reference's std::set order)
//   tri_isect  3 x float4 per triangle {a.xyz, c3} {e1.xyz, 0} {e2.xyz, 0}
//                 e1 = a-b, e2 = a-c (the ray-independent columns of rtx.hlsl:167-186),
//                 c3 = e1.y*e2.z - e2.y*e1.z (the one ray-independent cofactor)
//   leaf_tris  3 x float4 per LEAF REFERENCE: the tri_isect record of leaf_ids[k] copied in place k, with the
//                 triangle id in the free lane {e1.xyz, id bits}: a leaf's triangles are consecutive 48-byte
//                 records and the render path needs no id -> record indirection (tuning build variant)
//   rope_nodes, leaf_blob  stackless traversal: see kd_ropes.cpp
//   tri_shade  3 x float4 per triangle {n0.xyz, matId bits} {n1.xyz, 0} {n2.xyz, 0}
//   mats       3 x float4 per material {Ka, Ns} {Kd, Tr} {Ks, Ni}
//   emitters   3 x float4 per emitting triangle {a.xyz, area} {b-a, cdf} {c-a, triangle id bits}
#pragma once
#include <cuda_runtime.h>

#include <cstdint>
#include <string>
#include <vector>

#include "host_scene.h"

namespace mcpt {

struct DScene {
    const float4* ref_nodes;
    const uint2* kd_nodes;
    const uint32_t* leaf_ids;
    const float4* tri_isect;
    const float4* leaf_tris;
    const uint2* rope_nodes;
    const float4* leaf_blob;
    const float4* tri_shade;
    const float4* mats;
    const float4* emitters;
    float bmin[3], bmax[3];
    uint32_t num_nodes, num_tris, num_mats, num_emitters;
    float emitter_area;
};

// rows xaxis, yaxis, zaxis, eye of the inverse view matrix + the two projection scales
// (GraphicsRTX.cpp:173-184; closed forms in DESIGN.md).
struct DCamera {
    float inv[4][3];
    float p11, p22;
    float sx, sy;    // 2 / (width p11), 2 / (height p22): the render kernels' one-product form of rtx.hlsl:391-392
    float tan_half;  // CUDA-twin camera: tan(hfov / 2)
    int twin;        // 0: rtx.hlsl camera, 1: CUTracer.cu:189-212
};

struct PackedScene {
    std::vector<float4> ref_nodes;
    std::vector<uint2> kd_nodes, rope_nodes;
    std::vector<uint32_t> leaf_ids;
    std::vector<uint32_t> leaf_order;  // per node: depth-first ordinal of a leaf (host only)
    std::vector<float4> tri_isect, leaf_tris, leaf_blob, tri_shade, mats, emitters;
    float bmin[3], bmax[3];
    float emitter_area = 0.0f;
    size_t bytes() const {
        return ref_nodes.size() * 16 + (kd_nodes.size() + rope_nodes.size()) * 8 + leaf_ids.size() * 4 +
               (tri_isect.size() + leaf_tris.size() + leaf_blob.size() + tri_shade.size() + mats.size() + emitters.size()) * 16;
    }
};

bool pack_scene(const HostScene& scene, const HostTree& tree, PackedScene& out);  // false: scene too large for the leaf-block layout
void pack_small_tables(const HostScene& scene, const HostTree& tree, PackedScene& out);   // mats, emitters, scene box only (the device pack makes the rest)
void pack_parity_tables(const HostTree& tree, PackedScene& out);  // ref_nodes, leaf_tris: built on first use (parity hooks, A/B kernels)
bool build_ropes(const HostTree& tree, PackedScene& out);  // kd_ropes.cpp; needs kd_nodes, leaf_ids, tri_isect of pack_scene; false: too many leaf references
bool check_ropes(const HostTree& tree, const PackedScene& packed, std::string& err);  // kd_ropes.cpp: structural check of leaf_blob / rope_nodes
int walk_ropes_host(const PackedScene& packed, const float ray[6], int maxHops, uint32_t* leafOff, float* exitT, uint32_t* exitAxis);  // kd_ropes.cpp: diagnostic mirror of the walk
DCamera make_camera(const float eye[3], const float at[3], const float up[3], float fovy, int width, int height, int twin = 0);

}  // namespace mcpt
