// TEST INFRASTRUCTURE -- shared internals of the CPU oracle (see oracle.h).
#pragma once
#include <cstdint>
#include <vector>

#include "oracle.h"

namespace oracle {

struct Tri {
    uint32_t v[3], n[3], mat;
};

struct Material {
    float Ka[3], Kd[3], Ks[3], Ns, Tr, Ni;
};

// Flattened tree in the reference's BFS order (ShaderResource.hpp:128-179), with the
// fixed 64-slot id array replaced by (leafOffset, leafCount) into one id list.
struct Tree {
    std::vector<uint32_t> left, right, splitAxis, leafOffset, leafCount, leafIds;
    std::vector<float> aabb;  // 6 per node: min xyz, max xyz
    std::vector<float> splitValue;
    uint32_t maxDepth = 0, numLeaves = 0;
    void clear() {
        left.clear(); right.clear(); splitAxis.clear(); leafOffset.clear();
        leafCount.clear(); leafIds.clear(); aabb.clear(); splitValue.clear();
        maxDepth = 0; numLeaves = 0;
    }
};

struct Scene {
    std::vector<float> verts;    // 3 per vertex
    std::vector<float> normals;  // 3 per normal
    std::vector<Tri> tris;
    std::vector<Material> mats;
    Tree tree;
};

void build_tree(const Scene& scene, Tree& out);

// ---- device half (shader_restate.cpp) -------------------------------------------
struct V3 {
    float x, y, z;
};

struct Hit {
    uint32_t triID;  // 0xFFFFFFFF = miss (rtx.hlsl:147)
    float beta, gamma;
    V3 p;
    float t;  // not part of the shader's HitInfo; kept for the intersection-parity gate
};

struct Tally {
    uint64_t rays = 0, popped = 0, entered = 0, tris = 0;
};

// Random source handed through the path loop exactly where the shader hands `sd`.
// mode 0: rtx.hlsl:61-82 (TEA seed + Park-Miller LCG), sequential.
// mode 1: Philox4x32-10, key = (pixel, seed), counter = (sample, slot, block, 0);
//         slot 0 = camera jitter, slot b+1 = scattering event b; draws are taken
//         sequentially inside a slot; u = ((x >> 8) + 1) * 2^-24 in (0,1].
// mode 2: a fixed list of uniforms handed in by a test (`fixed`), returned as they are.
struct Rng {
    int mode = 0;
    const float* fixed = nullptr;
    uint32_t sd = 0;
    uint32_t key[2] = {0, 0};
    uint32_t sample = 0, slot = 0, draw = 0;
    uint32_t buf[4] = {0, 0, 0, 0};
    void set_slot(uint32_t s) { slot = s; draw = 0; }
    float next();
};

uint32_t rand_seed(uint32_t v0, uint32_t v1);
float rand_gen(uint32_t& sd);
void philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]);

Hit intersect(const Scene& scene, V3 from, V3 dir, Tally& tally);
V3 sample_hemi(Rng& rng, V3 normal);
V3 sample_phong(Rng& rng, V3 normal, V3 dir, float Ns);
V3 sample_fresnel(Rng& rng, V3 normal, V3 dir, float Tr, float Ni);
V3 sample_fresnel_twin(Rng& rng, V3 normal, V3 dir, float Tr, float Ni);
V3 sample_mc(const Scene& scene, Rng& rng, V3 from, V3 dir, uint32_t depth, Tally& tally);
V3 sample_mc_twin(const Scene& scene, Rng& rng, V3 from, V3 dir, uint32_t maxDepth, float illum, Tally& tally);
V3 normalize_twin(V3 v);

}  // namespace oracle
