// KD-tree construction on the GPU (MCPT_BUILD_GPU) -- SURVEY.md 8(f) rank 1, first option.
//
// The reference builds its tree on one host thread (KDTree.hpp:58-287: 25 s at 1 M triangles); this library's host
// SAH builder (kd_build_sah.cpp) needs 0.8 s there on 16 cores -- still twice the render it feeds (K4: 0.47 s).  Here the
// same surface-area heuristic runs level-synchronously on the device: all nodes of one depth are split by a handful of
// kernels, and because a level's children are allocated in the order of their parents the result IS the breadth-first,
// children-adjacent layout the rest of the pipeline expects (ShaderResource.hpp:128-179).  Same cost model and the same
// straddle rule as the host builders (flat on the plane -> left only, min < plane -> left, max > plane -> right):
//   * a reference is (triangle id, node); its box is the triangle's box cut to the node's cell on the fly (nested cells:
//     cutting to the final cell equals cutting at every level), so no per-reference boxes are stored;
//   * nodes of more than kSmall references take the best of the kBins - 1 equidistant planes per axis, counted with one
//     min-bin / max-bin histogram pass (Shevtsov et al. 2007); smaller nodes sweep every reference bound exactly;
//   * cost = 1 + Ci (P_L n_L + P_R n_R) against the leaf cost Ci n (Ci from tuning().sah_ci), depth cap as the host builder.
// The tree differs from the host SAH tree (no perfect splits: straddlers are cut as boxes, not as polygons), the hits do
// not: every traversal returns the same closest hit on any valid tree (tests/test_gpu_builder.py).
#include <cuda_runtime.h>

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <string>
#include <vector>

#include <cub/device/device_scan.cuh>

#include "context.h"
#include "host_scene.h"

namespace mcpt {
namespace {

constexpr int kBins = 32;
constexpr uint32_t kSmall = 32;

struct LevelNode {      // one node of the level being split
    float lo[3], hi[3]; // cell
    uint32_t begin, end;// its references in the level's reference array
};

struct Decision {       // what split_kernel decided for a level node
    int axis;           // -1 = leaf
    float plane;
};

#define GK_CHECK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { err = std::string(#x) + ": " + cudaGetErrorString(e_); return false; } } while (0)

__device__ __forceinline__ float surface(const float lo[3], const float hi[3]) {
    const float x = hi[0] - lo[0], y = hi[1] - lo[1], z = hi[2] - lo[2];
    return 2.0f * (x * y + y * z + z * x);
}

__global__ void tri_bounds_kernel(const float* __restrict__ verts, const uint32_t* __restrict__ tris, uint32_t nT,
                                  float4* __restrict__ triLo, float4* __restrict__ triHi, uint32_t* __restrict__ refTri,
                                  uint32_t* __restrict__ refNode) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= nT) return;
    float lo[3] = {3.4e38f, 3.4e38f, 3.4e38f}, hi[3] = {-3.4e38f, -3.4e38f, -3.4e38f};
    for (int k = 0; k < 3; ++k) {
        const float* v = verts + 3 * static_cast<size_t>(tris[7 * static_cast<size_t>(t) + k]);
        for (int i = 0; i < 3; ++i) { lo[i] = fminf(lo[i], v[i]); hi[i] = fmaxf(hi[i], v[i]); }
    }
    triLo[t] = make_float4(lo[0], lo[1], lo[2], 0.f);
    triHi[t] = make_float4(hi[0], hi[1], hi[2], 0.f);
    refTri[t] = t;
    refNode[t] = 0u;
}

// scene box = union of the triangle boxes (one block, enough for a one-off reduction over <= a few M triangles)
// Two stages: kBoxBlocks blocks reduce a slice each into partLo / partHi (root == nullptr), one block reduces those into the root node.
constexpr unsigned kBoxBlocks = 256;
__global__ void scene_box_kernel(const float4* __restrict__ triLo, const float4* __restrict__ triHi, uint32_t nT, LevelNode* root,
                                 float4* __restrict__ partLo, float4* __restrict__ partHi, uint32_t rootRefs) {
    __shared__ float slo[3][256], shi[3][256];
    float lo[3] = {3.4e38f, 3.4e38f, 3.4e38f}, hi[3] = {-3.4e38f, -3.4e38f, -3.4e38f};
    for (uint32_t t = blockIdx.x * 256u + threadIdx.x; t < nT; t += 256u * gridDim.x) {
        const float4 a = triLo[t], b = triHi[t];
        lo[0] = fminf(lo[0], a.x); lo[1] = fminf(lo[1], a.y); lo[2] = fminf(lo[2], a.z);
        hi[0] = fmaxf(hi[0], b.x); hi[1] = fmaxf(hi[1], b.y); hi[2] = fmaxf(hi[2], b.z);
    }
    for (int i = 0; i < 3; ++i) { slo[i][threadIdx.x] = lo[i]; shi[i][threadIdx.x] = hi[i]; }
    __syncthreads();
    for (int s = 128; s > 0; s >>= 1) {
        if (threadIdx.x < s)
            for (int i = 0; i < 3; ++i) {
                slo[i][threadIdx.x] = fminf(slo[i][threadIdx.x], slo[i][threadIdx.x + s]);
                shi[i][threadIdx.x] = fmaxf(shi[i][threadIdx.x], shi[i][threadIdx.x + s]);
            }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        if (root) {
            for (int i = 0; i < 3; ++i) { root->lo[i] = slo[i][0]; root->hi[i] = shi[i][0]; }
            root->begin = 0; root->end = rootRefs;
        } else {
            partLo[blockIdx.x] = make_float4(slo[0][0], slo[1][0], slo[2][0], 0.f);
            partHi[blockIdx.x] = make_float4(shi[0][0], shi[1][0], shi[2][0], 0.f);
        }
    }
}

// Histogram pass for the big nodes of the level: bins[node][axis][0][b] counts references whose (cut) lower bound falls in
// bin b, [1][b] those whose upper bound does.
__global__ void bin_kernel(const LevelNode* __restrict__ nodes, const uint32_t* __restrict__ refTri, const uint32_t* __restrict__ refNode,
                           uint32_t nRefs, const float4* __restrict__ triLo, const float4* __restrict__ triHi,
                           const uint32_t* __restrict__ binSlot, uint32_t* __restrict__ bins) {
    // A node's references are contiguous: most blocks work on one node.  The block counts that node (the node of its first
    // reference) in shared memory and adds the 192 totals to the global histogram once; other nodes' references go there directly.
    __shared__ uint32_t sh[3 * 2 * kBins];
    __shared__ uint32_t leadNode;
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    for (uint32_t k = threadIdx.x; k < 3 * 2 * kBins; k += blockDim.x) sh[k] = 0u;
    if (threadIdx.x == 0) leadNode = refNode[blockIdx.x * blockDim.x];
    __syncthreads();
    const uint32_t lead = leadNode;
    if (i < nRefs) {
        const uint32_t nd = refNode[i];
        const uint32_t slot = binSlot[nd];
        if (slot != 0xFFFFFFFFu) {   // else small node: exact sweep in split_kernel
            const LevelNode n = nodes[nd];
            const float4 a = triLo[refTri[i]], b = triHi[refTri[i]];
            const float tlo[3] = {a.x, a.y, a.z}, thi[3] = {b.x, b.y, b.z};
            uint32_t* base = nd == lead ? sh : bins + static_cast<size_t>(slot) * (3 * 2 * kBins);
#pragma unroll
            for (int ax = 0; ax < 3; ++ax) {
                const float w = n.hi[ax] - n.lo[ax];
                if (!(w > 0.0f)) continue;
                const float s = static_cast<float>(kBins) / w;
                const float lo = fmaxf(tlo[ax], n.lo[ax]), hi = fminf(thi[ax], n.hi[ax]);
                const int bl = min(kBins - 1, max(0, static_cast<int>((lo - n.lo[ax]) * s)));
                const int bh = min(kBins - 1, max(0, static_cast<int>((hi - n.lo[ax]) * s)));
                atomicAdd(base + (ax * 2 + 0) * kBins + bl, 1u);
                atomicAdd(base + (ax * 2 + 1) * kBins + bh, 1u);
            }
        }
    }
    __syncthreads();
    const uint32_t leadSlot = binSlot[lead];
    if (leadSlot != 0xFFFFFFFFu)
        for (uint32_t k = threadIdx.x; k < 3 * 2 * kBins; k += blockDim.x)
            if (sh[k]) atomicAdd(bins + static_cast<size_t>(leadSlot) * (3 * 2 * kBins) + k, sh[k]);
}

// One thread per level node: SAH over the binned planes (big nodes) or over every reference bound (small nodes).
__global__ void split_kernel(const LevelNode* __restrict__ nodes, uint32_t nNodes, const uint32_t* __restrict__ refTri,
                             const float4* __restrict__ triLo, const float4* __restrict__ triHi, const uint32_t* __restrict__ binSlot,
                             const uint32_t* __restrict__ bins, float ci, int canSplit, Decision* __restrict__ out) {
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= nNodes) return;
    const LevelNode n = nodes[k];
    const uint32_t cnt = n.end - n.begin;
    Decision d;
    d.axis = -1; d.plane = 0.0f;
    if (cnt == 0u || !canSplit) { out[k] = d; return; }
    const float area = surface(n.lo, n.hi);
    float best = ci * static_cast<float>(cnt);   // leaf cost: split only if strictly cheaper
    if (!(area > 0.0f)) { out[k] = d; return; }
    const float invArea = 1.0f / area;
    auto consider = [&](int ax, float s, uint32_t nl, uint32_t nr) {
        s += 0.0f;   // -0 -> +0: two references may offer the same plane under both signs, the tie-break below cannot tell them apart
        if (s < n.lo[ax] || s > n.hi[ax]) return;
        if ((s <= n.lo[ax] && nr == cnt) || (s >= n.hi[ax] && nl == cnt)) return;   // a child identical to its parent: no progress
        float llo[3] = {n.lo[0], n.lo[1], n.lo[2]}, lhi[3] = {n.hi[0], n.hi[1], n.hi[2]};
        float rlo[3] = {n.lo[0], n.lo[1], n.lo[2]}, rhi[3] = {n.hi[0], n.hi[1], n.hi[2]};
        lhi[ax] = s; rlo[ax] = s;
        const float c = 1.0f + ci * (surface(llo, lhi) * invArea * static_cast<float>(nl) + surface(rlo, rhi) * invArea * static_cast<float>(nr));
        // strict improvement, or the same cost at a smaller (axis, plane): the references of a node arrive in the order the scatter's
        // atomics left them, the choice must not depend on it
        if (c < best || (c == best && d.axis >= 0 && (ax < d.axis || (ax == d.axis && s < d.plane)))) { best = c; d.axis = ax; d.plane = s; }
    };
    const uint32_t slot = binSlot[k];
    if (slot != 0xFFFFFFFFu) {
        const uint32_t* base = bins + static_cast<size_t>(slot) * (3 * 2 * kBins);
        for (int ax = 0; ax < 3; ++ax) {
            const float w = n.hi[ax] - n.lo[ax];
            if (!(w > 0.0f)) continue;
            uint32_t nl = 0, nr = cnt;
            for (int b = 1; b < kBins; ++b) {     // plane between bin b-1 and bin b
                nl += base[(ax * 2 + 0) * kBins + (b - 1)];   // lower bound left of the plane -> goes left
                nr -= base[(ax * 2 + 1) * kBins + (b - 1)];   // upper bound left of the plane -> does not go right
                consider(ax, n.lo[ax] + w * (static_cast<float>(b) / static_cast<float>(kBins)), nl, nr);
            }
        }
    } else {
        // small node (<= kSmall references): every reference bound is a candidate plane, counted exactly against all references.
        // The bounds cut to the cell are fetched once per axis into a local array; the n^2 comparisons then run on it.
        float blo[kSmall], bhi[kSmall];
        for (int ax = 0; ax < 3; ++ax) {
            if (!(n.hi[ax] > n.lo[ax])) continue;
            for (uint32_t i = 0; i < cnt; ++i) {
                const uint32_t t = refTri[n.begin + i];
                const float4 a = triLo[t], b = triHi[t];
                blo[i] = fmaxf(ax == 0 ? a.x : (ax == 1 ? a.y : a.z), n.lo[ax]);
                bhi[i] = fminf(ax == 0 ? b.x : (ax == 1 ? b.y : b.z), n.hi[ax]);
            }
            for (uint32_t i = 0; i < cnt; ++i) {
                for (int side = 0; side < 2; ++side) {
                    const float s = side ? bhi[i] : blo[i];
                    uint32_t nl = 0, nr = 0;
                    for (uint32_t j = 0; j < cnt; ++j) {
                        const float l2 = blo[j], h2 = bhi[j];
                        if (l2 == h2 && l2 == s) { ++nl; continue; }   // flat on the plane: left only
                        nl += l2 < s ? 1u : 0u;
                        nr += h2 > s ? 1u : 0u;
                    }
                    consider(ax, s, nl, nr);
                }
            }
        }
    }
    out[k] = d;
}

// The references of a node are contiguous in the level's array, so the 32 lanes of a warp almost always work on ONE node: their
// atomics on that node's counters are folded into one per warp (the top levels would otherwise send a million atomics to a
// single address).  Returns true when every lane of the (full) warp is in range and names the same node.
__device__ __forceinline__ bool warp_same_node(bool valid, uint32_t nd) {
    const uint32_t lead = __shfl_sync(0xFFFFFFFFu, nd, 0);
    return __all_sync(0xFFFFFFFFu, valid && nd == lead);
}

// Big nodes choose among equidistant planes; a plane through the middle of a row of triangles cuts (duplicates) all of them,
// while the nearest reference bound next to it usually cuts none (exactly none on grid-aligned meshes such as K4's height
// field).  snap_kernel finds, per big node, the reference bound nearest to the chosen plane (64-bit atomicMin on
// (distance bits, bound bits)); snap_apply_kernel moves the plane there if it lies within one bin width.
__global__ void snap_kernel(const LevelNode* __restrict__ nodes, const Decision* __restrict__ dec, const uint32_t* __restrict__ refTri,
                            const uint32_t* __restrict__ refNode, uint32_t nRefs, const float4* __restrict__ triLo, const float4* __restrict__ triHi,
                            const uint32_t* __restrict__ binSlot, unsigned long long* __restrict__ snapKey) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    const bool valid = i < nRefs;
    const uint32_t nd = valid ? refNode[i] : 0xFFFFFFFFu;
    const bool uniform = warp_same_node(valid, nd);
    unsigned long long key = ~0ull;
    uint32_t slot = 0xFFFFFFFFu;
    if (valid) {
        slot = binSlot[nd];
        const Decision d = dec[nd];
        if (slot != 0xFFFFFFFFu && d.axis >= 0) {
            const LevelNode n = nodes[nd];
            const float4 a = triLo[refTri[i]], b = triHi[refTri[i]];
            const int ax = d.axis;
            const float lo = fmaxf(ax == 0 ? a.x : (ax == 1 ? a.y : a.z), n.lo[ax]);
            const float hi = fminf(ax == 0 ? b.x : (ax == 1 ? b.y : b.z), n.hi[ax]);
            for (int side = 0; side < 2; ++side) {
                const float c = (side ? hi : lo) + 0.0f;   // -0 -> +0
                if (!(c > n.lo[ax] && c < n.hi[ax])) continue;   // strictly inside the cell: both children shrink
                const unsigned long long k2 = (static_cast<unsigned long long>(__float_as_uint(fabsf(c - d.plane))) << 32) | __float_as_uint(c);
                key = k2 < key ? k2 : key;
            }
        }
    }
    if (uniform) {
        for (int o = 16; o > 0; o >>= 1) {
            const unsigned long long other = __shfl_xor_sync(0xFFFFFFFFu, key, o);
            key = other < key ? other : key;
        }
        if ((threadIdx.x & 31u) == 0u && key != ~0ull && slot != 0xFFFFFFFFu) atomicMin(snapKey + slot, key);
    } else if (key != ~0ull) {
        atomicMin(snapKey + slot, key);
    }
}
__global__ void snap_apply_kernel(const LevelNode* __restrict__ nodes, uint32_t nNodes, const uint32_t* __restrict__ binSlot,
                                  const unsigned long long* __restrict__ snapKey, Decision* __restrict__ dec) {
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= nNodes) return;
    const uint32_t slot = binSlot[k];
    if (slot == 0xFFFFFFFFu || dec[k].axis < 0) return;
    const unsigned long long key = snapKey[slot];
    if (key == ~0ull) return;
    const int ax = dec[k].axis;
    const float w = (nodes[k].hi[ax] - nodes[k].lo[ax]) * (1.0f / kBins);
    if (__uint_as_float(static_cast<uint32_t>(key >> 32)) <= w) dec[k].plane = __uint_as_float(static_cast<uint32_t>(key));
}

// classification of reference i against its node's plane: bit 0 left, bit 1 right (the host builders' straddle rule)
__device__ __forceinline__ uint32_t classify(const LevelNode& n, const Decision& d, float4 a, float4 b) {
    const int ax = d.axis;
    const float lo = fmaxf(ax == 0 ? a.x : (ax == 1 ? a.y : a.z), n.lo[ax]);
    const float hi = fminf(ax == 0 ? b.x : (ax == 1 ? b.y : b.z), n.hi[ax]);
    if (lo == hi && lo == d.plane) return 1u;
    return (lo < d.plane ? 1u : 0u) | (hi > d.plane ? 2u : 0u);
}

__global__ void count_kernel(const LevelNode* __restrict__ nodes, const Decision* __restrict__ dec, const uint32_t* __restrict__ refTri,
                             const uint32_t* __restrict__ refNode, uint32_t nRefs, const float4* __restrict__ triLo,
                             const float4* __restrict__ triHi, uint32_t* __restrict__ nL, uint32_t* __restrict__ nR) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    const bool valid = i < nRefs;
    const uint32_t nd = valid ? refNode[i] : 0xFFFFFFFFu;
    const bool uniform = warp_same_node(valid, nd);
    uint32_t c = 0u;
    if (valid) {
        const Decision d = dec[nd];
        if (d.axis >= 0) c = classify(nodes[nd], d, triLo[refTri[i]], triHi[refTri[i]]);
    }
    if (uniform) {
        const uint32_t l = __popc(__ballot_sync(0xFFFFFFFFu, (c & 1u) != 0u)), r = __popc(__ballot_sync(0xFFFFFFFFu, (c & 2u) != 0u));
        if ((threadIdx.x & 31u) == 0u) {
            if (l) atomicAdd(nL + nd, l);
            if (r) atomicAdd(nR + nd, r);
        }
    } else {
        if (c & 1u) atomicAdd(nL + nd, 1u);
        if (c & 2u) atomicAdd(nR + nd, 1u);
    }
}

// per level node: sizes that the three scans turn into offsets
__global__ void sizes_kernel(const LevelNode* __restrict__ nodes, const Decision* __restrict__ dec, uint32_t nNodes, const uint32_t* __restrict__ nL,
                             const uint32_t* __restrict__ nR, uint32_t* __restrict__ isSplit, uint32_t* __restrict__ childRefs, uint32_t* __restrict__ leafRefs) {
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= nNodes) return;
    const bool split = dec[k].axis >= 0;
    isSplit[k] = split ? 1u : 0u;
    childRefs[k] = split ? nL[k] + nR[k] : 0u;
    leafRefs[k] = split ? 0u : nodes[k].end - nodes[k].begin;
}

// Writes the level's nodes into the global (breadth-first) tree and creates the next level's nodes.
__global__ void emit_kernel(const LevelNode* __restrict__ nodes, const Decision* __restrict__ dec, uint32_t nNodes, uint32_t levelBase,
                            uint32_t nextBase, uint32_t leafIdBase, const uint32_t* __restrict__ nL, const uint32_t* __restrict__ nR,
                            const uint32_t* __restrict__ splitRank, const uint32_t* __restrict__ childRefOff, const uint32_t* __restrict__ leafRefOff,
                            LevelNode* __restrict__ next, uint32_t* __restrict__ gLeft, float* __restrict__ gBox, uint32_t* __restrict__ gAxis,
                            float* __restrict__ gPlane, uint32_t* __restrict__ gLeafOff, uint32_t* __restrict__ gLeafCnt) {
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= nNodes) return;
    const LevelNode n = nodes[k];
    const Decision d = dec[k];
    const uint32_t g = levelBase + k;
    for (int i = 0; i < 3; ++i) { gBox[6 * static_cast<size_t>(g) + i] = n.lo[i]; gBox[6 * static_cast<size_t>(g) + 3 + i] = n.hi[i]; }
    gLeafOff[g] = leafIdBase + leafRefOff[k];   // running count of leaf ids in front of this node (inner nodes too, as the host flatten does)
    if (d.axis < 0) {
        gLeft[g] = 0xFFFFFFFFu; gAxis[g] = 0u; gPlane[g] = 0.0f; gLeafCnt[g] = n.end - n.begin;
        return;
    }
    const uint32_t r = splitRank[k];
    gLeft[g] = nextBase + 2u * r; gAxis[g] = static_cast<uint32_t>(d.axis) + 1u; gPlane[g] = d.plane; gLeafCnt[g] = 0u;
    LevelNode l = n, rr = n;
    l.hi[d.axis] = d.plane; rr.lo[d.axis] = d.plane;
    l.begin = childRefOff[k]; l.end = l.begin + nL[k];
    rr.begin = l.end; rr.end = rr.begin + nR[k];
    next[2u * r] = l; next[2u * r + 1u] = rr;
}

// Moves every reference to its place: into the child ranges of the next level (both, for a straddler) or into the leaf id list.
__global__ void scatter_kernel(const LevelNode* __restrict__ nodes, const Decision* __restrict__ dec, const uint32_t* __restrict__ refTri,
                               const uint32_t* __restrict__ refNode, uint32_t nRefs, const float4* __restrict__ triLo, const float4* __restrict__ triHi,
                               const uint32_t* __restrict__ splitRank, const LevelNode* __restrict__ next, uint32_t* __restrict__ cursor /* 2 per split node */,
                               uint32_t* __restrict__ nextTri, uint32_t* __restrict__ nextNode, uint32_t leafIdBase, const uint32_t* __restrict__ leafRefOff,
                               uint32_t* __restrict__ leafCursor, uint32_t* __restrict__ leafIds) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    const bool valid = i < nRefs;
    const uint32_t nd = valid ? refNode[i] : 0xFFFFFFFFu;
    const bool uniform = warp_same_node(valid, nd);
    const uint32_t lane = threadIdx.x & 31u, below = (1u << lane) - 1u;
    if (uniform) {   // one node for the whole warp: one atomic per destination, the lanes take consecutive places
        const uint32_t t = refTri[i];
        const Decision d = dec[nd];
        if (d.axis < 0) {
            uint32_t base = 0;
            if (lane == 0u) base = atomicAdd(leafCursor + nd, 32u);
            base = __shfl_sync(0xFFFFFFFFu, base, 0);
            leafIds[leafIdBase + leafRefOff[nd] + base + lane] = t;
            return;
        }
        const uint32_t c = classify(nodes[nd], d, triLo[t], triHi[t]);
        const uint32_t r = splitRank[nd];
        const uint32_t ml = __ballot_sync(0xFFFFFFFFu, (c & 1u) != 0u), mr = __ballot_sync(0xFFFFFFFFu, (c & 2u) != 0u);
        uint32_t bl = 0, br = 0;
        if (lane == 0u) {
            if (ml) bl = atomicAdd(cursor + 2u * r, static_cast<uint32_t>(__popc(ml)));
            if (mr) br = atomicAdd(cursor + 2u * r + 1u, static_cast<uint32_t>(__popc(mr)));
        }
        bl = __shfl_sync(0xFFFFFFFFu, bl, 0); br = __shfl_sync(0xFFFFFFFFu, br, 0);
        if (c & 1u) {
            const uint32_t p = next[2u * r].begin + bl + __popc(ml & below);
            nextTri[p] = t; nextNode[p] = 2u * r;
        }
        if (c & 2u) {
            const uint32_t p = next[2u * r + 1u].begin + br + __popc(mr & below);
            nextTri[p] = t; nextNode[p] = 2u * r + 1u;
        }
        return;
    }
    if (!valid) return;
    const uint32_t t = refTri[i];
    const Decision d = dec[nd];
    if (d.axis < 0) {
        leafIds[leafIdBase + leafRefOff[nd] + atomicAdd(leafCursor + nd, 1u)] = t;
        return;
    }
    const uint32_t c = classify(nodes[nd], d, triLo[t], triHi[t]);
    const uint32_t r = splitRank[nd];
    if (c & 1u) {
        const uint32_t p = next[2u * r].begin + atomicAdd(cursor + 2u * r, 1u);
        nextTri[p] = t; nextNode[p] = 2u * r;
    }
    if (c & 2u) {
        const uint32_t p = next[2u * r + 1u].begin + atomicAdd(cursor + 2u * r + 1u, 1u);
        nextTri[p] = t; nextNode[p] = 2u * r + 1u;
    }
}

__global__ void bin_slots_kernel(const LevelNode* __restrict__ nodes, uint32_t nNodes, uint32_t* __restrict__ isBig) {
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k < nNodes) isBig[k] = (nodes[k].end - nodes[k].begin) > kSmall ? 1u : 0u;
}
__global__ void bin_slots_finish_kernel(const uint32_t* __restrict__ isBig, const uint32_t* __restrict__ rank, uint32_t nNodes, uint32_t* __restrict__ slot) {
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k < nNodes) slot[k] = isBig[k] ? rank[k] : 0xFFFFFFFFu;
}

// ascending ids inside every leaf (the scatter's atomics leave them in arrival order): the order the reference's std::set gives
__global__ void sort_leaves_kernel(const uint32_t* __restrict__ gAxis, const uint32_t* __restrict__ gLeafOff, const uint32_t* __restrict__ gLeafCnt,
                                   uint32_t nNodes, uint32_t* __restrict__ leafIds) {
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= nNodes || gAxis[k] != 0u) return;
    uint32_t* a = leafIds + gLeafOff[k];
    const uint32_t n = gLeafCnt[k];
    for (uint32_t i = 1; i < n; ++i) {
        const uint32_t v = a[i];
        uint32_t j = i;
        while (j > 0 && a[j - 1] > v) { a[j] = a[j - 1]; --j; }
        a[j] = v;
    }
}

// ---- device pack: the render-path tables (leaf blocks with neighbour links, location nodes, shading records) made where the
// tree already is, instead of on the host after a download and before an upload.  Same tables bit for bit as
// pack_scene + build_ropes (device_scene.cpp, kd_ropes.cpp) produce from the downloaded tree (tests compare them).
constexpr uint32_t kNoRope = 0xFFFFFFFFu, kRopeLeafBit = 0x80000000u, kHandleOffsetMask = 0x0FFFFFFFu;

// bottom-up, one level per launch: leaves and leaf references under every node
__global__ void under_kernel(uint32_t base, uint32_t n, const uint32_t* __restrict__ gLeft, const uint32_t* __restrict__ gAxis,
                             const uint32_t* __restrict__ gLeafCnt, uint32_t* __restrict__ lu, uint32_t* __restrict__ ru) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t g = base + i;
    if (gAxis[g] == 0u) { lu[g] = 1u; ru[g] = gLeafCnt[g]; return; }
    const uint32_t L = gLeft[g];
    lu[g] = lu[L] + lu[L + 1u]; ru[g] = ru[L] + ru[L + 1u];
}
// top-down: depth-first ordinal of the first leaf under a node and the references in front of it
__global__ void first_kernel(uint32_t base, uint32_t n, const uint32_t* __restrict__ gLeft, const uint32_t* __restrict__ gAxis,
                             const uint32_t* __restrict__ lu, const uint32_t* __restrict__ ru, uint32_t* __restrict__ first, uint32_t* __restrict__ off) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t g = base + i;
    if (gAxis[g] == 0u) return;
    const uint32_t L = gLeft[g];
    first[L] = first[g]; off[L] = off[g];
    first[L + 1u] = first[g] + lu[L]; off[L + 1u] = off[g] + ru[L];
}
// leaf handles (block offset in float4 units | min(count, 7) << 28) and the location-node table
__global__ void handle_kernel(uint32_t total, const uint32_t* __restrict__ gLeft, const uint32_t* __restrict__ gAxis, const float* __restrict__ gPlane,
                              const uint32_t* __restrict__ gLeafCnt, const uint32_t* __restrict__ first, const uint32_t* __restrict__ off,
                              uint32_t* __restrict__ handle, uint2* __restrict__ ropeNodes, uint32_t* __restrict__ overflow) {
    const uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= total) return;
    if (gAxis[g] != 0u) { ropeNodes[g] = make_uint2(__float_as_uint(gPlane[g]), (gLeft[g] << 2) | (gAxis[g] - 1u)); handle[g] = 0u; return; }
    const unsigned long long units = 4ull * first[g] + 3ull * off[g];
    if (units >= kHandleOffsetMask) atomicExch(overflow, 1u);
    const uint32_t cnt = gLeafCnt[g];
    const uint32_t h = static_cast<uint32_t>(units) | ((cnt < 7u ? cnt : 7u) << 28);
    handle[g] = h;
    ropeNodes[g] = make_uint2(h, (cnt << 2) | 3u);
}
// kd_ropes.cpp Builder::tighten
__device__ __forceinline__ uint32_t tighten_rope(uint32_t r, int fa, bool maxSide, const float* lo, const float* hi, const uint32_t* __restrict__ gLeft,
                                                 const uint32_t* __restrict__ gAxis, const float* __restrict__ gPlane) {
    while (r != kNoRope && gAxis[r] != 0u) {
        const int ra = static_cast<int>(gAxis[r]) - 1;
        const float s = gPlane[r];
        const uint32_t left = gLeft[r];
        if (ra == fa) r = maxSide ? left : left + 1u;
        else if (s >= hi[ra]) r = left;
        else if (s <= lo[ra]) r = left + 1u;
        else break;
    }
    return r;
}
// top-down, one level per launch: a node hands each child its six neighbour links, pushed down to the deepest covering node
__global__ void rope_level_kernel(uint32_t base, uint32_t n, const uint32_t* __restrict__ gLeft, const uint32_t* __restrict__ gAxis,
                                  const float* __restrict__ gPlane, const float* __restrict__ gBox, uint32_t* __restrict__ ropes) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t g = base + i;
    if (gAxis[g] == 0u) return;
    const int a = static_cast<int>(gAxis[g]) - 1;
    const uint32_t L = gLeft[g], R = L + 1u;
    uint32_t rl[6], rr[6];
    for (int f = 0; f < 6; ++f) { rl[f] = ropes[6 * static_cast<size_t>(g) + f]; rr[f] = rl[f]; }
    rl[2 * a + 1] = R;
    rr[2 * a] = L;
    float llo[3], lhi[3], rlo[3], rhi[3];
    for (int k = 0; k < 3; ++k) {
        llo[k] = gBox[6 * static_cast<size_t>(L) + k]; lhi[k] = gBox[6 * static_cast<size_t>(L) + 3 + k];
        rlo[k] = gBox[6 * static_cast<size_t>(R) + k]; rhi[k] = gBox[6 * static_cast<size_t>(R) + 3 + k];
    }
    for (int f = 0; f < 6; ++f) {
        ropes[6 * static_cast<size_t>(L) + f] = tighten_rope(rl[f], f >> 1, (f & 1) != 0, llo, lhi, gLeft, gAxis, gPlane);
        ropes[6 * static_cast<size_t>(R) + f] = tighten_rope(rr[f], f >> 1, (f & 1) != 0, rlo, rhi, gLeft, gAxis, gPlane);
    }
}
// one thread per leaf: header {cell.lo, first} {cell.hi, count} {links} + the triangle records of its id list
// (device_scene.cpp's tri_isect record: {a, c3} {a - b, id} {a - c, 0}, every operation rounded on its own like the host's)
__global__ void blob_kernel(uint32_t total, const uint32_t* __restrict__ gAxis, const float* __restrict__ gBox, const uint32_t* __restrict__ gLeafOff,
                            const uint32_t* __restrict__ gLeafCnt, const uint32_t* __restrict__ leafIds, const uint32_t* __restrict__ off,
                            const uint32_t* __restrict__ handle, const uint32_t* __restrict__ ropes, const float* __restrict__ verts,
                            const uint32_t* __restrict__ tris, float4* __restrict__ blob) {
    const uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= total || gAxis[g] != 0u) return;
    float4* blk = blob + (handle[g] & kHandleOffsetMask);
    const uint32_t cnt = gLeafCnt[g];
    const float* b = gBox + 6 * static_cast<size_t>(g);
    blk[0] = make_float4(b[0], b[1], b[2], __uint_as_float(off[g]));
    blk[1] = make_float4(b[3], b[4], b[5], __uint_as_float(cnt));
    uint32_t enc[6];
    for (int f = 0; f < 6; ++f) {
        const uint32_t r = ropes[6 * static_cast<size_t>(g) + f];
        enc[f] = (r != kNoRope && gAxis[r] == 0u) ? (kRopeLeafBit | handle[r]) : r;
    }
    blk[2] = make_float4(__uint_as_float(enc[0]), __uint_as_float(enc[1]), __uint_as_float(enc[2]), __uint_as_float(enc[3]));
    blk[3] = make_float4(__uint_as_float(enc[4]), __uint_as_float(enc[5]), 0.0f, 0.0f);
    for (uint32_t i = 0; i < cnt; ++i) {
        const uint32_t t = leafIds[gLeafOff[g] + i];
        const uint32_t* r = tris + 7 * static_cast<size_t>(t);
        const float* va = verts + 3 * static_cast<size_t>(r[0]);
        const float* vb = verts + 3 * static_cast<size_t>(r[1]);
        const float* vc = verts + 3 * static_cast<size_t>(r[2]);
        const float e1x = __fsub_rn(va[0], vb[0]), e1y = __fsub_rn(va[1], vb[1]), e1z = __fsub_rn(va[2], vb[2]);
        const float e2x = __fsub_rn(va[0], vc[0]), e2y = __fsub_rn(va[1], vc[1]), e2z = __fsub_rn(va[2], vc[2]);
        const float c3 = __fsub_rn(__fmul_rn(e1y, e2z), __fmul_rn(e2y, e1z));
        blk[4 + 3 * i] = make_float4(va[0], va[1], va[2], c3);
        blk[5 + 3 * i] = make_float4(e1x, e1y, e1z, __uint_as_float(t));
        blk[6 + 3 * i] = make_float4(e2x, e2y, e2z, 0.0f);
    }
}
__global__ void tri_shade_kernel(uint32_t nT, const float* __restrict__ normals, const uint32_t* __restrict__ tris, float4* __restrict__ out) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= nT) return;
    const uint32_t* r = tris + 7 * static_cast<size_t>(t);
    const float* n0 = normals + 3 * static_cast<size_t>(r[3]);
    const float* n1 = normals + 3 * static_cast<size_t>(r[4]);
    const float* n2 = normals + 3 * static_cast<size_t>(r[5]);
    out[3 * static_cast<size_t>(t)] = make_float4(n0[0], n0[1], n0[2], __uint_as_float(r[6]));
    out[3 * static_cast<size_t>(t) + 1] = make_float4(n1[0], n1[1], n1[2], 0.0f);
    out[3 * static_cast<size_t>(t) + 2] = make_float4(n2[0], n2[1], n2[2], 0.0f);
}

// Scratch comes from ONE grow-only device block kept between builds (cudaMalloc + cudaFree of the ~2 GB a 1 M-triangle build
// touches cost 250 ms per call -- five times the levels themselves); mcpt_shutdown releases it (release_gpu_builder_scratch).
struct Arena {
    char* base = nullptr;
    size_t cap = 0, used = 0;
} g_arena;

template <class T>
struct Dev {
    T* p = nullptr;
    size_t n = 0;
    cudaError_t alloc(size_t count) {   // sub-allocation: valid only after reserve_arena() sized the block for the whole build
        n = count;
        const size_t bytes = (std::max<size_t>(count, 1) * sizeof(T) + 255) & ~static_cast<size_t>(255);
        if (g_arena.used + bytes > g_arena.cap) return cudaErrorMemoryAllocation;
        p = reinterpret_cast<T*>(g_arena.base + g_arena.used);
        g_arena.used += bytes;
        return cudaSuccess;
    }
};

cudaError_t reserve_arena(size_t bytes) {
    g_arena.used = 0;
    if (bytes <= g_arena.cap) return cudaSuccess;
    if (g_arena.base) cudaFree(g_arena.base);
    g_arena.base = nullptr; g_arena.cap = 0;
    const cudaError_t e = cudaMalloc(reinterpret_cast<void**>(&g_arena.base), bytes);
    if (e == cudaSuccess) g_arena.cap = bytes;
    return e;
}

}  // namespace

void release_gpu_builder_scratch() {
    if (g_arena.base) cudaFree(g_arena.base);
    g_arena = Arena();
}

namespace {
// One attempt with room for `dup` references per triangle.  false + err: a CUDA error, or (tooSmall) the scene duplicates more
// than that -- the caller retries with more room.
bool build_gpu_tree_with(const HostScene& scene, HostTree& out, std::string& err, size_t dup, bool& tooSmall, DeviceTables* pack) {
    const auto t0 = std::chrono::steady_clock::now();
    const uint32_t nT = static_cast<uint32_t>(scene.num_triangles());
    const size_t refCap = dup * nT + 4096, nodeCap = (dup + 2) * nT + 4096;
    const float ci = tuning().sah_ci;
    const int maxDepth = std::min(31, static_cast<int>(8.0 + 1.3 * std::log2(static_cast<double>(std::max<uint32_t>(nT, 1)))));
    cudaStream_t s = nullptr;

    Dev<float> dVerts; Dev<uint32_t> dTris; Dev<float4> triLo, triHi;
    Dev<uint32_t> refTri[2], refNode[2];
    Dev<LevelNode> level[2];
    Dev<Decision> dec;
    Dev<uint32_t> nL, nR, isSplit, childRefs, leafRefs, splitRank, childRefOff, leafRefOff, cursor, leafCursor, isBig, bigRank, binSlot, bins;
    Dev<uint32_t> gLeft, gAxis, gLeafOff, gLeafCnt, leafIds;
    Dev<float> gBox, gPlane;
    Dev<unsigned char> scanTmp;
    Dev<unsigned long long> snapKey;
    const size_t levelCap = refCap + 2;   // a level cannot hold more non-empty nodes than references; empty children are bounded by the split count
    {
        size_t tmp = 0;
        cub::DeviceScan::ExclusiveSum(nullptr, tmp, static_cast<uint32_t*>(nullptr), static_cast<uint32_t*>(nullptr), static_cast<int>(2 * levelCap + 1), s);
        const size_t need = scene.vertices.size() * 4 + scene.triangles.size() * 4 + 32ull * nT + 4 * refCap * 4 + 2 * (2 * levelCap) * sizeof(LevelNode) +
                            2 * levelCap * sizeof(Decision) + 12 * (2 * levelCap + 1) * 4 + (4 * levelCap + 2) * 4 + (refCap / kSmall + 2) * (3 * 2 * kBins * 4 + 8) +
                            nodeCap * (4 * 4 + 24 + 4) + refCap * 4 + tmp + 256 + 128 * 256 + 2 * kBoxBlocks * 16 +
                            (pack ? scene.normals.size() * 4 + nodeCap * (5 * 4 + 24) + 16 * 256 : 0);
        GK_CHECK(reserve_arena(need));
    }
    GK_CHECK(dVerts.alloc(scene.vertices.size())); GK_CHECK(dTris.alloc(scene.triangles.size()));
    GK_CHECK(triLo.alloc(nT)); GK_CHECK(triHi.alloc(nT));
    Dev<float4> partLo, partHi;
    GK_CHECK(partLo.alloc(kBoxBlocks)); GK_CHECK(partHi.alloc(kBoxBlocks));
    for (int i = 0; i < 2; ++i) { GK_CHECK(refTri[i].alloc(refCap)); GK_CHECK(refNode[i].alloc(refCap)); GK_CHECK(level[i].alloc(2 * levelCap)); }
    GK_CHECK(dec.alloc(2 * levelCap));
    for (Dev<uint32_t>* d : {&nL, &nR, &isSplit, &childRefs, &leafRefs, &splitRank, &childRefOff, &leafRefOff, &leafCursor, &isBig, &bigRank, &binSlot})
        GK_CHECK(d->alloc(2 * levelCap + 1));
    GK_CHECK(cursor.alloc(4 * levelCap + 2));
    const size_t maxBig = refCap / kSmall + 2;
    GK_CHECK(bins.alloc(maxBig * 3 * 2 * kBins));
    GK_CHECK(gLeft.alloc(nodeCap)); GK_CHECK(gAxis.alloc(nodeCap)); GK_CHECK(gLeafOff.alloc(nodeCap)); GK_CHECK(gLeafCnt.alloc(nodeCap));
    GK_CHECK(gBox.alloc(6 * nodeCap)); GK_CHECK(gPlane.alloc(nodeCap)); GK_CHECK(leafIds.alloc(refCap));
    GK_CHECK(snapKey.alloc(refCap / kSmall + 2));
    size_t tmpBytes = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, tmpBytes, isSplit.p, splitRank.p, static_cast<int>(2 * levelCap + 1), s);
    GK_CHECK(scanTmp.alloc(tmpBytes + 256));

    const auto tAlloc = std::chrono::steady_clock::now();
    GK_CHECK(cudaMemcpyAsync(dVerts.p, scene.vertices.data(), scene.vertices.size() * 4, cudaMemcpyHostToDevice, s));
    GK_CHECK(cudaMemcpyAsync(dTris.p, scene.triangles.data(), scene.triangles.size() * 4, cudaMemcpyHostToDevice, s));
    const unsigned B = 256;
    auto grid = [&](size_t n) { return static_cast<unsigned>((n + B - 1) / B); };
    tri_bounds_kernel<<<grid(nT), B, 0, s>>>(dVerts.p, dTris.p, nT, triLo.p, triHi.p, refTri[0].p, refNode[0].p);
    scene_box_kernel<<<kBoxBlocks, 256, 0, s>>>(triLo.p, triHi.p, nT, nullptr, partLo.p, partHi.p, 0u);
    scene_box_kernel<<<1, 256, 0, s>>>(partLo.p, partHi.p, kBoxBlocks, level[0].p, nullptr, nullptr, nT);

    uint32_t nNodes = 1, nRefs = nT, levelBase = 0, leafIdBase = 0, depthReached = 0;
    std::vector<std::pair<uint32_t, uint32_t>> levels;   // (first node, node count) of every depth, for the device pack
    int cur = 0;
    for (int depth = 0; nNodes > 0; ++depth) {
        depthReached = static_cast<uint32_t>(depth);
        levels.emplace_back(levelBase, nNodes);
        const int nxt = cur ^ 1;
        if (static_cast<size_t>(levelBase) + nNodes > nodeCap) { err = "GPU KD build: node capacity exceeded"; tooSmall = true; return false; }
        // scan sizes include one extra element so that the last entry of each exclusive sum is the level's total
        GK_CHECK(cudaMemsetAsync(nL.p, 0, (nNodes + 1) * 4, s)); GK_CHECK(cudaMemsetAsync(nR.p, 0, (nNodes + 1) * 4, s));
        GK_CHECK(cudaMemsetAsync(leafCursor.p, 0, (nNodes + 1) * 4, s));
        GK_CHECK(cudaMemsetAsync(isBig.p + nNodes, 0, 4, s));
        bin_slots_kernel<<<grid(nNodes), B, 0, s>>>(level[cur].p, nNodes, isBig.p);
        cub::DeviceScan::ExclusiveSum(scanTmp.p, tmpBytes, isBig.p, bigRank.p, static_cast<int>(nNodes + 1), s);
        bin_slots_finish_kernel<<<grid(nNodes), B, 0, s>>>(isBig.p, bigRank.p, nNodes, binSlot.p);
        uint32_t numBig = 0;
        GK_CHECK(cudaMemcpyAsync(&numBig, bigRank.p + nNodes, 4, cudaMemcpyDeviceToHost, s));
        GK_CHECK(cudaStreamSynchronize(s));
        if (numBig > maxBig) { err = "GPU KD build: histogram capacity exceeded"; tooSmall = true; return false; }
        if (numBig) {
            GK_CHECK(cudaMemsetAsync(bins.p, 0, static_cast<size_t>(numBig) * 3 * 2 * kBins * 4, s));
            bin_kernel<<<grid(nRefs), B, 0, s>>>(level[cur].p, refTri[cur].p, refNode[cur].p, nRefs, triLo.p, triHi.p, binSlot.p, bins.p);
        }
        split_kernel<<<grid(nNodes), B, 0, s>>>(level[cur].p, nNodes, refTri[cur].p, triLo.p, triHi.p, binSlot.p, bins.p, ci, depth < maxDepth ? 1 : 0, dec.p);
        if (numBig && depth < maxDepth) {
            GK_CHECK(cudaMemsetAsync(snapKey.p, 0xFF, static_cast<size_t>(numBig) * 8, s));
            snap_kernel<<<grid(nRefs), B, 0, s>>>(level[cur].p, dec.p, refTri[cur].p, refNode[cur].p, nRefs, triLo.p, triHi.p, binSlot.p, snapKey.p);
            snap_apply_kernel<<<grid(nNodes), B, 0, s>>>(level[cur].p, nNodes, binSlot.p, snapKey.p, dec.p);
        }
        if (nRefs) count_kernel<<<grid(nRefs), B, 0, s>>>(level[cur].p, dec.p, refTri[cur].p, refNode[cur].p, nRefs, triLo.p, triHi.p, nL.p, nR.p);
        GK_CHECK(cudaMemsetAsync(isSplit.p + nNodes, 0, 4, s)); GK_CHECK(cudaMemsetAsync(childRefs.p + nNodes, 0, 4, s)); GK_CHECK(cudaMemsetAsync(leafRefs.p + nNodes, 0, 4, s));
        sizes_kernel<<<grid(nNodes), B, 0, s>>>(level[cur].p, dec.p, nNodes, nL.p, nR.p, isSplit.p, childRefs.p, leafRefs.p);
        cub::DeviceScan::ExclusiveSum(scanTmp.p, tmpBytes, isSplit.p, splitRank.p, static_cast<int>(nNodes + 1), s);
        cub::DeviceScan::ExclusiveSum(scanTmp.p, tmpBytes, childRefs.p, childRefOff.p, static_cast<int>(nNodes + 1), s);
        cub::DeviceScan::ExclusiveSum(scanTmp.p, tmpBytes, leafRefs.p, leafRefOff.p, static_cast<int>(nNodes + 1), s);
        uint32_t tot[3] = {0, 0, 0};
        GK_CHECK(cudaMemcpyAsync(&tot[0], splitRank.p + nNodes, 4, cudaMemcpyDeviceToHost, s));
        GK_CHECK(cudaMemcpyAsync(&tot[1], childRefOff.p + nNodes, 4, cudaMemcpyDeviceToHost, s));
        GK_CHECK(cudaMemcpyAsync(&tot[2], leafRefOff.p + nNodes, 4, cudaMemcpyDeviceToHost, s));
        GK_CHECK(cudaStreamSynchronize(s));
        const uint32_t numSplit = tot[0], nextRefs = tot[1], levelLeafRefs = tot[2];
        if (nextRefs > refCap || static_cast<size_t>(leafIdBase) + levelLeafRefs > refCap || 2ull * numSplit > 2 * levelCap) {
            err = "GPU KD build: reference capacity exceeded (duplication above " + std::to_string(dup) + "x)";
            tooSmall = true;
            return false;
        }
        const uint32_t nextBase = levelBase + nNodes;
        emit_kernel<<<grid(nNodes), B, 0, s>>>(level[cur].p, dec.p, nNodes, levelBase, nextBase, leafIdBase, nL.p, nR.p, splitRank.p, childRefOff.p, leafRefOff.p,
                                               level[nxt].p, gLeft.p, gBox.p, gAxis.p, gPlane.p, gLeafOff.p, gLeafCnt.p);
        if (numSplit) GK_CHECK(cudaMemsetAsync(cursor.p, 0, 2ull * numSplit * 4, s));
        if (nRefs) scatter_kernel<<<grid(nRefs), B, 0, s>>>(level[cur].p, dec.p, refTri[cur].p, refNode[cur].p, nRefs, triLo.p, triHi.p, splitRank.p, level[nxt].p,
                                                            cursor.p, refTri[nxt].p, refNode[nxt].p, leafIdBase, leafRefOff.p, leafCursor.p, leafIds.p);
        leafIdBase += levelLeafRefs;
        levelBase = nextBase;
        nNodes = 2u * numSplit;
        nRefs = nextRefs;
        cur = nxt;
    }
    const uint32_t total = levelBase;
    GK_CHECK(cudaStreamSynchronize(s));
    const auto tLevels = std::chrono::steady_clock::now();
    sort_leaves_kernel<<<grid(total), B, 0, s>>>(gAxis.p, gLeafOff.p, gLeafCnt.p, total, leafIds.p);
    GK_CHECK(cudaGetLastError());

    uint32_t packLeaves = 0;
    if (pack) {
        // ---- device pack (kernels above): the tree never leaves the device on its way into the render-path tables ----
        pack->done = false;
        const size_t nodeCapUsed = total;
        Dev<uint32_t> lu, ru, first, off, handle, ropes, overflow;
        Dev<float> dNormals;
        GK_CHECK(lu.alloc(nodeCapUsed)); GK_CHECK(ru.alloc(nodeCapUsed)); GK_CHECK(first.alloc(nodeCapUsed)); GK_CHECK(off.alloc(nodeCapUsed));
        GK_CHECK(handle.alloc(nodeCapUsed)); GK_CHECK(ropes.alloc(6 * nodeCapUsed)); GK_CHECK(overflow.alloc(1));
        GK_CHECK(dNormals.alloc(std::max<size_t>(scene.normals.size(), 1)));
        if (!scene.normals.empty()) GK_CHECK(cudaMemcpyAsync(dNormals.p, scene.normals.data(), scene.normals.size() * 4, cudaMemcpyHostToDevice, s));
        for (size_t d = levels.size(); d-- > 0;)
            under_kernel<<<grid(levels[d].second), B, 0, s>>>(levels[d].first, levels[d].second, gLeft.p, gAxis.p, gLeafCnt.p, lu.p, ru.p);
        GK_CHECK(cudaMemsetAsync(first.p, 0, 4, s)); GK_CHECK(cudaMemsetAsync(off.p, 0, 4, s)); GK_CHECK(cudaMemsetAsync(overflow.p, 0, 4, s));
        GK_CHECK(cudaMemsetAsync(ropes.p, 0xFF, 6 * 4, s));   // the root has no neighbours
        for (size_t d = 0; d < levels.size(); ++d) {
            first_kernel<<<grid(levels[d].second), B, 0, s>>>(levels[d].first, levels[d].second, gLeft.p, gAxis.p, lu.p, ru.p, first.p, off.p);
            rope_level_kernel<<<grid(levels[d].second), B, 0, s>>>(levels[d].first, levels[d].second, gLeft.p, gAxis.p, gPlane.p, gBox.p, ropes.p);
        }
        uint32_t counts[2] = {0, 0};
        GK_CHECK(cudaMemcpyAsync(&counts[0], lu.p, 4, cudaMemcpyDeviceToHost, s));
        GK_CHECK(cudaMemcpyAsync(&counts[1], ru.p, 4, cudaMemcpyDeviceToHost, s));
        GK_CHECK(cudaStreamSynchronize(s));
        const size_t units = 4ull * counts[0] + 3ull * counts[1];
        packLeaves = counts[0];
        if (counts[1] != leafIdBase) { err = "GPU KD pack: reference count mismatch"; return false; }
        if (units >= kHandleOffsetMask) { err = "scene too large: more than 4 GiB of leaf blocks (28-bit leaf handles)"; return false; }
        GK_CHECK(pack->leaf_blob->ensure(units + 3)); GK_CHECK(pack->rope_nodes->ensure(total)); GK_CHECK(pack->tri_shade->ensure(3ull * nT));
        handle_kernel<<<grid(total), B, 0, s>>>(total, gLeft.p, gAxis.p, gPlane.p, gLeafCnt.p, first.p, off.p, handle.p, pack->rope_nodes->ptr, overflow.p);
        blob_kernel<<<grid(total), B, 0, s>>>(total, gAxis.p, gBox.p, gLeafOff.p, gLeafCnt.p, leafIds.p, off.p, handle.p, ropes.p, dVerts.p, dTris.p, pack->leaf_blob->ptr);
        GK_CHECK(cudaMemsetAsync(pack->leaf_blob->ptr + units, 0, 3 * sizeof(float4), s));   // a lane may read one record past a leaf of 7+ triangles' known count
        tri_shade_kernel<<<grid(nT), B, 0, s>>>(nT, dNormals.p, dTris.p, pack->tri_shade->ptr);
        GK_CHECK(cudaGetLastError());
        pack->blob_float4s = units + 3;
        pack->done = true;     // complete once the stream has drained (the synchronize below)
    }

    out = HostTree();
    if (pack && pack->done && pack->tree_words && pack->tree_floats) {
        // the tree stays on the device (context buffers, so that it outlives the next build's scratch); the host gets the counts
        const size_t n = total;
        GK_CHECK(pack->tree_words->ensure(4 * n + leafIdBase)); GK_CHECK(pack->tree_floats->ensure(7 * n));
        uint32_t* w = pack->tree_words->ptr;
        float* f = pack->tree_floats->ptr;
        GK_CHECK(cudaMemcpyAsync(w, gLeft.p, n * 4, cudaMemcpyDeviceToDevice, s));
        GK_CHECK(cudaMemcpyAsync(w + n, gAxis.p, n * 4, cudaMemcpyDeviceToDevice, s));
        GK_CHECK(cudaMemcpyAsync(w + 2 * n, gLeafOff.p, n * 4, cudaMemcpyDeviceToDevice, s));
        GK_CHECK(cudaMemcpyAsync(w + 3 * n, gLeafCnt.p, n * 4, cudaMemcpyDeviceToDevice, s));
        if (leafIdBase) GK_CHECK(cudaMemcpyAsync(w + 4 * n, leafIds.p, leafIdBase * 4ull, cudaMemcpyDeviceToDevice, s));
        GK_CHECK(cudaMemcpyAsync(f, gBox.p, n * 24, cudaMemcpyDeviceToDevice, s));
        GK_CHECK(cudaMemcpyAsync(f + 6 * n, gPlane.p, n * 4, cudaMemcpyDeviceToDevice, s));
        GK_CHECK(cudaMemcpyAsync(out.root_box, gBox.p, 24, cudaMemcpyDeviceToHost, s));
        GK_CHECK(cudaStreamSynchronize(s));
        out.on_host = false; out.nodes = n; out.refs = leafIdBase;
        out.num_leaves = packLeaves;
    } else {
        out.left.resize(total); out.aabb.resize(6 * static_cast<size_t>(total)); out.split_axis.resize(total); out.split_value.resize(total);
        out.leaf_offset.resize(total); out.leaf_count.resize(total); out.leaf_ids.resize(leafIdBase);
        GK_CHECK(cudaMemcpyAsync(out.left.data(), gLeft.p, total * 4ull, cudaMemcpyDeviceToHost, s));
        GK_CHECK(cudaMemcpyAsync(out.aabb.data(), gBox.p, total * 24ull, cudaMemcpyDeviceToHost, s));
        GK_CHECK(cudaMemcpyAsync(out.split_axis.data(), gAxis.p, total * 4ull, cudaMemcpyDeviceToHost, s));
        GK_CHECK(cudaMemcpyAsync(out.split_value.data(), gPlane.p, total * 4ull, cudaMemcpyDeviceToHost, s));
        GK_CHECK(cudaMemcpyAsync(out.leaf_offset.data(), gLeafOff.p, total * 4ull, cudaMemcpyDeviceToHost, s));
        GK_CHECK(cudaMemcpyAsync(out.leaf_count.data(), gLeafCnt.p, total * 4ull, cudaMemcpyDeviceToHost, s));
        if (leafIdBase) GK_CHECK(cudaMemcpyAsync(out.leaf_ids.data(), leafIds.p, leafIdBase * 4ull, cudaMemcpyDeviceToHost, s));
        GK_CHECK(cudaStreamSynchronize(s));
        for (uint32_t k = 0; k < total; ++k) out.num_leaves += out.split_axis[k] == 0 ? 1u : 0u;
    }
    out.max_depth = depthReached;
    out.build_seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    if (tuning().build_timing) {
        auto sec = [](std::chrono::steady_clock::time_point a, std::chrono::steady_clock::time_point b) { return std::chrono::duration<double>(b - a).count(); };
        std::fprintf(stderr, "gpu kd build: %u nodes, %u leaf refs, depth %u, %.3f s = scratch allocation %.3f + H2D and levels %.3f + leaf sort and D2H %.3f\n", total,
                     leafIdBase, depthReached, out.build_seconds, sec(t0, tAlloc), sec(tAlloc, tLevels), sec(tLevels, std::chrono::steady_clock::now()));
    }
    return true;
}
}  // namespace

// Needs a bound CUDA device (the caller checks).  The scratch is sized for 6 references per triangle (tessellated surfaces
// duplicate 1.3-2.5x); a scene that duplicates more -- a soup of large overlapping triangles -- is built again with room for
// 16x, 40x, 100x.  false + err: beyond that, out of device memory, or a CUDA error.
bool build_gpu_tree(const HostScene& scene, HostTree& out, std::string& err, DeviceTables* pack) {
    const auto t0 = std::chrono::steady_clock::now();
    for (const size_t dup : {6u, 16u, 40u, 100u}) {
        bool tooSmall = false;
        if (build_gpu_tree_with(scene, out, err, dup, tooSmall, pack)) {
            out.build_seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();   // retries included
            return true;
        }
        cudaGetLastError();
        if (!tooSmall) return false;
    }
    return false;
}

}  // namespace mcpt
