// Neighbour links ("ropes") for stackless KD traversal (host, product code; no reference counterpart).
//
// The reference walks its tree with a 64-entry stack from the root for every ray (rtx.hlsl:144-211).  For
// the render path a ray instead starts in the LEAF that contains its origin and moves from leaf to leaf
// through the face it exits (Popov et al. 2007, "Stackless KD-tree traversal for high performance GPU ray
// tracing"): every leaf stores its cell and, per face, the deepest tree node that covers the whole face on
// the other side.  Long descents from the root -- the dominant source of warp divergence in the stack
// traversal (profiles/r1_tuning.md) -- disappear; what remains per hop is a short point location below the
// rope's target.
//
// Cells are derived from the split planes alone (root cell = root box), so this works for both builders:
// the reference-rule tree stores CLIPPED boxes per node (KDTree.hpp:154-155), which are not a partition.
//
//   rope_nodes  uint2 per node: inner = kd_nodes entry; leaf = {leaf ordinal, (count << 2) | 3}
//   leaf_hdr    4 x float4 per leaf: {cell.lo, first record in leaf_tris} {cell.hi, count}
//               {rope -x, +x, -y, +y} {rope -z, +z, -, -}   (0xFFFFFFFF = outside the scene; bit 31 set = the
//               neighbour is one leaf: its ordinal in the low bits, no node to load; else a node index)
#include <cstring>

#include "device_scene.h"

namespace mcpt {
namespace {

inline float as_float(uint32_t u) { float f; std::memcpy(&f, &u, 4); return f; }
const uint32_t kNone = 0xFFFFFFFFu;
const uint32_t kRopeIsLeaf = 0x80000000u;  // same constant as MCPT_ROPE_IS_LEAF in intersect.cuh

struct Cell {
    float lo[3], hi[3];
};

struct Builder {
    const HostTree& t;
    PackedScene& out;
    std::vector<uint32_t> leafOrdinal;

    // Pushes rope `r` of a face (axis fa, min side if !maxSide) of `cell` down to the deepest node that still
    // covers the whole face.
    uint32_t tighten(uint32_t r, int fa, bool maxSide, const Cell& cell) const {
        while (r != kNone && t.split_axis[r] != 0) {
            const int ra = static_cast<int>(t.split_axis[r]) - 1;
            const float s = t.split_value[r];
            const uint32_t left = t.left[r];
            if (ra == fa) r = maxSide ? left : left + 1;          // the child adjacent to the shared face
            else if (s >= cell.hi[ra]) r = left;                  // the face lies entirely below the plane
            else if (s <= cell.lo[ra]) r = left + 1;              // ... or entirely above it
            else break;
        }
        return r;
    }

    void walk(uint32_t node, const Cell& cell, const uint32_t ropes[6]) {
        if (t.split_axis[node] == 0) {
            const uint32_t k = leafOrdinal[node];
            out.leaf_hdr[4 * k] = make_float4(cell.lo[0], cell.lo[1], cell.lo[2], as_float(out.kd_nodes[node].x));
            out.leaf_hdr[4 * k + 1] = make_float4(cell.hi[0], cell.hi[1], cell.hi[2], as_float(t.leaf_count[node]));
            uint32_t enc[6];
            for (int f = 0; f < 6; ++f) {
                const uint32_t r = ropes[f];
                enc[f] = (r != kNone && t.split_axis[r] == 0) ? (kRopeIsLeaf | leafOrdinal[r]) : r;
            }
            out.leaf_hdr[4 * k + 2] = make_float4(as_float(enc[0]), as_float(enc[1]), as_float(enc[2]), as_float(enc[3]));
            out.leaf_hdr[4 * k + 3] = make_float4(as_float(enc[4]), as_float(enc[5]), 0.0f, 0.0f);
            return;
        }
        const int a = static_cast<int>(t.split_axis[node]) - 1;
        const float s = t.split_value[node];
        const uint32_t L = t.left[node], R = L + 1;
        Cell cl = cell, cr = cell;
        cl.hi[a] = s;
        cr.lo[a] = s;
        uint32_t rl[6], rr[6];
        for (int f = 0; f < 6; ++f) { rl[f] = ropes[f]; rr[f] = ropes[f]; }
        rl[2 * a + 1] = R;
        rr[2 * a] = L;
        for (int f = 0; f < 6; ++f) {
            rl[f] = tighten(rl[f], f >> 1, (f & 1) != 0, cl);
            rr[f] = tighten(rr[f], f >> 1, (f & 1) != 0, cr);
        }
        walk(L, cl, rl);
        walk(R, cr, rr);
    }
};

}  // namespace

void build_ropes(const HostTree& tree, PackedScene& out) {
    const size_t nN = tree.num_nodes();
    Builder b{tree, out, std::vector<uint32_t>(nN, 0u)};
    uint32_t leaves = 0;
    out.rope_nodes = out.kd_nodes;
    for (size_t k = 0; k < nN; ++k)
        if (tree.split_axis[k] == 0) {
            b.leafOrdinal[k] = out.leaf_order[k];  // depth-first ordinal (pack_scene)
            out.rope_nodes[k].x = out.leaf_order[k];
            ++leaves;
        }
    out.leaf_hdr.assign(4 * static_cast<size_t>(leaves), make_float4(0, 0, 0, 0));
    Cell root;
    for (int i = 0; i < 3; ++i) { root.lo[i] = tree.aabb[i]; root.hi[i] = tree.aabb[3 + i]; }
    const uint32_t none[6] = {kNone, kNone, kNone, kNone, kNone, kNone};
    b.walk(0u, root, none);
}

}  // namespace mcpt
