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
//   leaf_blob   one block per leaf, leaves in depth-first order: 4 x float4 of header
//               {cell.lo, first slot in leaf_ids} {cell.hi, count} {rope -x, +x, -y, +y} {rope -z, +z, -, -}
//               followed by the leaf's `count` triangle records (3 x float4 each, the tri_isect records of its ids).  A leaf is
//               named by its HANDLE: bits 0-27 = offset of the block in float4 units, bits 28-30 = min(count, 7).
//               With the handle alone a ray knows where the triangle records are and (for all but leaves of 7+
//               triangles) how many: the triangle tests do not wait for the header, which is only needed for the
//               exit face afterwards -- one dependent memory round trip less per leaf visit.
//               A rope is 0xFFFFFFFF = outside the scene; bit 31 set = the neighbour is one leaf: its handle in the
//               low bits, no node to load; else a node index.
//   rope_nodes  uint2 per node: inner = kd_nodes entry; leaf = {leaf handle, (count << 2) | 3}
#include <algorithm>
#include <cmath>
#include <cstring>
#include <string>
#include <utility>

#include "device_scene.h"

namespace mcpt {
namespace {

inline float as_float(uint32_t u) { float f; std::memcpy(&f, &u, 4); return f; }
const uint32_t kNone = 0xFFFFFFFFu;
const uint32_t kRopeIsLeaf = 0x80000000u;      // same constants as MCPT_ROPE_IS_LEAF, MCPT_LEAF_OFF, MCPT_LEAF_CNT in intersect.cuh
const uint32_t kLeafOffsetMask = 0x0FFFFFFFu;

struct Cell {
    float lo[3], hi[3];
};

// What a node hands its children: its cell and, per face, the deepest node covering that face from the other side.
struct NodeCtx {
    Cell cell;
    uint32_t rope[6];
};

struct Builder {
    const HostTree& t;
    PackedScene& out;
    std::vector<uint32_t> leafHandle;  // per node (leaves only)

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

    // A leaf's block: header {cell.lo, first} {cell.hi, count} {ropes} + its triangle records, read from tri_isect through
    // the leaf's id list (the record of leaf_ids[first + i] with the triangle id in its free lane).
    void write_leaf(uint32_t node, const NodeCtx& c) {
        float4* blk = out.leaf_blob.data() + (leafHandle[node] & kLeafOffsetMask);
        const uint32_t first = out.kd_nodes[node].x, count = t.leaf_count[node];
        blk[0] = make_float4(c.cell.lo[0], c.cell.lo[1], c.cell.lo[2], as_float(first));
        blk[1] = make_float4(c.cell.hi[0], c.cell.hi[1], c.cell.hi[2], as_float(count));
        uint32_t enc[6];
        for (int f = 0; f < 6; ++f) {
            const uint32_t r = c.rope[f];
            enc[f] = (r != kNone && t.split_axis[r] == 0) ? (kRopeIsLeaf | leafHandle[r]) : r;
        }
        blk[2] = make_float4(as_float(enc[0]), as_float(enc[1]), as_float(enc[2]), as_float(enc[3]));
        blk[3] = make_float4(as_float(enc[4]), as_float(enc[5]), 0.0f, 0.0f);
        for (uint32_t i = 0; i < count; ++i) {
            const float4* rec = out.tri_isect.data() + 3 * static_cast<size_t>(out.leaf_ids[first + i]);
            blk[4 + 3 * i] = rec[0]; blk[5 + 3 * i] = rec[1]; blk[6 + 3 * i] = rec[2];
        }
    }
    void split(uint32_t node, const NodeCtx& c, NodeCtx& l, NodeCtx& r) const {
        const int a = static_cast<int>(t.split_axis[node]) - 1;
        const float s = t.split_value[node];
        const uint32_t L = t.left[node], R = L + 1;
        l = c; r = c;
        l.cell.hi[a] = s;
        r.cell.lo[a] = s;
        l.rope[2 * a + 1] = R;
        r.rope[2 * a] = L;
        for (int f = 0; f < 6; ++f) {
            l.rope[f] = tighten(l.rope[f], f >> 1, (f & 1) != 0, l.cell);
            r.rope[f] = tighten(r.rope[f], f >> 1, (f & 1) != 0, r.cell);
        }
    }

    // Depth first, the upper levels fanned out as tasks: subtrees write disjoint, CONTIGUOUS runs of leaf blocks (the blocks
    // are laid out in depth-first leaf order), which is why this beats a level-synchronous sweep whose leaves of one depth
    // are scattered over the whole blob (measured at 1 M triangles: 150 ms against 250 ms on 8 cores).
    void walk(uint32_t node, const NodeCtx& c, int depth = 0) {
        if (t.split_axis[node] == 0) { write_leaf(node, c); return; }
        NodeCtx l, r;
        split(node, c, l, r);
        const uint32_t L = t.left[node];
        if (depth < 7) {
#pragma omp task firstprivate(L, l, depth)
            walk(L, l, depth + 1);
            walk(L + 1, r, depth + 1);
#pragma omp taskwait
        } else {
            walk(L, l, depth + 1);
            walk(L + 1, r, depth + 1);
        }
    }
};

}  // namespace

bool build_ropes(const HostTree& tree, PackedScene& out) {
    const size_t nN = tree.num_nodes();
    Builder b{tree, out, std::vector<uint32_t>(nN, 0u)};
    // blocks in depth-first leaf order (pack_scene's leaf_order): offsets by a prefix sum over the ordinals
    std::vector<uint32_t> nodeOfOrdinal;
    for (size_t k = 0; k < nN; ++k)
        if (tree.split_axis[k] == 0) {
            if (out.leaf_order[k] >= nodeOfOrdinal.size()) nodeOfOrdinal.resize(out.leaf_order[k] + 1u, 0u);
            nodeOfOrdinal[out.leaf_order[k]] = static_cast<uint32_t>(k);
        }
    out.rope_nodes = out.kd_nodes;
    size_t units = 0;
    for (const uint32_t k : nodeOfOrdinal) {
        if (units >= kLeafOffsetMask) return false;  // 4 GiB of leaf blocks: the 28-bit offset is exhausted
        const uint32_t count = tree.leaf_count[k];
        b.leafHandle[k] = static_cast<uint32_t>(units) | ((count < 7u ? count : 7u) << 28);
        out.rope_nodes[k].x = b.leafHandle[k];
        units += 4u + 3u * static_cast<size_t>(count);
    }
    // three float4 of padding: a lane may read one record past a leaf of 7+ triangles' known count, never past this
    out.leaf_blob.assign(units + 3, make_float4(0, 0, 0, 0));
    NodeCtx root;
    for (int i = 0; i < 3; ++i) { root.cell.lo[i] = tree.aabb[i]; root.cell.hi[i] = tree.aabb[3 + i]; }
    for (int f = 0; f < 6; ++f) root.rope[f] = kNone;
#pragma omp parallel if (nN > 8192)
#pragma omp single
    b.walk(0u, root);
    return true;
}

// Structural check of what build_ropes produced (host only; mcpt_check_stackless_layout, CPU tests):
//   * the blocks follow each other in leaf order, every handle carries its block's offset and min(count, 7), the
//     header repeats the tree's first/count and the records are the tri_isect records of the leaf's id list;
//   * for every leaf and face: a missing rope only where the face lies on the scene box; otherwise the rope's cell touches the face
//     from the other side and covers it in the two other axes (so the cell behind any exit point is below the rope),
//     a leaf rope names a real leaf, and a node rope is the deepest possible (its plane cuts the face).
// Cells are recomputed from the split planes, independently of Builder::walk's bookkeeping.
bool check_ropes(const HostTree& t, const PackedScene& out, std::string& err) {
    const size_t nN = t.num_nodes();
    std::vector<Cell> cell(nN);
    for (int i = 0; i < 3; ++i) { cell[0].lo[i] = t.aabb[i]; cell[0].hi[i] = t.aabb[3 + i]; }
    for (size_t k = 0; k < nN; ++k) {  // BFS order: a parent comes before its children
        if (t.split_axis[k] == 0) continue;
        const int a = static_cast<int>(t.split_axis[k]) - 1;
        const uint32_t L = t.left[k];
        if (L + 1 >= nN) { err = "child index out of range"; return false; }
        cell[L] = cell[k]; cell[L + 1] = cell[k];
        cell[L].hi[a] = t.split_value[k];
        cell[L + 1].lo[a] = t.split_value[k];
    }
    std::vector<std::pair<uint32_t, uint32_t>> byOrdinal;  // (ordinal, node)
    for (size_t k = 0; k < nN; ++k)
        if (t.split_axis[k] == 0) byOrdinal.emplace_back(out.leaf_order[k], static_cast<uint32_t>(k));
    std::sort(byOrdinal.begin(), byOrdinal.end());
    if (out.rope_nodes.size() != nN) { err = "rope_nodes size"; return false; }
    std::vector<uint32_t> nodeOfOffset;  // block offset -> leaf node (sparse)
    size_t units = 0;
    auto bits = [](float f) { uint32_t u; std::memcpy(&u, &f, 4); return u; };
    for (const auto& on : byOrdinal) {
        const uint32_t k = on.second, count = t.leaf_count[k], h = out.rope_nodes[k].x;
        if ((h & kLeafOffsetMask) != units || ((h >> 28) & 7u) != (count < 7u ? count : 7u) || (h & kRopeIsLeaf)) { err = "leaf handle of node " + std::to_string(k); return false; }
        if (units + 4 + 3 * static_cast<size_t>(count) > out.leaf_blob.size()) { err = "leaf block past the end"; return false; }
        const float4* blk = out.leaf_blob.data() + units;
        if (bits(blk[0].w) != out.kd_nodes[k].x || bits(blk[1].w) != count) { err = "leaf header first/count of node " + std::to_string(k); return false; }
        for (int i = 0; i < 3; ++i) {
            const float lo = i == 0 ? blk[0].x : (i == 1 ? blk[0].y : blk[0].z), hi = i == 0 ? blk[1].x : (i == 1 ? blk[1].y : blk[1].z);
            if (lo != cell[k].lo[i] || hi != cell[k].hi[i]) { err = "leaf cell of node " + std::to_string(k); return false; }
        }
        for (uint32_t i = 0; i < count; ++i)
            if (std::memcmp(blk + 4 + 3 * i, out.tri_isect.data() + 3 * static_cast<size_t>(out.leaf_ids[out.kd_nodes[k].x + i]), 48) != 0) { err = "leaf records of node " + std::to_string(k); return false; }
        if (nodeOfOffset.size() <= units) nodeOfOffset.resize(units + 1, kNone);
        nodeOfOffset[units] = k;
        units += 4 + 3 * static_cast<size_t>(count);
    }
    if (out.leaf_blob.size() != units + 3) { err = "leaf_blob size"; return false; }
    for (const auto& on : byOrdinal) {
        const uint32_t k = on.second;
        const Cell& c = cell[k];
        const float4* blk = out.leaf_blob.data() + (out.rope_nodes[k].x & kLeafOffsetMask);
        const uint32_t rope[6] = {bits(blk[2].x), bits(blk[2].y), bits(blk[2].z), bits(blk[2].w), bits(blk[3].x), bits(blk[3].y)};
        for (int f = 0; f < 6; ++f) {
            const int a = f >> 1;
            const bool maxSide = (f & 1) != 0;
            const std::string where = "rope " + std::to_string(f) + " of node " + std::to_string(k);
            const bool onBox = maxSide ? c.hi[a] == cell[0].hi[a] : c.lo[a] == cell[0].lo[a];
            if (rope[f] == kNone) {
                if (!onBox) { err = where + ": missing"; return false; }
                continue;
            }
            // (a face ON the scene box can still have a neighbour: a split plane that coincides with the box leaves a flat cell)
            uint32_t target;
            if (rope[f] & kRopeIsLeaf) {
                const uint32_t off = rope[f] & kLeafOffsetMask;
                if (off >= nodeOfOffset.size() || nodeOfOffset[off] == kNone) { err = where + ": not a leaf handle"; return false; }
                target = nodeOfOffset[off];
                if (out.rope_nodes[target].x != (rope[f] & ~kRopeIsLeaf)) { err = where + ": handle bits"; return false; }
            } else {
                target = rope[f];
                if (target >= nN || t.split_axis[target] == 0) { err = where + ": not an inner node"; return false; }
                const int sa = static_cast<int>(t.split_axis[target]) - 1;
                const float sv = t.split_value[target];
                if (sa == a || !(sv > c.lo[sa] && sv < c.hi[sa])) { err = where + ": not the deepest covering node"; return false; }
            }
            const Cell& n = cell[target];
            if (maxSide ? n.lo[a] != c.hi[a] : n.hi[a] != c.lo[a]) { err = where + ": target does not touch the face"; return false; }
            for (int b = 0; b < 3; ++b)
                if (b != a && !(n.lo[b] <= c.lo[b] && n.hi[b] >= c.hi[b])) { err = where + ": target does not cover the face"; return false; }
        }
    }
    return true;
}


// Host mirror of the WALK of intersect_ropes (intersect.cuh) -- no triangle tests, same float operations (IEEE
// reciprocal, separate multiply and add, fminf/fmaxf): the leaves a ray visits, in order, with the exit distance and axis
// of each.  A diagnostic for the tables (mcpt_debug_stackless_walk; tests use it to show WHICH cells a ray crosses), not
// a render path: the product traces rays on the GPU only.
int walk_ropes_host(const PackedScene& ps, const float ray[6], int maxHops, uint32_t* leafOff, float* exitT, uint32_t* exitAxis) {
    const float ox = ray[0], oy = ray[1], oz = ray[2], dx = ray[3], dy = ray[4], dz = ray[5];
    const float ix = 1.0f / dx, iy = 1.0f / dy, iz = 1.0f / dz;
    const uint32_t neg = (ix < 0.0f ? 1u : 0u) | (iy < 0.0f ? 2u : 0u) | (iz < 0.0f ? 4u : 0u);
    const float o[3] = {ox, oy, oz}, inv[3] = {ix, iy, iz};
    float t0 = 0.0f, t1 = 10000.0f;
    for (int a = 0; a < 3; ++a) {
        float tn = (ps.bmin[a] - o[a]) * inv[a], tf = (ps.bmax[a] - o[a]) * inv[a];
        if (inv[a] < 0.0f) std::swap(tn, tf);
        t0 = fmaxf(t0, tn); t1 = fminf(t1, tf);
    }
    if (!(t0 * (1.0f - 4e-7f) <= t1 * (1.0f + 4e-7f))) return 0;
    auto locate = [&](uint32_t node, float px, float py, float pz) {
        uint2 n = ps.rope_nodes[node];
        while ((n.y & 3u) != 3u) {
            const uint32_t axis = n.y & 3u;
            const float c = axis == 0u ? px : (axis == 1u ? py : pz);
            node = (n.y >> 2) + (c >= as_float(n.x) ? 1u : 0u);
            n = ps.rope_nodes[node];
        }
        return n.x;
    };
    float tcur = t0;
    float px = fminf(fmaxf(ox + t0 * dx, ps.bmin[0]), ps.bmax[0]), py = fminf(fmaxf(oy + t0 * dy, ps.bmin[1]), ps.bmax[1]),
          pz = fminf(fmaxf(oz + t0 * dz, ps.bmin[2]), ps.bmax[2]);
    uint32_t leaf = locate(0u, px, py, pz);
    int hops = 0;
    for (; hops < maxHops; ++hops) {
        const uint32_t off = leaf & kLeafOffsetMask;
        const float4 h0 = ps.leaf_blob[off], h1 = ps.leaf_blob[off + 1];
        const float plx = (neg & 1u) ? h0.x : h1.x, ply = (neg & 2u) ? h0.y : h1.y, plz = (neg & 4u) ? h0.z : h1.z;
        const float tx = (plx - ox) * ix, ty = (ply - oy) * iy, tz = (plz - oz) * iz;
        const float te = fminf(tx, fminf(ty, tz));
        const uint32_t axis = tx == te ? 0u : (ty == te ? 1u : 2u);
        const float plane = axis == 0u ? plx : (axis == 1u ? ply : plz);
        const uint32_t face = 2u * axis + (((neg >> axis) & 1u) ^ 1u);
        uint32_t rope;
        std::memcpy(&rope, reinterpret_cast<const char*>(&ps.leaf_blob[off + 2]) + 4 * face, 4);
        leafOff[hops] = off; exitT[hops] = te; exitAxis[hops] = axis;
        if (rope == kNone) { ++hops; break; }
        tcur = fmaxf(tcur, te);
        if (rope & kRopeIsLeaf) {
            leaf = rope & ~kRopeIsLeaf;
        } else {
            px = ox + tcur * dx; py = oy + tcur * dy; pz = oz + tcur * dz;
            if (axis == 0u) px = plane;
            if (axis == 1u) py = plane;
            if (axis == 2u) pz = plane;
            leaf = locate(rope, px, py, pz);
        }
    }
    return hops;
}

}  // namespace mcpt
