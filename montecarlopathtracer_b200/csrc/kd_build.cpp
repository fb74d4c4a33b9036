// Host KD-tree builder (product code).
//
// MCPT_BUILD_REFERENCE reproduces, node for node, the tree the reference builds in
// MCRT/QuinEngine/Utils/KDTree.hpp:58-287 and uploads in ShaderResource.hpp:128-179:
//   * breadth-first, depth cap 32 (:103), node order = BFS order, children adjacent;
//   * > 64 triangles: spatial median of the longest box axis (:108-122);
//   * otherwise exhaustive SAH over every vertex coordinate on the three axes, ordered by
//     (axis, value), cost (A_L n_L + A_R n_R)/A_0 + 0 against n, first strict minimum (:166-240);
//   * straddle rule: flat on the plane -> left only, min < plane -> left, max > plane -> right;
//   * child box = parent box cut at the plane, intersected with the union of the child's
//     triangle boxes (:154-155, :276-277).
// The intersection-parity gate needs the SAME leaf sets and boxes, so every float expression
// here keeps the reference's operand order (this file is compiled with -ffp-contract=off).
// Unlike the reference it is level-synchronous: all nodes of one depth are independent and are
// split in parallel (OpenMP), then appended in order, which yields the identical BFS numbering
// without std::set / std::list (1.0 M triangles: seconds instead of the reference's 25 s).
#include <omp.h>

#include <algorithm>
#include <cfloat>
#include <chrono>

#include "host_scene.h"

namespace mcpt {
namespace {

struct Box {
    float lo[3], hi[3];
};

inline Box empty_box() { return {{FLT_MAX, FLT_MAX, FLT_MAX}, {-FLT_MAX, -FLT_MAX, -FLT_MAX}}; }

inline void grow(Box& a, const Box& b) {
    for (int i = 0; i < 3; ++i) {
        a.lo[i] = std::min(a.lo[i], b.lo[i]);
        a.hi[i] = std::max(a.hi[i], b.hi[i]);
    }
}

inline void intersect_with(Box& a, const Box& b) {
    for (int i = 0; i < 3; ++i) {
        a.lo[i] = std::max(a.lo[i], b.lo[i]);
        a.hi[i] = std::min(a.hi[i], b.hi[i]);
    }
}

inline float extent_area(const Box& b) {
    float ex = b.hi[0] - b.lo[0], ey = b.hi[1] - b.lo[1], ez = b.hi[2] - b.lo[2];
    return ex * ey + ey * ez + ez * ex;
}

// bit 0: goes left, bit 1: goes right
inline unsigned classify(const Box& t, int axis, float plane) {
    const float lo = t.lo[axis], hi = t.hi[axis];
    if (lo == hi && lo == plane) return 1u;
    return (lo < plane ? 1u : 0u) | (hi > plane ? 2u : 0u);
}

struct BuildNode {
    Box box;
    std::vector<uint32_t> ids;
    int axis = -1;  // -1 leaf
    float plane = 0.0f;
    uint32_t child = 0;
};

struct Decision {
    int axis = -1;
    float plane = 0.0f;
    BuildNode l, r;
};

struct Params {
    uint32_t median_above = 64;  // KDTree.hpp:108
    uint32_t depth_cap = 32;     // KDTree.hpp:103
    float traversal_cost = 0.0f; // KDTree.hpp:166 (Cts)
};

void apply_split(const std::vector<Box>& triBox, const BuildNode& n, int axis, float plane, Decision& d) {
    d.axis = axis;
    d.plane = plane;
    d.l.box = n.box; d.l.box.hi[axis] = plane;
    d.r.box = n.box; d.r.box.lo[axis] = plane;
    Box bl = empty_box(), br = empty_box();
    d.l.ids.reserve(n.ids.size());
    d.r.ids.reserve(n.ids.size());
    for (uint32_t id : n.ids) {
        unsigned m = classify(triBox[id], axis, plane);
        if (m & 1u) { d.l.ids.push_back(id); grow(bl, triBox[id]); }
        if (m & 2u) { d.r.ids.push_back(id); grow(br, triBox[id]); }
    }
    intersect_with(d.l.box, bl);
    intersect_with(d.r.box, br);
}

void decide(const HostScene& scene, const std::vector<Box>& triBox, const BuildNode& n, const Params& p,
            Decision& d) {
    const size_t count = n.ids.size();
    if (count > p.median_above) {
        float ext[3] = {n.box.hi[0] - n.box.lo[0], n.box.hi[1] - n.box.lo[1], n.box.hi[2] - n.box.lo[2]};
        int axis = 0;
        if (ext[1] > ext[axis]) axis = 1;
        if (ext[2] > ext[axis]) axis = 2;
        apply_split(triBox, n, axis, 0.5f * (n.box.hi[axis] + n.box.lo[axis]), d);
        return;
    }
    // Small node: gather the node's triangle boxes once, then score every candidate plane.
    Box local[64];
    for (size_t i = 0; i < count; ++i) local[i] = triBox[n.ids[i]];
    const float A0 = extent_area(n.box);
    const float leafCost = static_cast<float>(count);
    float best = FLT_MAX, bestPlane = 0.0f;
    int bestAxis = -1;
    float cand[192];
    for (int axis = 0; axis < 3; ++axis) {
        size_t nc = 0;
        for (size_t i = 0; i < count; ++i)
            for (int k = 0; k < 3; ++k) cand[nc++] = scene.vertex(n.ids[i], k)[axis];
        std::sort(cand, cand + nc);
        nc = static_cast<size_t>(std::unique(cand, cand + nc) - cand);
        for (size_t c = 0; c < nc; ++c) {
            const float plane = cand[c];
            if (plane < n.box.lo[axis] || plane > n.box.hi[axis]) continue;
            uint32_t nl = 0, nr = 0;
            Box bl = empty_box(), br = empty_box();
            for (size_t i = 0; i < count; ++i) {
                unsigned m = classify(local[i], axis, plane);
                if (m & 1u) { ++nl; grow(bl, local[i]); }
                if (m & 2u) { ++nr; grow(br, local[i]); }
            }
            Box cl = n.box, cr = n.box;
            cl.hi[axis] = plane;
            cr.lo[axis] = plane;
            intersect_with(cl, bl);
            intersect_with(cr, br);
            // An empty side makes its area +inf and inf * 0 = NaN, which never compares below.
            float cost = (extent_area(cl) * nl + extent_area(cr) * nr) / A0 + p.traversal_cost;
            if (cost < best) { best = cost; bestAxis = axis; bestPlane = plane; }
        }
    }
    if (bestAxis >= 0 && best < leafCost) apply_split(triBox, n, bestAxis, bestPlane, d);
}

}  // namespace

void build_sah_tree(const HostScene& scene, HostTree& out);  // kd_build_sah.cpp

void build_kd_tree(const HostScene& scene, int mode, HostTree& out) {
    if (mode == 1) { build_sah_tree(scene, out); return; }
    auto t0 = std::chrono::steady_clock::now();
    Params params;
    const size_t nT = scene.num_triangles();
    std::vector<Box> triBox(nT);
#pragma omp parallel for schedule(static) if (nT >= 65536)
    for (int64_t t = 0; t < static_cast<int64_t>(nT); ++t) {
        Box b = empty_box();
        for (int k = 0; k < 3; ++k) {
            const float* v = scene.vertex(static_cast<uint32_t>(t), k);
            for (int i = 0; i < 3; ++i) { b.lo[i] = std::min(b.lo[i], v[i]); b.hi[i] = std::max(b.hi[i], v[i]); }
        }
        triBox[static_cast<size_t>(t)] = b;
    }

    std::vector<BuildNode> nodes(1);
    nodes[0].ids.resize(nT);
    nodes[0].box = empty_box();
    for (size_t t = 0; t < nT; ++t) { nodes[0].ids[t] = static_cast<uint32_t>(t); grow(nodes[0].box, triBox[t]); }

    out = HostTree();
    size_t levelBegin = 0;
    for (uint32_t depth = 0; levelBegin < nodes.size(); ++depth) {
        const size_t levelEnd = nodes.size();
        out.max_depth = depth;
        if (depth >= params.depth_cap) break;
        std::vector<Decision> decisions(levelEnd - levelBegin);
#pragma omp parallel for schedule(dynamic, 8) if (levelEnd - levelBegin >= 64)
        for (int64_t k = 0; k < static_cast<int64_t>(levelEnd - levelBegin); ++k)
            decide(scene, triBox, nodes[levelBegin + static_cast<size_t>(k)], params, decisions[static_cast<size_t>(k)]);
        for (size_t k = 0; k < decisions.size(); ++k) {
            Decision& d = decisions[k];
            if (d.axis < 0) continue;
            BuildNode& n = nodes[levelBegin + k];
            n.axis = d.axis;
            n.plane = d.plane;
            n.child = static_cast<uint32_t>(nodes.size());
            std::vector<uint32_t>().swap(n.ids);
            nodes.push_back(std::move(d.l));
            nodes.push_back(std::move(d.r));
        }
        levelBegin = levelEnd;
    }

    const size_t nN = nodes.size();
    out.left.resize(nN); out.aabb.resize(6 * nN); out.split_axis.resize(nN); out.split_value.resize(nN);
    out.leaf_offset.resize(nN); out.leaf_count.resize(nN);
    for (size_t k = 0; k < nN; ++k) {
        const BuildNode& n = nodes[k];
        for (int i = 0; i < 3; ++i) { out.aabb[6 * k + i] = n.box.lo[i]; out.aabb[6 * k + 3 + i] = n.box.hi[i]; }
        out.leaf_offset[k] = static_cast<uint32_t>(out.leaf_ids.size());
        if (n.axis >= 0) {
            out.left[k] = n.child;
            out.split_axis[k] = static_cast<uint32_t>(n.axis + 1);
            out.split_value[k] = n.plane;
            out.leaf_count[k] = 0;
        } else {
            out.left[k] = 0xFFFFFFFFu;
            out.split_axis[k] = 0;
            out.split_value[k] = 0.0f;
            out.leaf_count[k] = static_cast<uint32_t>(n.ids.size());
            out.leaf_ids.insert(out.leaf_ids.end(), n.ids.begin(), n.ids.end());
            ++out.num_leaves;
        }
    }
    out.build_seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}

}  // namespace mcpt
