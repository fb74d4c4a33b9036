// TEST INFRASTRUCTURE -- CPU oracle, host half.  NOT product code.
//
// Restatement of the reference's KD-tree builder and flatten:
//   MCRT/QuinEngine/Utils/KDTree.hpp:58-287      (BFS build, median / exhaustive SAH)
//   MCRT/QuinEngine/Utils/Structure.hpp:91-183   (QuinAABB / QuinTriangle arithmetic)
//   MCRT/QuinEngine/RTX/ShaderResource.hpp:128-179 (BFS flatten, children adjacent)
// It is pinned node-for-node (topology, boxes, split planes, leaf id lists) against the
// reference's own code compiled in oracle/_ref (tests/test_oracle_pins.py::test_restated_builder_equals_reference_builder).
// Compile with -ffp-contract=off: every float expression below keeps the reference's
// operand order, and must round the way the reference's x86 build does.
#include <algorithm>
#include <cfloat>
#include <cstddef>

#include "oracle_internal.h"

namespace oracle {
namespace {

struct Box {
    float lo[3], hi[3];
};

// Structure.hpp:94 -- default QuinAABB is the empty box.
Box empty_box() {
    Box b;
    for (int i = 0; i < 3; ++i) { b.lo[i] = FLT_MAX; b.hi[i] = -FLT_MAX; }
    return b;
}

// Structure.hpp:113-121 (Add) and :127-132 (operator+= adds both corners).
void unite(Box& a, const Box& b) {
    for (int i = 0; i < 3; ++i) {
        a.lo[i] = std::min(a.lo[i], b.lo[i]);
        a.hi[i] = std::max(a.hi[i], b.lo[i]);
    }
    for (int i = 0; i < 3; ++i) {
        a.lo[i] = std::min(a.lo[i], b.hi[i]);
        a.hi[i] = std::max(a.hi[i], b.hi[i]);
    }
}

// Structure.hpp:133-141 (operator*= is the intersection).
void clip(Box& a, const Box& b) {
    for (int i = 0; i < 3; ++i) {
        a.lo[i] = std::max(a.lo[i], b.lo[i]);
        a.hi[i] = std::min(a.hi[i], b.hi[i]);
    }
}

// KDTree.hpp:177-178 / :229-232 -- "area" is xy + yz + zx of the extent.
float half_area(const Box& b) {
    float e0 = b.hi[0] - b.lo[0], e1 = b.hi[1] - b.lo[1], e2 = b.hi[2] - b.lo[2];
    return e0 * e1 + e1 * e2 + e2 * e0;
}

enum : int { kLeft = 1, kRight = 2 };

// The straddle rule used at KDTree.hpp:131-151, :200-223 and :253-273:
// flat on the plane -> left only; min < plane -> left; max > plane -> right.
inline int sides(const Box& tb, int axis, float plane) {
    if (tb.lo[axis] == tb.hi[axis] && tb.lo[axis] == plane) return kLeft;
    int m = 0;
    if (tb.lo[axis] < plane) m |= kLeft;
    if (tb.hi[axis] > plane) m |= kRight;
    return m;
}

struct Work {
    Box box;
    std::vector<uint32_t> ids;  // ascending (the reference keeps a std::set<UINT>)
    uint32_t depth = 0;
    uint32_t axis = 0;  // 0 = leaf, 1..3 = x,y,z  (KDAxis)
    float plane = 0.0f;
    uint32_t child = 0;  // index of the left child; right = child + 1
};

struct Candidate {
    int axis;
    float value;
};

// KDTree.hpp:389-401 (GetNodeAABB): union of the node's triangle boxes, empty if none.
Box bounds_of(const std::vector<Box>& triBox, const std::vector<uint32_t>& ids) {
    Box b = empty_box();
    for (uint32_t id : ids) unite(b, triBox[id]);
    return b;
}

}  // namespace

void build_tree(const Scene& scene, Tree& out) {
    out.clear();
    const size_t nT = scene.tris.size();

    // KDTree.hpp:60-75 -- positions only, triangle id = running index in table order.
    std::vector<Box> triBox(nT);
    for (size_t t = 0; t < nT; ++t) {
        Box b = empty_box();
        for (int k = 0; k < 3; ++k) {
            const float* p = &scene.verts[3 * static_cast<size_t>(scene.tris[t].v[k])];
            for (int i = 0; i < 3; ++i) {
                b.lo[i] = std::min(b.lo[i], p[i]);
                b.hi[i] = std::max(b.hi[i], p[i]);
            }
        }
        triBox[t] = b;
    }

    std::vector<Work> nodes;
    nodes.reserve(2 * nT + 16);
    nodes.emplace_back();
    nodes[0].ids.resize(nT);
    for (size_t t = 0; t < nT; ++t) nodes[0].ids[t] = static_cast<uint32_t>(t);
    nodes[0].box = bounds_of(triBox, nodes[0].ids);  // :89

    // One helper for all three split sites (:123-161, :245-283).
    auto split_node = [&](size_t at, int axis, float plane) {
        Work l, r;
        l.box = nodes[at].box; l.box.hi[axis] = plane;
        r.box = nodes[at].box; r.box.lo[axis] = plane;
        l.depth = r.depth = nodes[at].depth + 1;
        for (uint32_t id : nodes[at].ids) {
            int m = sides(triBox[id], axis, plane);
            if (m & kLeft) l.ids.push_back(id);
            if (m & kRight) r.ids.push_back(id);
        }
        clip(l.box, bounds_of(triBox, l.ids));
        clip(r.box, bounds_of(triBox, r.ids));
        nodes[at].axis = static_cast<uint32_t>(axis + 1);
        nodes[at].plane = plane;
        nodes[at].child = static_cast<uint32_t>(nodes.size());
        nodes[at].ids.clear();
        nodes[at].ids.shrink_to_fit();
        nodes.push_back(std::move(l));
        nodes.push_back(std::move(r));
    };

    std::vector<Candidate> cands;
    // The work list of :91-286 is FIFO, children are appended left then right, so the
    // creation index IS the BFS index the flatten assigns.
    for (size_t at = 0; at < nodes.size(); ++at) {
        if (nodes[at].depth > out.maxDepth) out.maxDepth = nodes[at].depth;
        if (nodes[at].depth >= 32) continue;  // :103-106, no size check
        const size_t n = nodes[at].ids.size();
        const Box box = nodes[at].box;
        if (n > 64) {
            // :108-122 spatial median on the longest axis (first wins ties)
            float ext[3] = {box.hi[0] - box.lo[0], box.hi[1] - box.lo[1], box.hi[2] - box.lo[2]};
            float longest = ext[0];
            int axis = 0;
            for (int i = 1; i < 3; ++i) {
                if (ext[i] > longest) { longest = ext[i]; axis = i; }
            }
            float plane = 0.5f * (box.hi[axis] + box.lo[axis]);
            split_node(at, axis, plane);
            continue;
        }
        // :166-176 every vertex coordinate, ordered by (axis, value); equal keys collapse
        // onto the first one inserted (std::set semantics).
        cands.clear();
        for (int axis = 0; axis < 3; ++axis) {
            for (uint32_t id : nodes[at].ids) {
                for (int k = 0; k < 3; ++k) {
                    cands.push_back({axis, scene.verts[3 * static_cast<size_t>(scene.tris[id].v[k]) + axis]});
                }
            }
        }
        auto less = [](const Candidate& a, const Candidate& b) {
            if (a.axis != b.axis) return a.axis < b.axis;
            return a.value < b.value;
        };
        std::stable_sort(cands.begin(), cands.end(), less);
        cands.erase(std::unique(cands.begin(), cands.end(),
                                [&](const Candidate& a, const Candidate& b) { return !less(a, b) && !less(b, a); }),
                    cands.end());

        const float Cts = 0.0f;            // :166
        const float A0 = half_area(box);   // :177-178
        const float SAH0 = static_cast<float>(n);  // :179
        float best = FLT_MAX;
        Candidate bestCand{-1, 0.0f};
        for (const Candidate& c : cands) {
            if (c.value < box.lo[c.axis] || c.value > box.hi[c.axis]) continue;  // :186-189
            uint32_t numL = 0, numR = 0;
            Box bl = box, br = box;
            bl.hi[c.axis] = c.value;
            br.lo[c.axis] = c.value;
            Box tl = empty_box(), tr = empty_box();
            for (uint32_t id : nodes[at].ids) {
                int m = sides(triBox[id], c.axis, c.value);
                if (m & kLeft) { ++numL; unite(tl, triBox[id]); }
                if (m & kRight) { ++numR; unite(tr, triBox[id]); }
            }
            clip(bl, tl);
            clip(br, tr);
            float AL = half_area(bl), AR = half_area(br);
            // :233 -- an empty side gives inf * 0 = NaN here, which never compares less.
            float sah = (AL * numL + AR * numR) / A0 + Cts;
            if (sah < best) { best = sah; bestCand = c; }
        }
        if (best < SAH0) split_node(at, bestCand.axis, bestCand.value);  // :240
    }

    // ShaderResource.hpp:128-179.  Creation order == BFS pop order, children adjacent.
    const size_t nN = nodes.size();
    out.left.resize(nN); out.right.resize(nN); out.splitAxis.resize(nN);
    out.splitValue.resize(nN); out.leafOffset.resize(nN); out.leafCount.resize(nN);
    out.aabb.resize(6 * nN);
    for (size_t k = 0; k < nN; ++k) {
        const Work& w = nodes[k];
        for (int i = 0; i < 3; ++i) { out.aabb[6 * k + i] = w.box.lo[i]; out.aabb[6 * k + 3 + i] = w.box.hi[i]; }
        out.leafOffset[k] = static_cast<uint32_t>(out.leafIds.size());
        if (w.axis != 0) {
            out.left[k] = w.child;
            out.right[k] = w.child + 1;
            out.splitAxis[k] = w.axis;
            out.splitValue[k] = w.plane;
            out.leafCount[k] = 0;
        } else {
            out.left[k] = out.right[k] = 0xFFFFFFFFu;
            out.splitAxis[k] = 0;
            out.splitValue[k] = 0.0f;
            out.leafCount[k] = static_cast<uint32_t>(w.ids.size());
            out.leafIds.insert(out.leafIds.end(), w.ids.begin(), w.ids.end());
            ++out.numLeaves;
        }
    }
}

}  // namespace oracle
