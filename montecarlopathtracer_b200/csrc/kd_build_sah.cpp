// MCPT_BUILD_SAH: this library's own KD-tree builder (product code, no reference counterpart).
//
// The reference tree (kd_build.cpp, KDTree.hpp:58-287) bins a triangle by its WHOLE bounding box:
// a sphere triangle is referenced by every cell its box overlaps, median planes are used above 64
// triangles and Cts = 0 keeps splitting until duplication explodes.  It stays the parity
// anchor.  For rendering, the same hits can be found with far fewer node steps and triangle tests:
//   * surface-area heuristic on every level, sweep over sorted bound events (Wald & Havran 2006);
//   * "perfect splits": a straddling triangle is CLIPPED to each child cell (Sutherland-Hodgman in
//     double, bounds rounded outwards), so it is referenced only where it really is;
//   * empty-space bonus, traversal/intersection cost ratio tuned for the strict Cramer test;
//   * depth cap 31 so the traversal stack (MCPT_KD_STACK = 32) can never overflow.
// Output uses the same flattened layout as the reference tree (BFS order, children adjacent, ids
// ascending inside a leaf); the per-node box is the cell.
#include <omp.h>

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cfloat>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "host_scene.h"

namespace mcpt {
namespace {

struct Box {
    float lo[3], hi[3];
};

struct Ref {
    uint32_t id;
    Box b;  // bounds of the triangle clipped to the current cell
};

struct Node {
    Box cell;
    int axis = -1;
    float plane = 0.0f;
    int left = -1, right = -1;
    std::vector<uint32_t> ids;
};

const float kTraversalCost = 1.0f;
// the strict Cramer test + its loads against one plane step; defaults from profiles/r1_tuning.md.  Read once per
// build from tuning() (mcpt_set_tuning), never from the environment.
float kIntersectCost = 1.0f;
float kEmptyBonus = 1.0f;

inline float surface(const Box& b) {
    const float x = b.hi[0] - b.lo[0], y = b.hi[1] - b.lo[1], z = b.hi[2] - b.lo[2];
    return 2.0f * (x * y + y * z + z * x);
}

// Bounds of triangle `id` clipped to `cell` (conservative: rounded outwards, then clamped to the cell).
bool clipped_bounds(const HostScene& s, uint32_t id, const Box& cell, Box& out) {
    double poly[16][3], tmp[16][3];
    int n = 3;
    for (int k = 0; k < 3; ++k) {
        const float* v = s.vertex(id, k);
        for (int i = 0; i < 3; ++i) poly[k][i] = v[i];
    }
    for (int axis = 0; axis < 3 && n > 0; ++axis) {
        for (int side = 0; side < 2 && n > 0; ++side) {
            const double plane = side == 0 ? cell.lo[axis] : cell.hi[axis];
            const double sign = side == 0 ? 1.0 : -1.0;  // keep sign * (x - plane) >= 0
            bool cuts = false;  // most planes of a cell do not cut the polygon: it passes through unchanged
            for (int k = 0; k < n && !cuts; ++k) cuts = sign * (poly[k][axis] - plane) < 0;
            if (!cuts) continue;
            int m = 0;
            for (int k = 0; k < n; ++k) {
                const double* a = poly[k];
                const double* b = poly[(k + 1) % n];
                const double da = sign * (a[axis] - plane), db = sign * (b[axis] - plane);
                if (da >= 0) { for (int i = 0; i < 3; ++i) tmp[m][i] = a[i]; ++m; }
                if ((da > 0 && db < 0) || (da < 0 && db > 0)) {
                    const double t = da / (da - db);
                    for (int i = 0; i < 3; ++i) tmp[m][i] = a[i] + t * (b[i] - a[i]);
                    tmp[m][axis] = plane;
                    ++m;
                }
            }
            n = m;
            for (int k = 0; k < n; ++k) for (int i = 0; i < 3; ++i) poly[k][i] = tmp[k][i];
        }
    }
    if (n == 0) return false;
    for (int i = 0; i < 3; ++i) {
        double lo = poly[0][i], hi = poly[0][i];
        for (int k = 1; k < n; ++k) { lo = std::min(lo, poly[k][i]); hi = std::max(hi, poly[k][i]); }
        float flo = static_cast<float>(lo), fhi = static_cast<float>(hi);
        if (flo > lo) flo = std::nextafter(flo, -FLT_MAX);
        if (fhi < hi) fhi = std::nextafter(fhi, FLT_MAX);
        out.lo[i] = std::max(flo, cell.lo[i]);
        out.hi[i] = std::min(fhi, cell.hi[i]);
        if (out.lo[i] > out.hi[i]) return false;
    }
    return true;
}

// A bound event as one sortable integer: (order-preserving bits of the position << 2) | type with
// type 0 = end, 1 = planar, 2 = start -- plain integer compares sort by (position, type).  -0.0 is
// folded into +0.0 first so that equal positions have equal keys.
using Event = uint64_t;
inline Event make_event(float pos, unsigned type) {
    pos += 0.0f;
    uint32_t u;
    std::memcpy(&u, &pos, 4);
    u = (u & 0x80000000u) ? ~u : (u | 0x80000000u);
    return (static_cast<uint64_t>(u) << 2) | type;
}
inline float event_pos(Event e) {
    uint32_t u = static_cast<uint32_t>(e >> 2);
    u = (u & 0x80000000u) ? (u & 0x7FFFFFFFu) : ~u;
    float f;
    std::memcpy(&f, &u, 4);
    return f;
}

struct Split {
    float cost = FLT_MAX;
    int axis = -1;
    float plane = 0.0f;
    bool planarLeft = true;
};

// Events are 34-bit keys: three 12-bit passes of an LSD radix sort replace std::sort above a few thousand
// events (1 M triangles: 2 M events per axis at the root, 170 ms -> 25 ms per axis on one core; the sorts were
// two thirds of the build).  Same order as std::sort: the keys are plain integers.
void sort_events(std::vector<Event>& ev, std::vector<Event>& tmp, const int base = 0) {
    const size_t n = ev.size();
    if (n < 64) {
        std::sort(ev.begin(), ev.end());
        return;
    }
    tmp.resize(n);
    Event* src = ev.data();
    Event* dst = tmp.data();
    if (n < 4096) {
        // Small and middle nodes (most of the calls): 7-bit digits, counters on the stack, and a pass in which all keys
        // share the digit is skipped -- the events of a small cell agree in the upper bits of their positions.
        for (int pass = 0; pass < 5; ++pass) {
            const int shift = base + 7 * pass;
            uint32_t count[129] = {0};
            for (size_t i = 0; i < n; ++i) ++count[((src[i] >> shift) & 0x7Fu) + 1u];
            if (count[((src[0] >> shift) & 0x7Fu) + 1u] == n) continue;
            for (int k = 0; k < 128; ++k) count[k + 1] += count[k];
            for (size_t i = 0; i < n; ++i) dst[count[(src[i] >> shift) & 0x7Fu]++] = src[i];
            std::swap(src, dst);
        }
        if (src != ev.data()) ev.swap(tmp);
        return;
    }
    for (int pass = 0; pass < 3; ++pass) {
        const int shift = base + 12 * pass;
        uint32_t count[4097] = {0};
        for (size_t i = 0; i < n; ++i) ++count[((src[i] >> shift) & 0xFFFu) + 1u];
        for (int k = 0; k < 4096; ++k) count[k + 1] += count[k];
        for (size_t i = 0; i < n; ++i) dst[count[(src[i] >> shift) & 0xFFFu]++] = src[i];
        std::swap(src, dst);
    }
    // three passes: the sorted data ends in tmp
    ev.swap(tmp);
}

// Sweep over bound events sorted by key (Wald & Havran 2006, section 4.2).  SHIFT exposes the 34-bit key: 0 for plain
// events, kIdxBits for events that carry their reference's index in the low bits.
constexpr int kIdxBits = 30;
template <int SHIFT>
Split sweep_events(const Box& cell, int axis, int n, const uint64_t* ev, size_t m) {
    Split best;
    if (!(cell.hi[axis] > cell.lo[axis])) return best;
    const float invSA = 1.0f / surface(cell);
    int nl = 0, nr = n;
    for (size_t i = 0; i < m;) {
        const uint64_t group = ev[i] >> (SHIFT + 2);
        const float pos = event_pos(ev[i] >> SHIFT);
        int ends = 0, planar = 0, starts = 0;
        while (i < m && (ev[i] >> (SHIFT + 2)) == group && ((ev[i] >> SHIFT) & 3u) == 0u) { ++ends; ++i; }
        while (i < m && (ev[i] >> (SHIFT + 2)) == group && ((ev[i] >> SHIFT) & 3u) == 1u) { ++planar; ++i; }
        while (i < m && (ev[i] >> (SHIFT + 2)) == group && ((ev[i] >> SHIFT) & 3u) == 2u) { ++starts; ++i; }
        nr -= planar + ends;
        if (pos > cell.lo[axis] && pos < cell.hi[axis]) {
            Box l = cell, r = cell;
            l.hi[axis] = pos;
            r.lo[axis] = pos;
            const float pl = surface(l) * invSA, pr = surface(r) * invSA;
            for (int side = 0; side < 2; ++side) {
                const int cl = nl + (side == 0 ? planar : 0), cr = nr + (side == 1 ? planar : 0);
                float c = kTraversalCost + kIntersectCost * (pl * cl + pr * cr);
                if (cl == 0 || cr == 0) c *= kEmptyBonus;
                if (c < best.cost) { best.cost = c; best.axis = axis; best.plane = pos; best.planarLeft = side == 0; }
                if (planar == 0) break;
            }
        }
        nl += starts + planar;
    }
    return best;
}

// Best plane of one axis: events of the node's references, sorted, swept.
Split axis_split(const Box& cell, const std::vector<Ref>& refs, int axis, std::vector<Event>& ev, std::vector<Event>& tmp) {
    if (!(cell.hi[axis] > cell.lo[axis])) return Split();
    const int n = static_cast<int>(refs.size());
    ev.clear();
    ev.reserve(2 * refs.size());
    for (const Ref& r : refs) {
        if (r.b.lo[axis] == r.b.hi[axis]) {
            ev.push_back(make_event(r.b.lo[axis], 1u));
        } else {
            ev.push_back(make_event(r.b.lo[axis], 2u));
            ev.push_back(make_event(r.b.hi[axis], 0u));
        }
    }
    sort_events(ev, tmp);
    return sweep_events<0>(cell, axis, n, ev.data(), ev.size());
}

// Best plane over the three axes; ties go to the lower axis, then the lower position (the order a sequential
// sweep meets them in).  Big nodes run the axes as three tasks.
Split find_split(const Box& cell, const std::vector<Ref>& refs) {
    Split per[3];
    if (refs.size() >= (1u << 16)) {
        std::vector<Event> evs[3], tmps[3];
#pragma omp taskgroup
        {
            for (int axis = 0; axis < 3; ++axis) {
#pragma omp task shared(per, evs, tmps, cell, refs) firstprivate(axis)
                per[axis] = axis_split(cell, refs, axis, evs[axis], tmps[axis]);
            }
        }
    } else {
        // per-thread scratch: no scheduling point between here and the end of the three sweeps, and 1.8 M nodes would
        // otherwise allocate and free two event arrays each
        static thread_local std::vector<Event> ev, tmp;
        for (int axis = 0; axis < 3; ++axis) per[axis] = axis_split(cell, refs, axis, ev, tmp);
        if (ev.capacity() > (1u << 20)) { std::vector<Event>().swap(ev); std::vector<Event>().swap(tmp); }
    }
    Split best;
    for (int axis = 0; axis < 3; ++axis)
        if (per[axis].cost < best.cost) best = per[axis];
    return best;
}

void build_node(const HostScene& s, std::vector<Node>& nodes, int me, std::vector<Ref>& refs, int depth, int maxDepth) {
    const size_t n = refs.size();
    Split sp;
    if (n > 0 && depth < maxDepth) sp = find_split(nodes[me].cell, refs);
    if (sp.axis < 0 || sp.cost >= kIntersectCost * static_cast<float>(n)) {
        nodes[me].ids.reserve(n);
        for (const Ref& r : refs) nodes[me].ids.push_back(r.id);
        std::sort(nodes[me].ids.begin(), nodes[me].ids.end());
        return;
    }
    const int a = sp.axis;
    Box lc = nodes[me].cell, rc = nodes[me].cell;
    lc.hi[a] = sp.plane;
    rc.lo[a] = sp.plane;
    std::vector<Ref> lrefs, rrefs;
    for (const Ref& r : refs) {
        if (r.b.lo[a] == r.b.hi[a] && r.b.lo[a] == sp.plane) {
            (sp.planarLeft ? lrefs : rrefs).push_back(r);
        } else if (r.b.hi[a] <= sp.plane) {
            lrefs.push_back(r);
        } else if (r.b.lo[a] >= sp.plane) {
            rrefs.push_back(r);
        } else {
            Ref c{r.id, {}};
            if (clipped_bounds(s, r.id, lc, c.b)) lrefs.push_back(c);
            if (clipped_bounds(s, r.id, rc, c.b)) rrefs.push_back(c);
        }
    }
    std::vector<Ref>().swap(refs);
    int li = -1, ri = -1;
#pragma omp critical(mcpt_sah_nodes)
    if (nodes.size() + 2 <= nodes.capacity()) {  // never reallocate: other tasks read their own nodes concurrently
        li = static_cast<int>(nodes.size());
        ri = li + 1;
        nodes.emplace_back();
        nodes.emplace_back();
        nodes[li].cell = lc;
        nodes[ri].cell = rc;
        nodes[me].axis = a;
        nodes[me].plane = sp.plane;
        nodes[me].left = li;
        nodes[me].right = ri;
    }
    if (li < 0) {  // node budget exhausted (pathological duplication): stop here
        for (const Ref& r : lrefs) nodes[me].ids.push_back(r.id);
        for (const Ref& r : rrefs) nodes[me].ids.push_back(r.id);
        std::sort(nodes[me].ids.begin(), nodes[me].ids.end());
        nodes[me].ids.erase(std::unique(nodes[me].ids.begin(), nodes[me].ids.end()), nodes[me].ids.end());
        return;
    }
    const bool spawn = lrefs.size() + rrefs.size() > 4096;
    // lrefs is shared on purpose (it would be copied into the task otherwise): this frame waits at the taskwait below
#pragma omp task shared(s, nodes, lrefs) firstprivate(li, depth, maxDepth) if (spawn)
    {
        std::vector<Ref> mine;
        mine.swap(lrefs);
        build_node(s, nodes, li, mine, depth + 1, maxDepth);
    }
    build_node(s, nodes, ri, rrefs, depth + 1, maxDepth);
#pragma omp taskwait
}

// ---- sort once, keep sorted ------------------------------------------------------------------------------------
// build_node sorts the bound events of a node afresh on every level: half of the build (6.0 of 11.5 G cycles on the
// 1 M-triangle scene).  Above kResortBelow references a node instead INHERITS its three event lists already sorted
// (Wald & Havran 2006, section 4.3): an event carries the index of its reference in the node's array; after the
// split the events of references that go to one side only are appended to that child's list in order, the few
// references the plane cuts get new events from their clipped bounds, sorted by themselves and merged in.  Same
// references, same bounds, same event keys as build_node would produce, so the same sweep result and the same tree
// (tests/test_host_library.py::test_sah_tree_is_frozen, tools/host_build_bench.cpp digest).
using SEvent = uint64_t;  // (34-bit event key << kIdxBits) | index of the reference in the node's array
constexpr size_t kResortBelow = 48;
constexpr uint32_t kGone = 0xFFFFFFFFu;

inline void push_ref_events(std::vector<SEvent>* ev, const Ref& r, uint32_t idx) {
    for (int axis = 0; axis < 3; ++axis) {
        if (r.b.lo[axis] == r.b.hi[axis]) {
            ev[axis].push_back((make_event(r.b.lo[axis], 1u) << kIdxBits) | idx);
        } else {
            ev[axis].push_back((make_event(r.b.lo[axis], 2u) << kIdxBits) | idx);
            ev[axis].push_back((make_event(r.b.hi[axis], 0u) << kIdxBits) | idx);
        }
    }
}

void build_node_sorted(const HostScene& s, std::vector<Node>& nodes, int me, std::vector<Ref>& refs, std::vector<SEvent>* ev, int depth,
                       int maxDepth) {
    const size_t n = refs.size();
    if (n < kResortBelow) {
        for (int axis = 0; axis < 3; ++axis) std::vector<SEvent>().swap(ev[axis]);
        build_node(s, nodes, me, refs, depth, maxDepth);
        return;
    }
    const bool big = n >= (1u << 16);
    Split sp;
    if (depth < maxDepth) {
        Split per[3];
        const Box cell = nodes[me].cell;
        if (big) {
#pragma omp taskgroup
            for (int axis = 0; axis < 3; ++axis) {
#pragma omp task shared(per, ev) firstprivate(axis, cell, n)
                per[axis] = sweep_events<kIdxBits>(cell, axis, static_cast<int>(n), ev[axis].data(), ev[axis].size());
            }
        } else {
            for (int axis = 0; axis < 3; ++axis) per[axis] = sweep_events<kIdxBits>(cell, axis, static_cast<int>(n), ev[axis].data(), ev[axis].size());
        }
        for (int axis = 0; axis < 3; ++axis)
            if (per[axis].cost < sp.cost) sp = per[axis];
    }
    if (sp.axis < 0 || sp.cost >= kIntersectCost * static_cast<float>(n)) {
        nodes[me].ids.reserve(n);
        for (const Ref& r : refs) nodes[me].ids.push_back(r.id);
        std::sort(nodes[me].ids.begin(), nodes[me].ids.end());
        return;
    }
    const int a = sp.axis;
    Box lc = nodes[me].cell, rc = nodes[me].cell;
    lc.hi[a] = sp.plane;
    rc.lo[a] = sp.plane;
    // where each reference goes: its index in the left / right child (kGone = not there); cut references get new bounds
    std::vector<Ref> lrefs, rrefs;
    std::vector<uint32_t> li(n, kGone), ri(n, kGone);
    std::vector<SEvent> newL[3], newR[3];
    for (size_t i = 0; i < n; ++i) {
        const Ref& r = refs[i];
        if (r.b.lo[a] == r.b.hi[a] && r.b.lo[a] == sp.plane) {
            if (sp.planarLeft) { li[i] = static_cast<uint32_t>(lrefs.size()); lrefs.push_back(r); }
            else { ri[i] = static_cast<uint32_t>(rrefs.size()); rrefs.push_back(r); }
        } else if (r.b.hi[a] <= sp.plane) {
            li[i] = static_cast<uint32_t>(lrefs.size()); lrefs.push_back(r);
        } else if (r.b.lo[a] >= sp.plane) {
            ri[i] = static_cast<uint32_t>(rrefs.size()); rrefs.push_back(r);
        } else {
            Ref c{r.id, {}};
            if (clipped_bounds(s, r.id, lc, c.b)) { push_ref_events(newL, c, static_cast<uint32_t>(lrefs.size())); lrefs.push_back(c); }
            if (clipped_bounds(s, r.id, rc, c.b)) { push_ref_events(newR, c, static_cast<uint32_t>(rrefs.size())); rrefs.push_back(c); }
        }
    }
    int lidx = -1, ridx = -1;
#pragma omp critical(mcpt_sah_nodes)
    if (nodes.size() + 2 <= nodes.capacity()) {
        lidx = static_cast<int>(nodes.size());
        ridx = lidx + 1;
        nodes.emplace_back();
        nodes.emplace_back();
        nodes[lidx].cell = lc;
        nodes[ridx].cell = rc;
        nodes[me].axis = a;
        nodes[me].plane = sp.plane;
        nodes[me].left = lidx;
        nodes[me].right = ridx;
    }
    if (lidx < 0) {  // node budget exhausted (pathological duplication): stop here, as build_node does
        for (const Ref& r : lrefs) nodes[me].ids.push_back(r.id);
        for (const Ref& r : rrefs) nodes[me].ids.push_back(r.id);
        std::sort(nodes[me].ids.begin(), nodes[me].ids.end());
        nodes[me].ids.erase(std::unique(nodes[me].ids.begin(), nodes[me].ids.end()), nodes[me].ids.end());
        return;
    }
    // hand the sorted lists down: uncut references keep their events (new index), cut ones were regenerated above
    std::vector<SEvent> L[3], R[3];
    auto hand_down = [&](int axis) {
        const SEvent idxMask = (SEvent(1) << kIdxBits) - 1;
        std::vector<SEvent> l, r;
        l.reserve(std::min(ev[axis].size(), 2 * lrefs.size()));  // at most two events per reference and axis
        r.reserve(std::min(ev[axis].size(), 2 * rrefs.size()));
        for (const SEvent e : ev[axis]) {
            const uint32_t i = static_cast<uint32_t>(e & idxMask);
            if (li[i] != kGone) l.push_back((e & ~idxMask) | li[i]);
            else if (ri[i] != kGone) r.push_back((e & ~idxMask) | ri[i]);
        }
        std::vector<SEvent>().swap(ev[axis]);
        auto byKey = [](SEvent x, SEvent y) { return (x >> kIdxBits) < (y >> kIdxBits); };
        auto merge_in = [&](std::vector<SEvent>& base, std::vector<SEvent>& fresh, std::vector<SEvent>& out) {
            if (fresh.empty()) { out.swap(base); return; }
            std::sort(fresh.begin(), fresh.end());
            out.resize(base.size() + fresh.size());
            std::merge(base.begin(), base.end(), fresh.begin(), fresh.end(), out.begin(), byKey);
        };
        merge_in(l, newL[axis], L[axis]);
        merge_in(r, newR[axis], R[axis]);
    };
    if (big) {
#pragma omp taskgroup
        for (int axis = 0; axis < 3; ++axis) {
#pragma omp task shared(hand_down) firstprivate(axis)
            hand_down(axis);
        }
    } else {
        for (int axis = 0; axis < 3; ++axis) hand_down(axis);
    }
    std::vector<Ref>().swap(refs);
    std::vector<uint32_t>().swap(li);
    std::vector<uint32_t>().swap(ri);
    const bool spawn = lrefs.size() + rrefs.size() > 4096;
#pragma omp task shared(s, nodes, lrefs, L) firstprivate(lidx, depth, maxDepth) if (spawn)
    {
        std::vector<Ref> mine;
        mine.swap(lrefs);
        std::vector<SEvent> mineEv[3];
        for (int axis = 0; axis < 3; ++axis) mineEv[axis].swap(L[axis]);
        build_node_sorted(s, nodes, lidx, mine, mineEv, depth + 1, maxDepth);
    }
    build_node_sorted(s, nodes, ridx, rrefs, R, depth + 1, maxDepth);
#pragma omp taskwait
}

}  // namespace

void build_sah_tree(const HostScene& scene, HostTree& out) {
    kIntersectCost = tuning().sah_ci;
    kEmptyBonus = tuning().sah_eb;
    auto t0 = std::chrono::steady_clock::now();
    const size_t nT = scene.num_triangles();
    std::vector<Ref> refs(nT);
    Box root = {{FLT_MAX, FLT_MAX, FLT_MAX}, {-FLT_MAX, -FLT_MAX, -FLT_MAX}};
    for (size_t t = 0; t < nT; ++t) {
        Box b = {{FLT_MAX, FLT_MAX, FLT_MAX}, {-FLT_MAX, -FLT_MAX, -FLT_MAX}};
        for (int k = 0; k < 3; ++k) {
            const float* v = scene.vertex(static_cast<uint32_t>(t), k);
            for (int i = 0; i < 3; ++i) { b.lo[i] = std::min(b.lo[i], v[i]); b.hi[i] = std::max(b.hi[i], v[i]); }
        }
        refs[t] = {static_cast<uint32_t>(t), b};
        for (int i = 0; i < 3; ++i) { root.lo[i] = std::min(root.lo[i], b.lo[i]); root.hi[i] = std::max(root.hi[i], b.hi[i]); }
    }
    const int maxDepth = std::min(31, static_cast<int>(8.0 + 1.3 * std::log2(static_cast<double>(std::max<size_t>(nT, 1)))));
    // std::vector<Node> would move nodes while tasks hold indices only (never references across the
    // critical section), so growth is safe; reserve to keep it cheap.
    std::vector<Node> nodes;
    nodes.reserve(6 * nT + 1024);
    nodes.emplace_back();
    nodes[0].cell = root;
    // Small scenes are built by the calling thread alone: waking a team of 16-24 threads costs more than the 862-triangle
    // Cornell box takes to build (K1's end-to-end step is 2.7 ms of rendering).
    const bool team = nT > 4096;
    if (tuning().sah_resort) {  // A/B knob: the re-sorting builder on every level
#pragma omp parallel if (team)
#pragma omp single
        build_node(scene, nodes, 0, refs, 0, maxDepth);
    } else {
        std::vector<SEvent> ev[3];
        for (int axis = 0; axis < 3; ++axis) ev[axis].reserve(2 * nT);
        for (size_t t = 0; t < nT; ++t) push_ref_events(ev, refs[t], static_cast<uint32_t>(t));
#pragma omp parallel if (team)
#pragma omp single
        {
#pragma omp taskgroup
            for (int axis = 0; axis < 3; ++axis) {
#pragma omp task shared(ev) firstprivate(axis)
                {
                    std::vector<SEvent> tmp;
                    sort_events(ev[axis], tmp, kIdxBits);
                }
            }
            build_node_sorted(scene, nodes, 0, refs, ev, 0, maxDepth);
        }
    }
    if (tuning().build_timing) std::fprintf(stderr, "sah: recursion %.3f s\n", std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count());

    // breadth-first renumbering, children adjacent (same layout as the reference flatten)
    std::vector<int> order;
    order.reserve(nodes.size());
    std::vector<uint32_t> depthOf;
    order.push_back(0);
    depthOf.push_back(0);
    std::vector<int> newIndex(nodes.size(), -1);
    for (size_t k = 0; k < order.size(); ++k) {
        const Node& n = nodes[static_cast<size_t>(order[k])];
        newIndex[static_cast<size_t>(order[k])] = static_cast<int>(k);
        if (n.axis >= 0) {
            order.push_back(n.left); depthOf.push_back(depthOf[k] + 1);
            order.push_back(n.right); depthOf.push_back(depthOf[k] + 1);
        }
    }
    out = HostTree();
    const size_t nN = order.size();
    out.left.resize(nN); out.aabb.resize(6 * nN); out.split_axis.resize(nN); out.split_value.resize(nN);
    out.leaf_offset.resize(nN); out.leaf_count.resize(nN);
    // children of node k are the next two unassigned BFS slots; a leaf's ids start where the previous leaves' end:
    // two running sums (serial, cheap), then every node is filled independently
    {
        size_t next = 1, idsSoFar = 0;
        for (size_t k = 0; k < nN; ++k) {
            const Node& n = nodes[static_cast<size_t>(order[k])];
            out.leaf_offset[k] = static_cast<uint32_t>(idsSoFar);
            if (n.axis >= 0) { out.left[k] = static_cast<uint32_t>(next); next += 2; }
            else { out.left[k] = 0xFFFFFFFFu; idsSoFar += n.ids.size(); ++out.num_leaves; }
            out.max_depth = std::max(out.max_depth, depthOf[k]);
        }
        out.leaf_ids.resize(idsSoFar);
    }
#pragma omp parallel for schedule(static) if (nN > 65536)
    for (size_t k = 0; k < nN; ++k) {
        const Node& n = nodes[static_cast<size_t>(order[k])];
        for (int i = 0; i < 3; ++i) { out.aabb[6 * k + i] = n.cell.lo[i]; out.aabb[6 * k + 3 + i] = n.cell.hi[i]; }
        if (n.axis >= 0) {
            out.split_axis[k] = static_cast<uint32_t>(n.axis + 1);
            out.split_value[k] = n.plane;
            out.leaf_count[k] = 0;
        } else {
            out.split_axis[k] = 0;
            out.split_value[k] = 0.0f;
            out.leaf_count[k] = static_cast<uint32_t>(n.ids.size());
            std::copy(n.ids.begin(), n.ids.end(), out.leaf_ids.begin() + out.leaf_offset[k]);
        }
    }
    if (tuning().build_timing) std::fprintf(stderr, "sah: recursion + flatten %.3f s\n", std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count());
    out.build_seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}

}  // namespace mcpt
