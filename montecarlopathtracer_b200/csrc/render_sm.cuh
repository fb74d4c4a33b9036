// Path-sampling kernel, persistent-lane variant (the render path).
//
// ncu on the first kernel (render.cuh: one whole ray per loop iteration) showed 4.9 of 32 lanes
// active per issued instruction (profiles/r1_tuning.md): incoherent bounce rays need very different numbers
// of node steps and leaf visits, and at every trace call the warp waited for its slowest lane.
// Here a lane keeps its traversal state (node, interval, stack, closest hit) in registers ACROSS
// loop iterations and the warp alternates two phases:
//   TRAVERSE  every lane with a ray in flight advances it by one leaf visit (descend to the next
//             leaf, test its triangles, pop); after each visit the warp votes (ballot + popc) and
//             leaves the phase as soon as `shadeThreshold` lanes have finished their rays;
//   SHADE     those lanes consume their hit (Russian roulette, emitter test, BSDF sample, optional
//             shadow ray, or the next camera sample) and start a new ray; the others keep theirs.
// A finished lane therefore waits for at most a few leaf visits of its neighbours instead of for
// their whole traversal, and shading runs with at least `shadeThreshold` lanes.  Same rays, same
// arithmetic, same random streams as render_kernel: the images are bit-identical.
#pragma once
#include "render.cuh"

namespace mcpt {

enum : uint32_t { ST_INNER = 0u, ST_LEAF = 1u, ST_SHADE = 2u, ST_DEAD = 3u };

// VAR (tuning knobs, profiles/r1_tuning.md):
//   bit 0  leaf triangles are read as consecutive records from leaf_tris (no id -> record indirection);
//   bit 1  stackless traversal over neighbour links (kd_ropes.cpp): a ray starts in the leaf that contains its
//          origin (hint: the leaf where the previous ray of the path ended) and walks leaf to leaf;
//   bit 2  tri_test<true> (division-free rejections, intersect.cuh);
//   bit 4  camera-ray queue: the next sample's primary ray is generated ahead of time, for many lanes at once,
//          and parked in shared memory (in-place generation runs with ~2.6 of 32 lanes);
//   bit 5  div3_rn: the three IEEE quotients of a surviving triangle share one reciprocal and one range guard;
//   bit 6  (with bit 1) at most `tri_cap` triangle tests per lane and round: a lane in a bigger leaf stays in it
//          (ST_LEAF) and goes on next round, so the round's triangle loop is not as long as the warp's biggest leaf.
//   bit 7  (with bit 1) a bounce or shadow ray starts its walk in the leaf of the hit it leaves, 0.01 BEHIND its origin
//          (the hit point lies in that leaf's cell): no clip against the scene box, no containment test of the hint, no
//          point location from the root when the 0.01 offset has left the cell -- the walk crosses those cells itself;
//   bit 8  tri_test<.., LEAN>: inside div3_guard's range the `> 0` comparisons of the three quotients are implied.
//   bit 9  (with bits 1, 6) PACK: ballot/prefix compaction of the round's triangle tests.  Lanes sit in leaves of 0..4 pending
//          triangles; the per-lane loop runs as long as the fullest leaf (3.9 iterations at K2) with 2.8 tests per lane on
//          average.  Here the round's (lane, triangle) pairs are numbered by a warp prefix sum, pair q is tested by lane
//          q mod 32 (the owner's ray arrives by shuffle), ceil(P / 32) iterations with every lane busy (2.0 at K2), and a
//          64-bit shared-memory atomicMin on (t bits, q) hands each owner its closest accepted hit -- the same hit the
//          sequential loop keeps (smallest t, first in leaf order on a tie).
template <class RNG, bool NEE, bool MOMENTS, int MIN_BLOCKS, int VAR, int BLOCK = 256>
__global__ void __launch_bounds__(BLOCK, MIN_BLOCKS) render_kernel_sm(const RenderParams p, const int shadeThreshold) {
    constexpr bool LEAFREC = (VAR & 1) != 0, ROPES = (VAR & 2) != 0, FASTREJ = (VAR & 4) != 0, STATS = (VAR & 8) != 0, CAMQ = (VAR & 16) != 0,
                   DIV3 = (VAR & 32) != 0, TRICAP = ROPES && (VAR & 64) != 0, STARTHINT = ROPES && (VAR & 128) != 0,
                   LEAN = (VAR & 256) != 0, PACK = TRICAP && (VAR & 512) != 0;
    __shared__ float camq[CAMQ ? 7 * BLOCK : 1];  // CAMQ: parked primary ray {o, d, rng state} per thread
    __shared__ unsigned long long packKey[PACK ? BLOCK : 1];     // PACK: per owner lane, min over its pairs of (t bits << 32 | pair index)
    __shared__ unsigned char packOwner[PACK ? 4 * BLOCK : 1];    // PACK: per warp, pair index -> owner lane | (k << 5); <= 4 pairs per lane
    bool camReady = false;
    uint32_t st[14] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};  // STATS: per-warp round statistics (tuning builds)
    uint32_t px, py;
    thread_pixel<BLOCK>(px, py);
    const bool inside = px < static_cast<uint32_t>(p.width) && py < static_cast<uint32_t>(p.height);
    const uint32_t pixel = py * static_cast<uint32_t>(p.width) + px;
    const int chunk = blockIdx.z;
    const int base = p.sample_count / p.chunks, rem = p.sample_count % p.chunks;
    const int local0 = chunk * base + min(chunk, rem);
    const int localN = base + (chunk < rem ? 1 : 0);
    int s = local0;
    const int sEnd = local0 + localN;

    // ---- per-lane state -------------------------------------------------------------------
    RNG rng;
    float ox = 0, oy = 0, oz = 0, dx = 0, dy = 0, dz = 1;  // the ray being traced
    float ix = 0, iy = 0, iz = 1;                           // 1 / d
    uint32_t nearBits = 0;                                  // bit a: d[a] < 0 -> the right child is nearer
    float tmin = 0, tmax = 0, best = 10000.0f, hbeta = 0, hgamma = 0;
    uint32_t htri = 0xFFFFFFFFu, node = 0, li = 0, lend = 0;
    uint32_t leaf = 0, hintLeaf = 0xFFFFFFFFu, camLeaf = 0xFFFFFFFFu;              // ROPES: current leaf; leaf where the path's last hit was found
    int top = 0;
    uint2 stack[MCPT_KD_STACK];
    F3 thr = f3(1, 1, 1), cur = f3(0, 0, 0), acc = f3(0, 0, 0), acc2 = f3(0, 0, 0);
    F3 nextOrg = f3(0, 0, 0), nextDir = f3(0, 0, 1), shVal = f3(0, 0, 0);  // NEE: path ray parked behind the shadow ray
    uint32_t shTri = 0, bounce = 0, nrays = 0;
    bool shadow = false, lastWasNee = false;
    uint32_t state = ST_DEAD;

    // Prepares traversal of the ray (ox..dz): reciprocal, child order bits, clip against the root box.
    auto begin_ray = [&]() {
        // 1/d steers the walk only (which cells, when to stop); hits are computed from o and d.  The stackless walk needs
        // no more than the hardware reciprocal (<= 1 ulp, 3 instructions instead of ~30 for three IEEE reciprocals);
        // the stack traversal keeps the correctly rounded one.
        if (ROPES) {
            asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(ix) : "f"(dx));
            asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(iy) : "f"(dy));
            asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(iz) : "f"(dz));
        } else {
            ix = __frcp_rn(dx); iy = __frcp_rn(dy); iz = __frcp_rn(dz);
        }
        nearBits = (ix < 0.0f ? 1u : 0u) | (iy < 0.0f ? 2u : 0u) | (iz < 0.0f ? 4u : 0u);
        if (STARTHINT && !(bounce == 0u && !shadow) && hintLeaf != 0xFFFFFFFFu) {
            // The ray leaves a hit at o - 0.01 d, inside the cell of `hintLeaf` (the walk that found the hit stopped there).
            // Starting the walk at that point costs a leaf visit or two when the offset has left the cell; triangles
            // behind the origin fail `t > 0` as they do in the origin's own leaf.
            best = 10000.0f; htri = 0xFFFFFFFFu; hbeta = 0; hgamma = 0;
            leaf = hintLeaf; tmin = -0.0101f; state = ST_INNER;
            return;
        }
        tmin = 0.0f; tmax = 10000.0f;
        float tn = (p.sc.bmin[0] - ox) * ix, tf = (p.sc.bmax[0] - ox) * ix;
        if (ix < 0.0f) { const float sw = tn; tn = tf; tf = sw; }
        tmin = fmaxf(tmin, tn); tmax = fminf(tmax, tf);  // fmaxf/fminf drop the NaN of 0 * inf
        tn = (p.sc.bmin[1] - oy) * iy; tf = (p.sc.bmax[1] - oy) * iy;
        if (iy < 0.0f) { const float sw = tn; tn = tf; tf = sw; }
        tmin = fmaxf(tmin, tn); tmax = fminf(tmax, tf);
        tn = (p.sc.bmin[2] - oz) * iz; tf = (p.sc.bmax[2] - oz) * iz;
        if (iz < 0.0f) { const float sw = tn; tn = tf; tf = sw; }
        tmin = fmaxf(tmin, tn); tmax = fminf(tmax, tf);
        tmin *= (1.0f - 4e-7f); tmax *= (1.0f + 4e-7f);
        best = 10000.0f; htri = 0xFFFFFFFFu; hbeta = 0; hgamma = 0;
        node = 0; top = 0;
        state = (tmin <= tmax) ? ST_INNER : ST_SHADE;  // a ray that misses the scene box is shaded as a miss
        if (ROPES && state == ST_INNER) {
            // tmin doubles as the distance already covered.  The walk starts in the leaf that contains the
            // entry point q: the hinted leaf when q lies in its cell (a bounce ray starts where the previous
            // ray ended; a camera ray enters where this pixel's previous sample entered), else located from
            // the root.
            const bool cam = bounce == 0u && !shadow;
            const uint32_t hint = cam ? camLeaf : hintLeaf;
            float qx = ox, qy = oy, qz = oz;
            if (tmin > 0.0f) {
                qx = fminf(fmaxf(ox + tmin * dx, p.sc.bmin[0]), p.sc.bmax[0]);
                qy = fminf(fmaxf(oy + tmin * dy, p.sc.bmin[1]), p.sc.bmax[1]);
                qz = fminf(fmaxf(oz + tmin * dz, p.sc.bmin[2]), p.sc.bmax[2]);
            }
            bool located = false;
            if (hint != 0xFFFFFFFFu) {
                const float4 h0 = ldg4(p.sc.leaf_blob + MCPT_LEAF_OFF(hint)), h1 = ldg4(p.sc.leaf_blob + MCPT_LEAF_OFF(hint) + 1);
                located = qx >= h0.x && qx <= h1.x && qy >= h0.y && qy <= h1.y && qz >= h0.z && qz <= h1.z;
            }
            if (STATS) { st[3] += 1; st[4] += located ? 1u : 0u; }
            leaf = located ? hint : rope_locate(p.sc, 0u, qx, qy, qz, nearBits);
            if (cam) camLeaf = leaf;
        }
    };
    auto start_sample = [&]() {
        rng.start(pixel, p.seed, static_cast<uint32_t>(p.sample_begin + s), p.frame_seeds ? p.frame_seeds[s] : 0u);
        float u1 = 0.5f, u2 = 0.5f;
        if (p.jitter) { u1 = rng.next(); u2 = rng.next(); }
        F3 o, d;
        camera_ray<!MCPT_FAST_SHADE>(p.cam, p.width, p.height, px, py, u1, u2, o, d);
        ox = o.x; oy = o.y; oz = o.z; dx = d.x; dy = d.y; dz = d.z;
        thr = f3(1, 1, 1); cur = f3(0, 0, 0);
        bounce = 0; lastWasNee = false; shadow = false;
    };
    // CAMQ: generate sample s+1's primary ray for every lane that has none parked, once enough lanes need one.
    auto refill_camera_queue = [&]() {
        const bool need = !camReady && inside && state != ST_DEAD && s + 1 < sEnd;
        if (__popc(__ballot_sync(0xFFFFFFFFu, need)) < p.cam_threshold) return;
        if (need) {
            RNG r;
            r.start(pixel, p.seed, static_cast<uint32_t>(p.sample_begin + s + 1), p.frame_seeds ? p.frame_seeds[s + 1] : 0u);
            float u1 = 0.5f, u2 = 0.5f;
            if (p.jitter) { u1 = r.next(); u2 = r.next(); }
            F3 o, d;
            camera_ray<!MCPT_FAST_SHADE>(p.cam, p.width, p.height, px, py, u1, u2, o, d);
            float* q = camq + threadIdx.x;
            q[0] = o.x; q[BLOCK] = o.y; q[2 * BLOCK] = o.z; q[3 * BLOCK] = d.x; q[4 * BLOCK] = d.y; q[5 * BLOCK] = d.z;
            q[6 * BLOCK] = __uint_as_float(r.save());
            camReady = true;
        }
    };
    auto take_parked_sample = [&]() {
        const float* q = camq + threadIdx.x;
        ox = q[0]; oy = q[BLOCK]; oz = q[2 * BLOCK]; dx = q[3 * BLOCK]; dy = q[4 * BLOCK]; dz = q[5 * BLOCK];
        rng.resume(pixel, p.seed, static_cast<uint32_t>(p.sample_begin + s), __float_as_uint(q[6 * BLOCK]));
        thr = f3(1, 1, 1); cur = f3(0, 0, 0);
        bounce = 0; lastWasNee = false; shadow = false;
        camReady = false;
    };
    // After a leaf: stop if the closest hit lies before the end of this cell, else continue with the
    // next cell along the ray (cells are contiguous: its tmin is this cell's tmax).
    auto pop_or_finish = [&]() {
        if (best < tmax || top == 0) {
            state = ST_SHADE;
        } else {
            --top;
            node = stack[top].x;
            tmin = tmax;
            tmax = __uint_as_float(stack[top].y);
            state = ST_INNER;
        }
    };

    // From `node` down to the next leaf the ray enters within [tmin, tmax]; far children that the ray
    // also crosses are pushed with the interval end they inherit.  The node step is predicated (no
    // divergent branch): 28 SASS instructions.  Leaf triangles [li, lend) pending on return.
    uint32_t myNodes = 0, myTris = 0;
    auto descend = [&]() {
        uint2 n = __ldg(p.sc.kd_nodes + node);
        while ((n.y & 3u) != 3u) {
            if (STATS) ++myNodes;
            const uint32_t axis = n.y & 3u;
            float oa = oz, ia = iz;
            if (axis == 0u) { oa = ox; ia = ix; }
            if (axis == 1u) { oa = oy; ia = iy; }
            const float tp = (__uint_as_float(n.x) - oa) * ia;
            const uint32_t nb = (nearBits >> axis) & 1u, left = n.y >> 2;
            const uint32_t nearChild = left + nb, farChild = left + (nb ^ 1u);
            const bool nearOnly = tp > tmax, farOnly = tp < tmin;
            if (!(nearOnly || farOnly)) {
                stack[top] = make_uint2(farChild, __float_as_uint(tmax));
                ++top;
                tmax = tp;
            }
            node = (farOnly && !nearOnly) ? farChild : nearChild;
            n = __ldg(p.sc.kd_nodes + node);
        }
        li = n.x; lend = n.x + (n.y >> 2);
    };
    auto test_leaf = [&](uint32_t stop) {
        for (; li < stop; li += ROPES ? 3u : 1u) {
            float t, be, ga;
            if (STATS) ++myTris;
            if (ROPES) {
                // li walks the leaf's block in float4 units; results written in place, inside the test's (rare) accept branch
                tri_test<FASTREJ, DIV3, LEAN>(p.sc.leaf_blob + li, ox, oy, oz, dx, dy, dz, best, best, hbeta, hgamma, &htri);
            } else if (LEAFREC) {
                tri_test<FASTREJ, DIV3>(p.sc.leaf_tris + 3 * li, ox, oy, oz, dx, dy, dz, best, best, hbeta, hgamma, &htri);
            } else {
                const uint32_t id = __ldg(p.sc.leaf_ids + li);
                if (tri_test<FASTREJ, DIV3>(p.sc.tri_isect + 3 * id, ox, oy, oz, dx, dy, dz, best, t, be, ga)) {
                    best = t; htri = id; hbeta = be; hgamma = ga;
                }
            }
        }
    };

    if (inside && s < sEnd) {
        start_sample(); begin_ray();
    }

    while (true) {
        {
            // ================= SHADE: consume the hit, produce the next ray ====================
            if (CAMQ) refill_camera_queue();
            if (STATS) { st[7] += 1; st[8] += __popc(__ballot_sync(0xFFFFFFFFu, state == ST_SHADE)); }
            if (state == ST_SHADE) {
                bool done = false, haveRay = false;
                // Without NEE a path's only contribution is the emitter it ends on: `cur` is zero until the shading
                // event that ends the path, so it need not live in a register between events.
                if (!NEE) cur = f3(0, 0, 0);
                if (NEE && shadow) {
                    if (htri == shTri) cur = cur + shVal;  // the sampled emitter point is visible
                    shadow = false;
                    ox = nextOrg.x; oy = nextOrg.y; oz = nextOrg.z; dx = nextDir.x; dy = nextDir.y; dz = nextDir.z;
                    haveRay = true;
                } else if (htri == 0xFFFFFFFFu) {
                    done = true;  // rtx.hlsl:370
                } else {
                    if (ROPES) hintLeaf = leaf;  // the walk stopped in the cell that contains the hit point
                    rng.set_slot(bounce + 1u);
                    rng.fill();  // once for the warp: roulette and every BSDF sampler below draw from this block
                    if (bounce >= p.kill_at) {
                        done = true;  // :313-316
                    } else if (bounce >= p.depth) {
                        const float illum = fmaxf(fmaxf(thr.x, thr.y), thr.z);  // :317-328
                        if (illum > rng.next()) thr = MCPT_FAST_SHADE ? thr * shade_rcp(illum) : thr / illum; else done = true;
                    }
                    if (!done) {
                        const float4 s0 = ldg4(p.sc.tri_shade + 3 * htri);
                        const uint32_t mat = __float_as_uint(s0.w);
                        const float4 mKa = ldg4(p.sc.mats + 3 * mat);
                        if (mKa.x > 0 || mKa.y > 0 || mKa.z > 0) {  // :329-334
                            if (!(NEE && lastWasNee)) cur = cur + thr * f3(mKa.x, mKa.y, mKa.z) * p.emission_scale;
                            done = true;
                        } else if (bounce >= p.gather_at) {
                            done = true;  // CUDA twin: the last intersect only gathers emission (CUTracer.cu:162-175)
                        } else {
                            const float4 s1 = ldg4(p.sc.tri_shade + 3 * htri + 1), s2 = ldg4(p.sc.tri_shade + 3 * htri + 2);
                            const float4 mKd = ldg4(p.sc.mats + 3 * mat + 1);
                            F3 normal = f3(s0.x, s0.y, s0.z) * (1.0f - hbeta - hgamma) + f3(s1.x, s1.y, s1.z) * hbeta +
                                        f3(s2.x, s2.y, s2.z) * hgamma;  // :336-340
                            normal = normalize(normal);
                            const F3 hp = f3(ox + best * dx, oy + best * dy, oz + best * dz);  // :197-199
                            F3 dir = f3(dx, dy, dz);
                            lastWasNee = false;
                            if (mKd.w > 0) {  // Tr > 0, :342-346
                                const float4 mKs = ldg4(p.sc.mats + 3 * mat + 2);
                                dir = sample_fresnel(rng, normal, dir, mKd.w, mKs.w, p.twin != 0u);  // twin: Utils.hpp:97-137
                                if (p.twin) thr = thr * f3(mKd.x, mKd.y, mKd.z);  // CUTracer.cu:131-133
                            } else if (mKa.w > 1) {  // Ns > 1, :348-352
                                const float4 mKs = ldg4(p.sc.mats + 3 * mat + 2);
                                dir = sample_phong(rng, normal, dir, p.twin ? static_cast<float>(static_cast<uint32_t>(mKa.w)) : mKa.w);  // Utils.hpp:72
                                thr = thr * f3(mKs.x, mKs.y, mKs.z);
                            } else {  // :354-365
                                const bool flip = dot(dir, normal) > 0;
                                const F3 lobe = sample_hemi(rng, normal);  // one copy for both signs
                                dir = flip ? -lobe : lobe;
                                thr = thr * f3(mKd.x, mKd.y, mKd.z);
                                if (NEE && p.sc.num_emitters > 0 && bounce + 1u < p.kill_at) {
                                    // what the reference would collect at bounce+1, gathered by area sampling
                                    // of the emitters (DESIGN.md "next-event estimation")
                                    rng.set_slot(0x80000000u | (bounce + 1u));
                                    const float u0 = rng.next(), ub = rng.next(), uc = rng.next();
                                    uint32_t e = 0;
                                    while (e + 1 < p.sc.num_emitters && u0 > ldg4(p.sc.emitters + 3 * e + 1).w) ++e;
                                    const float4 ea = ldg4(p.sc.emitters + 3 * e), eu = ldg4(p.sc.emitters + 3 * e + 1),
                                                 ev = ldg4(p.sc.emitters + 3 * e + 2);
                                    const float su = shade_sqrt(ub);
                                    const float b1 = 1.0f - su, b2 = uc * su;
                                    const F3 y = f3(ea.x + b1 * eu.x + b2 * ev.x, ea.y + b1 * eu.y + b2 * ev.y, ea.z + b1 * eu.z + b2 * ev.z);
                                    F3 w = y - hp;
                                    const float r2 = dot(w, w);
                                    w = w * shade_rsqrt(r2);
                                    const F3 ng = f3(eu.y * ev.z - eu.z * ev.y, eu.z * ev.x - eu.x * ev.z, eu.x * ev.y - eu.y * ev.x);
                                    const float cosy = fabsf(dot(ng, w)) * shade_rcp(2.0f * ea.w);
                                    const float cosx = dot(flip ? -normal : normal, w);
                                    if (cosx > 0 && cosy > 0 && r2 > 0) {
                                        shTri = __float_as_uint(ev.w);
                                        const uint32_t lmat = __float_as_uint(ldg4(p.sc.tri_shade + 3 * shTri).w);
                                        const float4 lKa = ldg4(p.sc.mats + 3 * lmat);
                                        const float g = cosx * cosy * p.sc.emitter_area * shade_rcp(MCPT_PI * r2);
                                        shVal = thr * f3(lKa.x, lKa.y, lKa.z) * (g * p.emission_scale);
                                        shadow = true;
                                        ++nrays;
                                        nextOrg = hp + dir * 0.01f;  // the path ray waits behind the shadow ray
                                        nextDir = dir;
                                        const F3 so = hp + w * 0.01f;
                                        ox = so.x; oy = so.y; oz = so.z; dx = w.x; dy = w.y; dz = w.z;
                                    }
                                    lastWasNee = true;
                                }
                            }
                            if (!(NEE && shadow)) {
                                const F3 no = hp + dir * 0.01f;  // :366
                                ox = no.x; oy = no.y; oz = no.z; dx = dir.x; dy = dir.y; dz = dir.z;
                            }
                            ++bounce;
                            haveRay = true;
                        }
                    }
                }
                if (STATS && done) st[9] += 1;
                if (done) {
                    nrays += bounce + 1u;  // the camera ray and one ray per scattering event (shadow rays are counted where they start)
                    acc = acc + cur;
                    if (MOMENTS) acc2 = acc2 + cur * cur;
                    ++s;
                    if (s < sEnd) {
                        if (CAMQ && camReady) take_parked_sample(); else start_sample();
                        haveRay = true;
                    } else {
                        state = ST_DEAD;
                    }
                }
                if (haveRay) begin_ray();
            }
        }
        if (__ballot_sync(0xFFFFFFFFu, state != ST_DEAD) == 0u) break;
        // ================= TRAVERSE: one leaf visit per round, until enough lanes wait ==========
        while (true) {
            if (STATS) { myNodes = 0; myTris = 0; st[1] += 1; st[2] += __popc(__ballot_sync(0xFFFFFFFFu, state == ST_INNER)); }
            if (ROPES) {
                // The leaf's handle says where its block is and (up to 6) how many triangles it holds: the triangle records
                // are requested without waiting for the header, which is read once, after the tests, for the exit face.
                const uint32_t off = MCPT_LEAF_OFF(leaf);
                if (PACK) {
                    const uint32_t lane = threadIdx.x & 31u;
                    if (state == ST_INNER) {
                        uint32_t cnt = MCPT_LEAF_CNT(leaf);
                        if (cnt == 7u) cnt = __float_as_uint(ldg4(p.sc.leaf_blob + off + 1).w);
                        li = off + 4u; lend = cnt;  // PACK: lend counts the triangles of this leaf still to test
                    }
                    const bool trav = state == ST_INNER || state == ST_LEAF;
                    const uint32_t n = trav ? min(lend, min(static_cast<uint32_t>(p.tri_cap), 4u)) : 0u;
                    if (STATS) myTris = n;
                    uint32_t incl = n;  // inclusive prefix sum of n over the warp
#pragma unroll
                    for (int d = 1; d < 32; d <<= 1) {
                        const uint32_t v = __shfl_up_sync(0xFFFFFFFFu, incl, d);
                        if (lane >= static_cast<uint32_t>(d)) incl += v;
                    }
                    const uint32_t pbase = incl - n, P = __shfl_sync(0xFFFFFFFFu, incl, 31);
                    unsigned char* own = packOwner + 4u * (threadIdx.x & ~31u);
#pragma unroll
                    for (uint32_t j = 0; j < 4u; ++j)
                        if (j < n) own[pbase + j] = static_cast<unsigned char>(lane | (j << 5));
                    packKey[threadIdx.x] = ~0ull;
                    __syncwarp();
                    for (uint32_t q0 = 0; q0 < P; q0 += 32u) {
                        const uint32_t q = q0 + lane;
                        const bool act = q < P;
                        const uint32_t e = act ? own[q] : lane;
                        const uint32_t o = e & 31u;
                        const float rox = __shfl_sync(0xFFFFFFFFu, ox, o), roy = __shfl_sync(0xFFFFFFFFu, oy, o), roz = __shfl_sync(0xFFFFFFFFu, oz, o);
                        const float rdx = __shfl_sync(0xFFFFFFFFu, dx, o), rdy = __shfl_sync(0xFFFFFFFFu, dy, o), rdz = __shfl_sync(0xFFFFFFFFu, dz, o);
                        const float rbest = __shfl_sync(0xFFFFFFFFu, best, o);
                        const uint32_t rli = __shfl_sync(0xFFFFFFFFu, li, o);
                        float t = 0, be = 0, ga = 0;
                        uint32_t id = 0;
                        if (act && tri_test<FASTREJ, DIV3, LEAN>(p.sc.leaf_blob + rli + 3u * (e >> 5), rox, roy, roz, rdx, rdy, rdz, rbest, t, be, ga, &id))
                            atomicMin(packKey + (threadIdx.x & ~31u) + o, (static_cast<unsigned long long>(__float_as_uint(t)) << 32) | q);
                        __syncwarp();
                        const unsigned long long key = packKey[threadIdx.x];
                        const uint32_t kq = static_cast<uint32_t>(key);
                        const bool mine = key != ~0ull && kq - q0 < 32u;  // set in THIS iteration: pull the record from the lane that found it
                        const uint32_t src = mine ? (kq & 31u) : lane;
                        const float nbeta = __shfl_sync(0xFFFFFFFFu, be, src), ngamma = __shfl_sync(0xFFFFFFFFu, ga, src);
                        const uint32_t nid = __shfl_sync(0xFFFFFFFFu, id, src);
                        if (mine) { best = __uint_as_float(static_cast<uint32_t>(key >> 32)); hbeta = nbeta; hgamma = ngamma; htri = nid; }
                    }
                    if (trav) { li += 3u * n; lend -= n; state = lend > 0u ? ST_LEAF : ST_INNER; }
                } else if (state == ST_INNER || (TRICAP && state == ST_LEAF)) {
                    if (!TRICAP || state == ST_INNER) {
                        uint32_t cnt = MCPT_LEAF_CNT(leaf);
                        if (cnt == 7u) cnt = __float_as_uint(ldg4(p.sc.leaf_blob + off + 1).w);
                        li = off + 4u; lend = li + 3u * cnt;
                    }
                    test_leaf(TRICAP ? min(lend, li + 3u * static_cast<uint32_t>(p.tri_cap)) : lend);
                    if (TRICAP) state = li < lend ? ST_LEAF : ST_INNER;
                }
                if (state == ST_INNER) {
                    const float4 h0 = ldg4(p.sc.leaf_blob + off), h1 = ldg4(p.sc.leaf_blob + off + 1);
                    const LeafExit e = leaf_exit(p.sc, off, h0, h1, ox, oy, oz, ix, iy, iz, nearBits);
                    // No hop counter here (it costs a register the kernel does not have; intersect_ropes keeps one):
                    // the walk terminates structurally -- GIVEN consistent tables.  That is a precondition this kernel depends on:
                    // kd_ropes.cpp builds the tables, check_ropes (mcpt_check_stackless_layout) verifies them for every scene
                    // and both builders in the CPU suite, and mcpt_set_tuning("verify_layout", 1) runs the same check on
                    // every geometry install (0.9 s at 1 M triangles, off by default).  A warp-uniform stall counter as a
                    // run-time net was measured: -5 % at K2 (profiles/r2_tuning.md), not shipped.  The face taken on each axis is fixed by the ray's sign bits,
                    // a rope's subtree lies wholly beyond that face, so the cell's bound on the exit axis moves
                    // strictly in the sign direction every hop whatever the float values (NaN included) and the
                    // number of hops is bounded by the number of split planes.
                    if (best <= e.t || e.rope == 0xFFFFFFFFu) {
                        state = ST_SHADE;
                    } else {
                        tmin = fmaxf(tmin, e.t);
                        if (e.rope & MCPT_ROPE_IS_LEAF) {
                            leaf = e.rope & ~MCPT_ROPE_IS_LEAF;  // one leaf behind the whole face: nothing to locate
                        } else {
                            float qx = ox + tmin * dx, qy = oy + tmin * dy, qz = oz + tmin * dz;
                            if (e.axis == 0u) qx = e.plane;
                            if (e.axis == 1u) qy = e.plane;
                            if (e.axis == 2u) qz = e.plane;
                            uint32_t steps = 0;
                            leaf = rope_locate(p.sc, e.rope, qx, qy, qz, nearBits, STATS ? &steps : nullptr);
                            if (STATS) { myNodes = steps; st[10] += 1; }
                        }
                    }
                }
            } else if (state == ST_INNER) {
                descend();
                test_leaf(lend);
                pop_or_finish();
            }
            if (STATS && ROPES) {  // point location behind exit faces: steps summed over lanes, and the round's longest walk
                const uint32_t mx = __reduce_max_sync(0xFFFFFFFFu, myNodes);
                st[11] += myNodes; st[12] += mx > 0u ? 1u : 0u; st[13] += mx;
            }
            if (STATS) {
                st[3] += __reduce_add_sync(0xFFFFFFFFu, ROPES ? 0u : myNodes); st[4] += __reduce_max_sync(0xFFFFFFFFu, ROPES ? 0u : myNodes);
                st[5] += __reduce_add_sync(0xFFFFFFFFu, myTris); st[6] += __reduce_max_sync(0xFFFFFFFFu, myTris);
            }
            if (__popc(__ballot_sync(0xFFFFFFFFu, state != ST_INNER && state != ST_LEAF)) >= shadeThreshold) break;
        }
    }

    if (inside) {
        const size_t o = (static_cast<size_t>(chunk) * p.height + py) * p.width + px;
        p.sum[o] = make_float4(acc.x, acc.y, acc.z, static_cast<float>(localN));
        if (MOMENTS) p.sumsq[o] = make_float4(acc2.x, acc2.y, acc2.z, static_cast<float>(localN));
    }
    if (STATS) {
        st[9] = __reduce_add_sync(0xFFFFFFFFu, st[9]);
        if (ROPES) {
            st[3] = __reduce_add_sync(0xFFFFFFFFu, st[3]); st[4] = __reduce_add_sync(0xFFFFFFFFu, st[4]);
            st[10] = __reduce_add_sync(0xFFFFFFFFu, st[10]); st[11] = __reduce_add_sync(0xFFFFFFFFu, st[11]);
        }
        if ((threadIdx.x & 31u) == 0u)
            for (int k = 1; k < 14; ++k) atomicAdd(p.ray_counter + k, static_cast<unsigned long long>(st[k]));
    }
    for (int off = 16; off > 0; off >>= 1) nrays += __shfl_xor_sync(0xFFFFFFFFu, nrays, off);
    if ((threadIdx.x & 31u) == 0u && nrays) atomicAdd(p.ray_counter, static_cast<unsigned long long>(nrays));
}

}  // namespace mcpt
