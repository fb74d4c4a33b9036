// Path-sampling kernel, pooled-lane variant (round 2; emission-on-hit estimator, no moments).
//
// render_kernel_sm ties a path to a lane: a lane whose ray has finished waits until `shadeThreshold` lanes of its warp
// have finished too (a TRAVERSE round runs with 22 of 32 lanes on average) and a SHADE phase runs with the ~16 lanes that
// wait.  Here a warp owns a POOL of 32 * SETS paths and a lane is only an execution slot for whichever path needs one:
//   * a lane whose ray finished writes the hit into the path's record, queues the path for shading and at once takes
//     the next waiting ray from the warp's ray queue -- TRAVERSE rounds run with every lane busy while rays wait;
//   * when a lane finds the ray queue empty the warp shades a batch of up to 32 queued hits, one per lane; a shaded path
//     puts its bounce ray back into the ray queue, or -- its sample finished -- joins the camera queue;
//   * the camera queue is served 16+ paths at a time (the next sample's primary ray: Philox block, camera map, clip
//     against the scene box, entry leaf), again one path per lane.
// The three queues are FIFO rings of path ids private to the warp: no atomics, no block-level synchronisation.
// Where the state lives: the ray a lane picks up (origin, direction, start distance, start leaf: 8 words per path) in
// shared memory; everything only SHADE touches (hit record, throughput, bounce, sample index, camera-leaf hint) in a
// global scratch array that stays in L2 (ld.cg / st.cg, 40 bytes per path, read and written once per ray, one coalesced
// row per word and warp) -- kept out of shared memory because L1 is carved from the same 256 KB and the scene's
// working set (171 KB at K2) must stay in it (with the whole record in shared memory the L1 hit rate fell from 99.5 %
// to 92 % and the kernel lost what the pooling gained: profiles/r2_tuning.md).
// Warps are persistent: the grid is MIN_BLOCKS blocks per SM, every WARP draws (8x4-pixel tile, slice set) items from a
// global counter until none is left, so the scratch is sized by the resident warps, not by the image.
// A path is (pixel, slice): set k of an item renders slice  z * SETS + k  of its 32 pixels, the samples of a path run in
// order and each path accumulates into its own slice entry, so the sums are the same floating-point sums as
// render_kernel_sm's with the same number of slices: images are bit-identical (tests).
// Traversal (stackless walk over neighbour links, <= tri_cap triangle tests per lane and round, bit-exact triangle test)
// and shading are the code of render_sm.cuh / VAR 503.
#pragma once
#include "render.cuh"

#include <type_traits>

namespace mcpt {

enum : uint32_t { PL_WALK = 0u, PL_LEAF = 1u, PL_DONE = 2u, PL_IDLE = 3u };
// words of the global per-path record (structure of arrays over a warp's paths)
enum { PG_T = 0, PG_TRI, PG_BETA, PG_GAMMA, PG_THRX, PG_THRY, PG_THRZ, PG_BOUNCE, PG_SAMPLE, PG_CAMLEAF, PG_RNG, PG_WORDS };
constexpr int kPoolWorkCounter = 15;   // ray_counter[15]: next work item (zeroed with the other counters before the launch)

template <class RNG, int MIN_BLOCKS, int SETS, int BLOCK = 256>
__global__ void __launch_bounds__(BLOCK, MIN_BLOCKS) render_kernel_pool(const RenderParams p) {
    constexpr int WARPS = BLOCK / 32, S = 32 * SETS;
    constexpr bool STATEFUL_RNG = !std::is_same<RNG, RngPhilox>::value;
    enum { W_OX = 0, W_OY, W_OZ, W_DX, W_DY, W_DZ, W_TMIN, W_LEAF /* start leaf; after the walk: the leaf of the hit */, W_WORDS };
    constexpr int PITCH = S + SETS;   // path j sits at j + (j >> 5): the sets start in different banks
    __shared__ float rays[WARPS * W_WORDS * PITCH];
    __shared__ unsigned char rings[WARPS * 3 * S];
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    float* const pw = rays + warp * (W_WORDS * PITCH);
    unsigned char* const rayQ = rings + warp * (3 * S);
    unsigned char* const hitQ = rayQ + S;
    unsigned char* const camQ = hitQ + S;
    float* const gs = p.pool_state + static_cast<size_t>(blockIdx.x * WARPS + warp) * (PG_WORDS * S);
    auto at = [&](int word, uint32_t j) -> float& { return pw[word * PITCH + j + (j >> 5)]; };
    auto atu = [&](int word, uint32_t j) -> uint32_t& { return reinterpret_cast<uint32_t*>(pw)[word * PITCH + j + (j >> 5)]; };
    auto gld = [&](int word, uint32_t j) { return __ldcg(gs + word * S + j); };
    auto gldu = [&](int word, uint32_t j) { return __float_as_uint(__ldcg(gs + word * S + j)); };
    auto gst = [&](int word, uint32_t j, float v) { __stcg(gs + word * S + j, v); };
    auto gstu = [&](int word, uint32_t j, uint32_t v) { __stcg(gs + word * S + j, __uint_as_float(v)); };
    const uint32_t ltMask = (1u << lane) - 1u;
    const int base = p.sample_count / p.chunks, rem = p.sample_count % p.chunks;
    const uint32_t tilesX = (static_cast<uint32_t>(p.width) + 15u) / 16u, tilesY = (static_cast<uint32_t>(p.height) + 15u) / 16u;
    const uint32_t zN = (static_cast<uint32_t>(p.chunks) + SETS - 1u) / SETS;
    const unsigned long long items = static_cast<unsigned long long>(tilesX) * tilesY * zN * WARPS;
    uint32_t nrays = 0;

    // Appends the paths of the lanes with `has` to a ring, in lane order.
    auto push = [&](unsigned char* ring, uint32_t& tail, bool has, uint32_t j) {
        const uint32_t m = __ballot_sync(0xFFFFFFFFu, has);
        if (has) ring[(tail + __popc(m & ltMask)) % S] = static_cast<unsigned char>(j);
        tail += __popc(m);
    };

    while (true) {
        unsigned long long item = 0;
        if (lane == 0u) item = atomicAdd(p.ray_counter + kPoolWorkCounter, 1ull);
        item = __shfl_sync(0xFFFFFFFFu, item, 0);
        if (item >= items) break;
        // item -> (tile, warp of the tile, slice set); the pixel map of thread_pixel()
        const uint32_t tw = static_cast<uint32_t>(item % WARPS);
        uint32_t rest = static_cast<uint32_t>(item / WARPS);
        const uint32_t tx = rest % tilesX; rest /= tilesX;
        const uint32_t ty = rest % tilesY, tz = rest / tilesY;
        auto pixel_of = [&](uint32_t j, uint32_t& qx, uint32_t& qy) {
            const uint32_t l = j & 31u;
            qx = tx * 16u + (tw & 1u) * 8u + (l & 7u);
            qy = ty * static_cast<uint32_t>(BLOCK / 16) + (tw >> 1) * 4u + (l >> 3);
        };
        auto slice_of = [&](uint32_t j, int& chunk, int& sEnd) {   // -> first sample
            chunk = static_cast<int>(tz) * SETS + static_cast<int>(j >> 5);
            const int local0 = chunk * base + min(chunk, rem);
            sEnd = local0 + base + (chunk < rem ? 1 : 0);
            return local0;
        };
        uint32_t rHead = 0, rTail = 0, hHead = 0, hTail = 0, cHead = 0, cTail = 0;   // warp-uniform ring positions (free-running, used mod S)

        // ---- start: every live path waits for its first camera ray ----------------------------------------------------
        {
            uint32_t px, py;
            pixel_of(lane, px, py);
            const bool inside = px < static_cast<uint32_t>(p.width) && py < static_cast<uint32_t>(p.height);
#pragma unroll 1
            for (int k = 0; k < SETS; ++k) {
                const uint32_t j = static_cast<uint32_t>(k) * 32u + lane;
                int chunk, sEnd;
                const int s0 = slice_of(j, chunk, sEnd);
                const bool live = inside && chunk < p.chunks && s0 < sEnd;
                if (inside && chunk < p.chunks)
                    __stcg(p.sum + (static_cast<size_t>(chunk) * p.height + py) * p.width + px, make_float4(0.0f, 0.0f, 0.0f, static_cast<float>(sEnd - s0)));
                if (live) { gstu(PG_SAMPLE, j, static_cast<uint32_t>(s0)); gstu(PG_CAMLEAF, j, 0xFFFFFFFFu); }
                push(camQ, cTail, live, j);
            }
            __syncwarp();
        }

        // ---- per-lane traversal state ---------------------------------------------------------------------------------
        float ox = 0, oy = 0, oz = 0, dx = 0, dy = 0, dz = 1, ix = 0, iy = 0, iz = 1;
        uint32_t nearBits = 0;
        float tmin = 0, best = 10000.0f, hbeta = 0, hgamma = 0;
        uint32_t htri = 0xFFFFFFFFu, li = 0, lend = 0, leaf = 0, path = 0;
        uint32_t state = PL_IDLE;

        while (true) {
            // ---- (A) idle lanes take waiting rays ------------------------------------------------------------------
            {
                const uint32_t m = __ballot_sync(0xFFFFFFFFu, state == PL_IDLE);
                const uint32_t waiting = rTail - rHead;
                if (m != 0u && waiting != 0u) {
                    const uint32_t rank = __popc(m & ltMask);
                    if (state == PL_IDLE && rank < waiting) {
                        path = rayQ[(rHead + rank) % S];
                        ox = at(W_OX, path); oy = at(W_OY, path); oz = at(W_OZ, path); dx = at(W_DX, path); dy = at(W_DY, path); dz = at(W_DZ, path);
                        tmin = at(W_TMIN, path); leaf = atu(W_LEAF, path);
                        asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(ix) : "f"(dx));
                        asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(iy) : "f"(dy));
                        asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(iz) : "f"(dz));
                        nearBits = (ix < 0.0f ? 1u : 0u) | (iy < 0.0f ? 2u : 0u) | (iz < 0.0f ? 4u : 0u);
                        best = 10000.0f; htri = 0xFFFFFFFFu; hbeta = 0; hgamma = 0;
                        state = leaf == 0xFFFFFFFFu ? PL_DONE : PL_WALK;   // a camera ray that misses the scene box ends as a miss
                    }
                    rHead += min(static_cast<uint32_t>(__popc(m)), waiting);
                }
            }
            const uint32_t idleMask = __ballot_sync(0xFFFFFFFFu, state == PL_IDLE);
            if (idleMask != 0u) {   // a lane found the ray queue empty: make rays
                const uint32_t queuedHits = hTail - hHead, queuedCams = cTail - cHead;
                // ---- (B1) camera rays for the paths whose sample finished, a batch at a time ---------------------------
                if (queuedCams >= static_cast<uint32_t>(p.cam_threshold) || (queuedCams != 0u && queuedHits == 0u)) {
                    const uint32_t n = min(queuedCams, 32u);
                    const bool mine = lane < n;
                    uint32_t j = 0;
                    if (mine) {
                        j = camQ[(cHead + lane) % S];
                        uint32_t qx, qy;
                        pixel_of(j, qx, qy);
                        const int s = static_cast<int>(gldu(PG_SAMPLE, j));
                        const uint32_t camLeaf = gldu(PG_CAMLEAF, j);
                        RNG rng;
                        rng.start(qy * static_cast<uint32_t>(p.width) + qx, p.seed, static_cast<uint32_t>(p.sample_begin + s), p.frame_seeds ? p.frame_seeds[s] : 0u);
                        float u1 = 0.5f, u2 = 0.5f;
                        if (p.jitter) { u1 = rng.next(); u2 = rng.next(); }
                        F3 o, d;
                        camera_ray<!MCPT_FAST_SHADE>(p.cam, p.width, p.height, qx, qy, u1, u2, o, d);
                        float jx, jy, jz;
                        asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(jx) : "f"(d.x));
                        asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(jy) : "f"(d.y));
                        asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(jz) : "f"(d.z));
                        float t0 = 0.0f, t1 = 10000.0f;
                        float tn = (p.sc.bmin[0] - o.x) * jx, tf = (p.sc.bmax[0] - o.x) * jx;
                        if (jx < 0.0f) { const float sw = tn; tn = tf; tf = sw; }
                        t0 = fmaxf(t0, tn); t1 = fminf(t1, tf);  // fmaxf/fminf drop the NaN of 0 * inf
                        tn = (p.sc.bmin[1] - o.y) * jy; tf = (p.sc.bmax[1] - o.y) * jy;
                        if (jy < 0.0f) { const float sw = tn; tn = tf; tf = sw; }
                        t0 = fmaxf(t0, tn); t1 = fminf(t1, tf);
                        tn = (p.sc.bmin[2] - o.z) * jz; tf = (p.sc.bmax[2] - o.z) * jz;
                        if (jz < 0.0f) { const float sw = tn; tn = tf; tf = sw; }
                        t0 = fmaxf(t0, tn); t1 = fminf(t1, tf);
                        t0 *= (1.0f - 4e-7f); t1 *= (1.0f + 4e-7f);
                        uint32_t entry = 0xFFFFFFFFu;   // stays so when the ray misses the scene box: the path ends as a miss without a walk
                        if (t0 <= t1) {
                            // the leaf that contains the entry point: where this pixel's previous sample entered when the point lies
                            // in that cell, else located from the root (render_sm.cuh begin_ray)
                            float ex = o.x, ey = o.y, ez = o.z;
                            if (t0 > 0.0f) {
                                ex = fminf(fmaxf(o.x + t0 * d.x, p.sc.bmin[0]), p.sc.bmax[0]);
                                ey = fminf(fmaxf(o.y + t0 * d.y, p.sc.bmin[1]), p.sc.bmax[1]);
                                ez = fminf(fmaxf(o.z + t0 * d.z, p.sc.bmin[2]), p.sc.bmax[2]);
                            }
                            bool located = false;
                            if (camLeaf != 0xFFFFFFFFu) {
                                const float4 h0 = ldg4(p.sc.leaf_blob + MCPT_LEAF_OFF(camLeaf)), h1 = ldg4(p.sc.leaf_blob + MCPT_LEAF_OFF(camLeaf) + 1);
                                located = ex >= h0.x && ex <= h1.x && ey >= h0.y && ey <= h1.y && ez >= h0.z && ez <= h1.z;
                            }
                            entry = located ? camLeaf : rope_locate(p.sc, 0u, ex, ey, ez, 0u);
                            if (!located) gstu(PG_CAMLEAF, j, entry);
                        }
                        at(W_OX, j) = o.x; at(W_OY, j) = o.y; at(W_OZ, j) = o.z; at(W_DX, j) = d.x; at(W_DY, j) = d.y; at(W_DZ, j) = d.z;
                        at(W_TMIN, j) = t0; atu(W_LEAF, j) = entry;
                        gst(PG_THRX, j, 1.0f); gst(PG_THRY, j, 1.0f); gst(PG_THRZ, j, 1.0f);
                        gstu(PG_BOUNCE, j, 0u);
                        if (STATEFUL_RNG) gstu(PG_RNG, j, rng.save());
                    }
                    cHead += n;
                    push(rayQ, rTail, mine, j);
                    __syncwarp();
                    continue;
                }
                // ---- (B2) shade a batch of queued hits ----------------------------------------------------------------
                if (queuedHits != 0u) {
                    const uint32_t n = min(queuedHits, 32u);
                    const bool mine = lane < n;
                    bool haveRay = false, nextSample = false;
                    uint32_t j = 0;
                    if (mine) {
                        j = hitQ[(hHead + lane) % S];
                        const uint32_t tri = gldu(PG_TRI, j);
                        const uint32_t bounce = gldu(PG_BOUNCE, j);
                        const int s = static_cast<int>(gldu(PG_SAMPLE, j));
                        uint32_t qx, qy;
                        pixel_of(j, qx, qy);
                        F3 cur = f3(0, 0, 0);
                        bool done = false;
                        if (tri == 0xFFFFFFFFu) {
                            done = true;  // rtx.hlsl:370
                        } else {
                            F3 thr = f3(gld(PG_THRX, j), gld(PG_THRY, j), gld(PG_THRZ, j));
                            RNG rng;
                            rng.resume(qy * static_cast<uint32_t>(p.width) + qx, p.seed, static_cast<uint32_t>(p.sample_begin + s), STATEFUL_RNG ? gldu(PG_RNG, j) : 0u);
                            rng.set_slot(bounce + 1u);
                            rng.fill();
                            if (bounce >= p.kill_at) {
                                done = true;  // :313-316
                            } else if (bounce >= p.depth) {
                                const float illum = fmaxf(fmaxf(thr.x, thr.y), thr.z);  // :317-328
                                if (illum > rng.next()) thr = MCPT_FAST_SHADE ? thr * shade_rcp(illum) : thr / illum; else done = true;
                            }
                            if (!done) {
                                const float4 s0 = ldg4(p.sc.tri_shade + 3 * tri);
                                const uint32_t mat = __float_as_uint(s0.w);
                                const float4 mKa = ldg4(p.sc.mats + 3 * mat);
                                if (mKa.x > 0 || mKa.y > 0 || mKa.z > 0) {  // :329-334
                                    cur = cur + thr * f3(mKa.x, mKa.y, mKa.z) * p.emission_scale;
                                    done = true;
                                } else if (bounce >= p.gather_at) {
                                    done = true;  // CUDA twin (CUTracer.cu:162-175)
                                } else {
                                    const float4 s1 = ldg4(p.sc.tri_shade + 3 * tri + 1), s2 = ldg4(p.sc.tri_shade + 3 * tri + 2);
                                    const float4 mKd = ldg4(p.sc.mats + 3 * mat + 1);
                                    const float hb = gld(PG_BETA, j), hg = gld(PG_GAMMA, j), t = gld(PG_T, j);
                                    const float rox = at(W_OX, j), roy = at(W_OY, j), roz = at(W_OZ, j), rdx = at(W_DX, j), rdy = at(W_DY, j), rdz = at(W_DZ, j);
                                    F3 normal = f3(s0.x, s0.y, s0.z) * (1.0f - hb - hg) + f3(s1.x, s1.y, s1.z) * hb + f3(s2.x, s2.y, s2.z) * hg;  // :336-340
                                    normal = normalize(normal);
                                    const F3 hp = f3(rox + t * rdx, roy + t * rdy, roz + t * rdz);  // :197-199
                                    F3 dir = f3(rdx, rdy, rdz);
                                    if (mKd.w > 0) {  // Tr > 0, :342-346
                                        const float4 mKs = ldg4(p.sc.mats + 3 * mat + 2);
                                        dir = sample_fresnel(rng, normal, dir, mKd.w, mKs.w, p.twin != 0u);
                                        if (p.twin) thr = thr * f3(mKd.x, mKd.y, mKd.z);
                                    } else if (mKa.w > 1) {  // Ns > 1, :348-352
                                        const float4 mKs = ldg4(p.sc.mats + 3 * mat + 2);
                                        dir = sample_phong(rng, normal, dir, p.twin ? static_cast<float>(static_cast<uint32_t>(mKa.w)) : mKa.w);
                                        thr = thr * f3(mKs.x, mKs.y, mKs.z);
                                    } else {  // :354-365
                                        const bool flip = dot(dir, normal) > 0;
                                        const F3 lobe = sample_hemi(rng, normal);
                                        dir = flip ? -lobe : lobe;
                                        thr = thr * f3(mKd.x, mKd.y, mKd.z);
                                    }
                                    const F3 no = hp + dir * 0.01f;  // :366
                                    at(W_OX, j) = no.x; at(W_OY, j) = no.y; at(W_OZ, j) = no.z; at(W_DX, j) = dir.x; at(W_DY, j) = dir.y; at(W_DZ, j) = dir.z;
                                    at(W_TMIN, j) = -0.0101f;   // the walk starts 0.01 behind the new origin, in the leaf of the hit (W_LEAF as the walk left it)
                                    gst(PG_THRX, j, thr.x); gst(PG_THRY, j, thr.y); gst(PG_THRZ, j, thr.z);
                                    gstu(PG_BOUNCE, j, bounce + 1u);
                                    if (STATEFUL_RNG) gstu(PG_RNG, j, rng.save());
                                    haveRay = true;
                                }
                            }
                        }
                        if (done) {
                            nrays += bounce + 1u;
                            int chunk, sEnd;
                            slice_of(j, chunk, sEnd);
                            if (cur.x != 0.0f || cur.y != 0.0f || cur.z != 0.0f) {   // adding +0 changes nothing (the sums start at +0, radiance is >= 0)
                                float4* const dst = p.sum + (static_cast<size_t>(chunk) * p.height + qy) * p.width + qx;
                                float4 a = __ldcg(dst);   // L2: whichever lane shades this path next reads it back after a __syncwarp
                                a.x += cur.x; a.y += cur.y; a.z += cur.z;
                                __stcg(dst, a);
                            }
                            if (s + 1 < sEnd) { gstu(PG_SAMPLE, j, static_cast<uint32_t>(s + 1)); nextSample = true; }
                        }
                    }
                    hHead += n;
                    push(rayQ, rTail, haveRay, j);
                    push(camQ, cTail, nextSample, j);
                    __syncwarp();
                    continue;
                }
                if (idleMask == 0xFFFFFFFFu) break;   // nothing in flight, nothing queued anywhere: the item is done
            }
            // ---- (C) one TRAVERSE round: every lane with a ray advances by one leaf visit (<= tri_cap triangle tests) ------
            if (state == PL_WALK || state == PL_LEAF) {
                const uint32_t off = MCPT_LEAF_OFF(leaf);
                if (state == PL_WALK) {
                    uint32_t cnt = MCPT_LEAF_CNT(leaf);
                    if (cnt == 7u) cnt = __float_as_uint(ldg4(p.sc.leaf_blob + off + 1).w);
                    li = off + 4u; lend = li + 3u * cnt;
                }
                const uint32_t stop = min(lend, li + 3u * static_cast<uint32_t>(p.tri_cap));
                for (; li < stop; li += 3u)
                    tri_test<true, true, true>(p.sc.leaf_blob + li, ox, oy, oz, dx, dy, dz, best, best, hbeta, hgamma, &htri);
                state = li < lend ? PL_LEAF : PL_WALK;
                if (state == PL_WALK) {
                    const float4 h0 = ldg4(p.sc.leaf_blob + off), h1 = ldg4(p.sc.leaf_blob + off + 1);
                    const LeafExit e = leaf_exit(p.sc, off, h0, h1, ox, oy, oz, ix, iy, iz, nearBits);
                    if (best <= e.t || e.rope == 0xFFFFFFFFu) {
                        state = PL_DONE;
                    } else {
                        tmin = fmaxf(tmin, e.t);
                        if (e.rope & MCPT_ROPE_IS_LEAF) {
                            leaf = e.rope & ~MCPT_ROPE_IS_LEAF;
                        } else {
                            float qx = ox + tmin * dx, qy = oy + tmin * dy, qz = oz + tmin * dz;
                            if (e.axis == 0u) qx = e.plane;
                            if (e.axis == 1u) qy = e.plane;
                            if (e.axis == 2u) qz = e.plane;
                            leaf = rope_locate(p.sc, e.rope, qx, qy, qz, nearBits);
                        }
                    }
                }
            }
            // ---- (D) finished rays: the hit goes into the path's record, the path into the shading queue, the lane is free ---
            {
                const bool fin = state == PL_DONE;
                if (fin) {
                    gst(PG_T, path, best); gstu(PG_TRI, path, htri); gst(PG_BETA, path, hbeta); gst(PG_GAMMA, path, hgamma);
                    atu(W_LEAF, path) = leaf;
                    state = PL_IDLE;
                }
                push(hitQ, hTail, fin, path);
                __syncwarp();
            }
        }
        __syncwarp();
    }
    for (int o = 16; o > 0; o >>= 1) nrays += __shfl_xor_sync(0xFFFFFFFFu, nrays, o);
    if (lane == 0u && nrays) atomicAdd(p.ray_counter, static_cast<unsigned long long>(nrays));
}

}  // namespace mcpt
