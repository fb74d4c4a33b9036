"""Random triangle soups -- content the shipped scenes do not have: long slivers, zero-area triangles, duplicates,
triangles in a split plane, triangles far larger and far smaller than their neighbours -- through every tree builder and
every traversal against the oracle's brute force over ALL triangles (the reference's arithmetic, no tree): triangle ids
equal and t / beta / gamma bit-equal, except exact ties between coincident triangles (same t bits, other id).  The
reference-rule tree is additionally compared with the oracle's restatement node for node (CPU)."""
import numpy as np
import pytest

from oracle.pyoracle import SceneArrays, TreeArrays


def _soup(seed, n):
    g = np.random.Generator(np.random.Philox(seed))
    c = g.uniform(-4, 4, (n, 1, 3))
    size = 10.0 ** g.uniform(-2.5, 0.7, (n, 1, 1))                 # 0.003 .. 5 units across
    v = c + size * g.normal(size=(n, 3, 3))
    k = n // 10
    v[:k, 2] = v[:k, 1] + 1e-3 * (v[:k, 0] - v[:k, 1])            # slivers: the third corner almost on an edge
    v[k:k + 3, 2] = v[k:k + 3, 1]                                 # zero area: two corners coincide
    v[k + 3:k + 6] = v[k + 6:k + 9]                               # exact duplicates
    v[k + 9:2 * k, :, 1] = np.round(v[k + 9:2 * k, :1, 1])        # flat in y on integer planes (triangles lying IN candidate split planes)
    verts = v.reshape(-1, 3).astype(np.float32)
    tris = np.zeros((n, 7), np.uint32)
    tris[:, :3] = np.arange(3 * n, dtype=np.uint32).reshape(n, 3)
    mats = np.zeros((1, 12), np.float32); mats[0, 3:6] = 0.7
    return SceneArrays(verts, np.array([[0, 1, 0]], np.float32), tris, mats)


def _rays(s, n, seed):
    g = np.random.Generator(np.random.Philox(seed))
    lo, hi = s.verts.min(0), s.verts.max(0)
    o = g.uniform(lo - 0.5, hi + 0.5, (n, 3))
    d = g.normal(size=(n, 3)); d /= np.linalg.norm(d, axis=1, keepdims=True)
    d[: n // 16, 0] = 0.0                                         # axis-parallel components: 1/0 in the slab tests
    d[n // 16: n // 8, 1] = 0.0
    d[: n // 8] /= np.maximum(np.linalg.norm(d[: n // 8], axis=1, keepdims=True), 1e-6)
    return np.concatenate([o, d], 1).astype(np.float32)


def _bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def _brute(oracle, s, rays):
    n = len(s.tris)
    box = np.concatenate([s.verts.min(0) - 1, s.verts.max(0) + 1]).astype(np.float32)[None]
    one = TreeArrays(np.array([0xFFFFFFFF], np.uint32), np.array([0xFFFFFFFF], np.uint32), box, np.zeros(1, np.uint32),
                     np.zeros(1, np.float32), np.zeros(1, np.uint32), np.array([n], np.uint32), np.arange(n, dtype=np.uint32), 0, 1)
    return oracle.scene(s, one).intersect(rays)


# The reference's split rule (no traversal cost: KDTree.hpp:98-153) keeps splitting overlapping triangles down to its depth
# cap, so its tree of a dense soup explodes (3 000 of these triangles: more than 4 GiB of leaf blocks, which the library
# refuses with MCPT_ERR_INVALID): the reference-rule cases stay small, the SAH builders take the big ones.
@pytest.mark.parametrize("seed,n", [(1, 60), (2, 400)])
def test_reference_rule_tree_of_a_soup_equals_the_oracles(tracer, oracle, seed, n):
    s = _soup(seed, n)
    tracer.create_geometry(s.verts, s.normals, s.tris, s.mats, tracer.BUILD_REFERENCE)
    t, o = tracer.get_tree(), oracle.scene(s).tree()
    assert np.array_equal(t.left, o.left) and np.array_equal(t.right, o.right)
    assert np.array_equal(_bits(t.aabb), _bits(o.aabb)) and np.array_equal(t.split_axis, o.split_axis)
    assert np.array_equal(_bits(t.split_value), _bits(o.split_value))
    assert np.array_equal(t.leaf_offset, o.leaf_offset) and np.array_equal(t.leaf_count, o.leaf_count) and np.array_equal(t.leaf_ids, o.leaf_ids)
    for mode in (tracer.BUILD_REFERENCE, tracer.BUILD_SAH):
        tracer.create_geometry(s.verts, s.normals, s.tris, s.mats, mode)
        tracer.check_stackless_layout()
    tracer.destroy_geometry()


@pytest.mark.gpu
@pytest.mark.parametrize("seed,n", [(1, 60), (2, 400), (3, 3000), (4, 20000)])
def test_soup_hits_equal_brute_force_for_every_builder_and_traversal(gpu, oracle, seed, n):
    T = gpu
    s = _soup(seed, n)
    rays = _rays(s, 1 << 15, seed + 100)
    btri, bt, bb, bg, _ = _brute(oracle, s, rays)
    assert (btri >= 0).mean() > 0.05
    for mode in ((T.BUILD_REFERENCE, T.BUILD_SAH, T.BUILD_GPU) if n <= 400 else (T.BUILD_SAH, T.BUILD_GPU)):
        T.create_geometry(s.verts, s.normals, s.tris, s.mats, mode)
        T.check_stackless_layout()
        for variant in (T.ISECT_REFERENCE, T.ISECT_KD, T.ISECT_ROPES):
            if variant == T.ISECT_REFERENCE and mode != T.BUILD_REFERENCE:
                continue     # the reference-order traversal is defined on the reference's clipped boxes
            tri, t, b, g = T.intersect_batch(rays, variant)
            if variant == T.ISECT_REFERENCE:
                # the reference's own traversal may drop a hit its clipped-box slab test rounds away: it must
                # agree with the ORACLE's reference-order traversal exactly, and with brute force on all but a few rays
                otri, ot, _, _, _ = oracle.scene(s).intersect(rays)
                assert np.array_equal(tri, otri) and np.array_equal(_bits(t)[otri >= 0], _bits(ot)[otri >= 0])
                assert ((tri != btri) & (_bits(t) != _bits(bt))).mean() < 2e-3
                continue
            hit = btri >= 0
            assert np.array_equal(tri >= 0, hit), (mode, variant, int(((tri >= 0) != hit).sum()))
            assert np.array_equal(_bits(t)[hit], _bits(bt)[hit]), (mode, variant)
            same = tri == btri
            assert same[hit].mean() > 0.995                                      # other id only on exact ties (duplicates, coincident edges)
            ok = hit & same
            assert np.array_equal(_bits(b)[ok], _bits(bb)[ok]) and np.array_equal(_bits(g)[ok], _bits(bg)[ok])
    T.destroy_geometry()


def _dressed_soup(seed, n):
    """The soup with per-vertex shading normals that do NOT follow the geometry and four kinds of material: emitters,
    Lambert, glossy Phong (Ns 30) and glass (Tr 0.8, Ni 1.5) -- every branch of sampleMC on content nobody arranged."""
    s = _soup(seed, n)
    g = np.random.Generator(np.random.Philox(seed + 1000))
    nn = g.normal(size=(3 * n, 3)); nn /= np.linalg.norm(nn, axis=1, keepdims=True)
    mats = np.zeros((4, 12), np.float32)
    mats[0, 0:3] = (6, 5, 4)                                  # Ka > 0: emitter
    mats[1, 3:6] = (0.7, 0.6, 0.5)                            # Lambert
    mats[2, 3:6] = 0.1; mats[2, 6:9] = 0.8; mats[2, 9] = 30   # Phong, Ns = 30
    mats[3, 3:6] = 0.9; mats[3, 10] = 0.8; mats[3, 11] = 1.5  # glass
    tris = s.tris.copy()
    tris[:, 3:6] = np.arange(3 * n, dtype=np.uint32).reshape(n, 3)
    tris[:, 6] = g.choice(4, n, p=[0.15, 0.45, 0.2, 0.2]).astype(np.uint32)
    return SceneArrays(s.verts, nn.astype(np.float32), tris, mats)


@pytest.mark.gpu
@pytest.mark.parametrize("seed,n,mode", [(5, 300, "reference"), (6, 2500, "sah"), (7, 2500, "gpu")])
def test_soup_renders_follow_the_oracles_paths(gpu, oracle, seed, n, mode):
    """Shared Philox stream: the GPU and the oracle trace the same paths through the soup; pixels agree to rounding except
    where an ulp (hardware rsqrt / sincospif against libm) flips a discrete decision -- and then the path, not the estimator,
    differs: ray counts within 0.5 %, image mean within 1 %."""
    T = gpu
    s = _dressed_soup(seed, n)
    T.create_geometry(s.verts, s.normals, s.tris, s.mats, {"reference": T.BUILD_REFERENCE, "sah": T.BUILD_SAH, "gpu": T.BUILD_GPU}[mode])
    W, H, spp, depth = 80, 64, 6, 5
    eye, at = (0.0, 0.0, 9.0), (0.0, 0.0, 8.0)
    osum, osq, _, ocnt = oracle.scene(s).render(oracle.cfg(W, H, spp, depth, rng_mode=1, seed=5, eye=eye, at=at))
    mean = T.render_scene(T.render_cfg(W, H, spp, depth, seed=5, eye=eye, at=at))
    c = T.counters()
    assert c.paths == W * H * spp == ocnt.paths
    assert abs(int(c.rays) - int(ocnt.rays)) <= 0.005 * ocnt.rays, (c.rays, ocnt.rays)
    assert mean.max() > 0 and np.isfinite(mean).all()
    close = np.isclose(mean, osum / spp, rtol=1e-3, atol=1e-5).all(-1)
    assert close.mean() > 0.95, close.mean()
    assert abs(mean.astype(np.float64).mean() - (osum / spp).mean()) < 0.01 * (osum / spp).mean() + 1e-6
    # and the kernels agree among themselves on it: per-ray kernel, reference-order traversal
    simple = T.render_scene(T.render_cfg(W, H, spp, depth, seed=5, eye=eye, at=at, flags=T.FLAG_SIMPLE_KERNEL))
    assert np.isclose(mean, simple, rtol=1e-4, atol=1e-6).all(-1).mean() > 0.999
    T.destroy_geometry()
