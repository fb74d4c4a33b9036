"""CPU tests that pin the oracle (no GPU): known-answer vectors and the reference's own host code."""
import ctypes as C
import os

import numpy as np
import pytest

from conftest import heightfield

# SURVEY.md 8(a5) / BASELINE.md section 2: statistics of the reference's own builder, measured with
# the reference's KDTree.hpp compiled unmodified.
TREE_STATS = {"scene01": (862, 4395, 2198, 4079, 20), "scene02": (1732, 7947, 3974, 7526, 32),
              "scene03": (3024, 39201, 19601, 40055, 32)}


def test_philox_known_answers(oracle):
    # Random123 kat_vectors, philox4x32-10
    kat = [([0, 0, 0, 0], [0, 0], [0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]),
           ([0xffffffff] * 4, [0xffffffff] * 2, [0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd]),
           ([0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344], [0xa4093822, 0x299f31d0],
            [0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1])]
    for ctr, key, want in kat:
        assert oracle.philox(ctr, key).tolist() == want


def test_park_miller_known_answer(oracle):
    # rtx.hlsl:74-82 is the minimal-standard generator; Park & Miller's check value:
    # starting from 1, the 10000th state is 1043618065.
    sd = C.c_uint32(1)
    for _ in range(10000):
        oracle.L.oracle_rand_gen(C.byref(sd))
    assert sd.value == 1043618065
    sd = C.c_uint32(1)
    assert oracle.L.oracle_rand_gen(C.byref(sd)) == np.float32(16807) / np.float32(2147483648.0)


def test_tea_seed_avalanche(oracle):
    # rtx.hlsl:61-72: neighbouring pixels / frames must decorrelate
    a = np.array([oracle.L.oracle_rand_seed(i, 12345) for i in range(4096)], dtype=np.uint32)
    assert len(np.unique(a)) == 4096
    bits = np.unpackbits(a.view(np.uint8)).mean()
    assert 0.48 < bits < 0.52


def test_frame_seeds_are_mt19937_1234(oracle):
    # GraphicsRTX.cpp:170-171,192
    bg = np.random.MT19937()
    bg._legacy_seeding(1234)
    want = bg.random_raw(32).astype(np.uint32)
    assert np.array_equal(oracle.frame_seeds(0, 32), want)
    assert np.array_equal(oracle.frame_seeds(7, 5), want[7:12])


@pytest.mark.parametrize("name", ["scene01", "scene02", "scene03"])
def test_restated_builder_matches_survey_statistics(oracle, scenes, name):
    t = oracle.scene(scenes[name]).tree()
    ntri, nodes, leaves, refs, depth = TREE_STATS[name]
    assert len(scenes[name].tris) == ntri
    assert (len(t.left), t.num_leaves, len(t.leaf_ids), t.max_depth) == (nodes, leaves, refs, depth)


@pytest.mark.parametrize("name,path", [
    ("scene01", "/root/reference/MCRT/QuinEngine/Res/scene01.obj"),
    ("scene02", "/root/reference/CVMCTracer/CVMCTracer/Resources/scene02.obj"),
    ("scene03", "/root/reference/CVMCTracer/CVMCTracer/Resources/scene03.obj")])
def test_restated_builder_equals_reference_builder(oracle, reference, scenes, name, path):
    """Node for node (topology, boxes, planes, leaf id lists) against KDTree.hpp compiled from the reference."""
    if not os.path.exists(path):
        pytest.skip("reference not mounted")
    h, s, _ = reference.load_obj(path)
    # the committed fixture is what the reference's tinyobj produces
    for a, b in ((s.verts, scenes[name].verts), (s.normals, scenes[name].normals), (s.tris, scenes[name].tris),
                 (s.mats, scenes[name].mats)):
        assert np.array_equal(a.view(np.uint32), b.view(np.uint32))
    tr, _ = reference.build_tree(h)
    assert tr.same_as(oracle.scene(s).tree())
    reference.free(h)


def test_restated_builder_equals_reference_on_heightfield(oracle, reference):
    from oracle.pyoracle import SceneArrays
    verts, tri_v = heightfield(40)
    h = reference.from_arrays(verts, tri_v)
    tr, _ = reference.build_tree(h)
    tris = np.zeros((len(tri_v), 7), np.uint32)
    tris[:, :3] = tri_v
    s = SceneArrays(verts, np.array([[0, 1, 0]], np.float32), tris, np.zeros((1, 12), np.float32))
    assert tr.same_as(oracle.scene(s).tree())
    reference.free(h)


def test_samplers_are_unit_and_in_lobe(oracle):
    rng = np.random.default_rng(0)
    out = np.zeros(3, np.float32)
    for _ in range(200):
        n = rng.normal(size=3)
        n = (n / np.linalg.norm(n)).astype(np.float32)
        u1, u2 = rng.uniform(0.01, 0.99, 2)
        oracle.L.oracle_sample_hemi(u1, u2, n.ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p))
        assert abs(np.linalg.norm(out) - 1) < 1e-4 and out @ n > -1e-5
        # cos(theta) = sqrt(1 - u1)  (rtx.hlsl:283)
        assert abs(out @ n - np.sqrt(1 - np.float32(u1))) < 2e-4
    # the two special cases of the rotation (rtx.hlsl:287-292)
    for n in ([0, 1, 0], [0, -1, 0]):
        n = np.array(n, np.float32)
        oracle.L.oracle_sample_hemi(0.3, 0.6, n.ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p))
        assert out @ n > 0


def test_fresnel_reflects_or_refracts(oracle):
    n = np.array([0, 1, 0], np.float32)
    d = np.array([0.6, -0.8, 0], np.float32)
    out = np.zeros(3, np.float32)
    p = lambda a: a.ctypes.data_as(C.c_void_p)
    # u above Tr' -> mirror reflection (rtx.hlsl:246-249)
    oracle.L.oracle_sample_fresnel(0.999, p(n), p(d), 0.9, 1.5, p(out))
    assert np.allclose(out, [0.6, 0.8, 0], atol=1e-6)
    # u below Tr' -> Snell refraction entering (rtx.hlsl:223-227): sin_t = sin_i / Ni
    oracle.L.oracle_sample_fresnel(0.001, p(n), p(d), 0.9, 1.5, p(out))
    assert abs(out[0] - 0.6 / 1.5) < 1e-5 and out[1] < 0
    assert abs(np.linalg.norm(out) - 1) < 1e-5


def test_intersect_pixel_centre_rays_hit_expected_shapes(oracle, scenes):
    """tinyobj shape order of scene01: ids 0-5 pWall1, 6-7 pWall2, 8-9 pWall3,
    10-429 mirror sphere, 430-441 light, 442-861 glass sphere."""
    sc = oracle.scene(scenes["scene01"])
    cfg = oracle.cfg(64, 64, 1, 5, jitter=0)
    rays = oracle.primary_rays(cfg)
    assert np.allclose(rays[:, 2], 16.0)  # origin on the z = 16 plane
    tri, t, b, g, cnt = sc.intersect(rays)
    tri = tri.reshape(64, 64)
    assert set(np.unique(tri[32, 2:6])) <= {8, 9}                                # red wall pWall3
    assert set(np.unique(tri[32, -6:-2])) <= {6, 7}                              # blue wall pWall2
    assert ((tri >= 430) & (tri <= 441)).any()                                   # the light is visible
    assert ((tri >= 10) & (tri <= 429)).any() and (tri >= 442).any()             # both spheres
    hit = tri >= 0
    assert (t.reshape(64, 64)[hit] > 0).all() and cnt.rays == 64 * 64


def test_render_is_deterministic_and_energy_bounded(oracle, scenes):
    sc = oracle.scene(scenes["scene01"])
    cfg = oracle.cfg(32, 32, 4, 5, threads=2)
    s1, q1, d1, c1 = sc.render(cfg, display=True)
    cfg.threads = 1
    s2, q2, d2, c2 = sc.render(cfg, display=True)
    assert np.array_equal(s1, s2) and np.array_equal(d1, d2) and c1.rays == c2.rays
    assert s1.max() / 4 <= 0.8 + 1e-6      # no path can carry more than Ka = 0.8 (throughput <= 1 before RR)
    assert c1.paths == 32 * 32 * 4 and c1.rays >= c1.paths
    # the display value replays rtx.hlsl:401-402: after 4 frames it is (mean)^(1/2.2)
    assert np.allclose(d1[..., :3], (s1 / 4) ** (1 / 2.2), atol=2e-3)
    # sample ranges compose: frames [0,4) == [0,2) + [2,4)
    a = oracle.cfg(32, 32, 2, 5)
    b = oracle.cfg(32, 32, 2, 5, sample_offset=2)
    assert np.allclose(sc.render(a)[0] + sc.render(b)[0], s1, rtol=0, atol=1e-12)


def test_triangle_test_equals_the_cuda_twins_arithmetic_with_reference_det(oracle):
    """The one explicit operation order the reference gives for the Cramer system is Matrix33f::det
    (CVMCTracer/.../Framework/Math.hpp:169-175) applied to the matrices of CUTracer.cu:58-81.  The reference's
    own det(), compiled in oracle/_ref/libmcpt_refcv.so, must give the oracle's t / beta / gamma bit for bit."""
    from oracle.pyoracle import ReferenceCV, SceneArrays, TreeArrays
    if not ReferenceCV.available():
        pytest.skip("oracle/_ref/libmcpt_refcv.so not built")
    cv = ReferenceCV()
    g = np.random.default_rng(5)
    f32 = np.float32
    checked = 0
    for _ in range(300):
        tri = g.uniform(-3, 3, (3, 3)).astype(f32)
        o = g.uniform(-3, 3, 3).astype(f32)
        target = (tri[0] * f32(0.3) + tri[1] * f32(0.3) + tri[2] * f32(0.4)).astype(f32) + g.normal(0, 0.3, 3).astype(f32)
        d = (target - o).astype(f32)
        d = (d / np.linalg.norm(d)).astype(f32)
        a, b, c = tri
        col = lambda x, y, z: [x[0], y[0], z[0], x[1], y[1], z[1], x[2], y[2], z[2]]   # three column vectors, row-major
        e1, e2, s = (a - b).astype(f32), (a - c).astype(f32), (a - o).astype(f32)
        detA = cv.det33(col(e1, e2, d))
        with np.errstate(all="ignore"):
            beta = f32(cv.det33(col(s, e2, d)) / detA)
            gamma = f32(cv.det33(col(e1, s, d)) / detA)
            t = f32(cv.det33(col(e1, e2, s)) / detA)
        hit = bool(f32(beta + gamma) < 1 and beta > 0 and gamma > 0 and t > 0 and t < 10000)
        scene = SceneArrays(tri, np.array([[0, 1, 0]], f32), np.array([[0, 1, 2, 0, 0, 0, 0]], np.uint32), np.zeros((1, 12), f32))
        box = np.array([[-10, -10, -10, 10, 10, 10]], f32)
        one = TreeArrays(np.array([0xFFFFFFFF], np.uint32), np.array([0xFFFFFFFF], np.uint32), box, np.zeros(1, np.uint32),
                         np.zeros(1, f32), np.zeros(1, np.uint32), np.array([1], np.uint32), np.arange(1, dtype=np.uint32), 0, 1)
        otri, ot, ob, og, _ = oracle.scene(scene, one).intersect(np.concatenate([o, d])[None].astype(f32))
        assert (otri[0] == 0) == hit
        if hit:
            assert (ot[0].view(np.uint32), ob[0].view(np.uint32), og[0].view(np.uint32)) == (t.view(np.uint32), beta.view(np.uint32), gamma.view(np.uint32))
            checked += 1
    assert checked > 50


def test_cuda_twin_mode_semantics(oracle, scenes):
    """CUTracer.cu:98-218 restated: origin at the eye, horizontal fov 60 deg with square pixels, at most
    depth + 1 = 8 intersect calls per path, no roulette (throughput <= 1, so radiance <= Ka * ILLUM)."""
    cfg = oracle.cfg(80, 60, 1, 7, tracer_mode=1, emission_scale=10.0, jitter=0, fovy=np.float32(60 * np.float32(3.14159265359) / 180))
    rays = oracle.primary_rays(cfg).reshape(60, 80, 6)
    assert np.array_equal(rays[..., :3], np.broadcast_to(np.float32([0, 5, 17]), (60, 80, 3)))
    assert np.allclose(np.linalg.norm(rays[..., 3:], axis=-1), 1, atol=1e-6)
    # jitter off: u = 0.5 -> bias = pixel index exactly; column 0 looks along -tan(30 deg), column w/2 straight ahead
    d = rays[30, :, 3:]
    assert np.allclose(d[0, 0] / -d[0, 2], -np.tan(np.pi / 6), atol=1e-6) and abs(d[40, 0]) < 1e-7
    assert abs(rays[30, 40, 4]) < 1e-7 and np.allclose(rays[0, 40, 4] / -rays[0, 40, 5], 0.75 * np.tan(np.pi / 6), atol=1e-6)
    sc = oracle.scene(scenes["scene01"])
    cfg = oracle.cfg(40, 30, 8, 7, tracer_mode=1, emission_scale=10.0, rng_mode=1, seed=2, fovy=np.float32(60 * np.float32(3.14159265359) / 180))
    s, _, _, c = sc.render(cfg)
    assert s.max() / 8 <= 8.0 + 1e-5 and s.max() > 0
    assert c.paths == 40 * 30 * 8 and c.paths <= c.rays <= 8 * c.paths
    # the MCRT estimator on the same scene may take up to 3 * depth + 1 rays and applies roulette: a different sample set
    s2, _, _, c2 = sc.render(oracle.cfg(40, 30, 8, 7, rng_mode=1, seed=2))
    assert c2.rays != c.rays


# ---- the shading half, pinned to the reference's own code ---------------------------------------------------
# CVMCTracer/CVMCTracer/CUDA/Utils.hpp:46-137 is the same math as rtx.hlsl:213-302 and compiles
# unmodified on the host (oracle/ref_bridge_cv.cpp).  The bridge hands back the uniforms each call drew from the
# reference's own cuRAND stream; the restatement is fed those numbers as they are.
def _sampler_inputs(n, seed):
    """Unit normals and incoming directions incl. the rotation's special cases (|n.y +- 1| < FLT_EPSILON,
    rtx.hlsl:262-267 / :287-292), near-special normals, grazing and normal incidence."""
    g = np.random.default_rng(seed)
    nrm = g.normal(size=(n, 3))
    nrm /= np.linalg.norm(nrm, axis=1, keepdims=True)
    nrm = nrm.astype(np.float32)
    nrm[0], nrm[1] = [0, 1, 0], [0, -1, 0]
    nrm[2] = [0, np.float32(1) - np.float32(2.0 ** -24), 0]      # inside the epsilon
    nrm[3] = [3e-4, np.sqrt(1 - 9e-8), 0]                         # just outside it
    nrm[4] = [0, -1 + 2.0 ** -24, 1e-5]
    nrm[5], nrm[6], nrm[7] = [1, 0, 0], [0, 0, -1], [0.6, 0.0, 0.8]
    d = g.normal(size=(n, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    d = d.astype(np.float32)
    d[8] = -nrm[8]                                                # normal incidence
    t = np.cross(nrm[9], [0.3, 0.5, 0.8]); t /= np.linalg.norm(t)
    d[9] = (t + 1e-3 * nrm[9]).astype(np.float32)                # grazing
    return nrm, d


def _bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


@pytest.fixture(scope="module")
def refcv():
    from oracle.pyoracle import ReferenceCV, build
    build()
    if not ReferenceCV.available():
        pytest.skip("oracle/_ref/libmcpt_refcv.so not built (needs /root/reference)")
    return ReferenceCV()


def test_sample_hemi_bit_equal_to_reference_utils_hpp(oracle, refcv):
    """rtx.hlsl:278-302 restated == CUDA/Utils.hpp:46-70 compiled from /root/reference, bit for bit."""
    nrm, _ = _sampler_inputs(30000, 11)
    want, u = refcv.sample_hemi(1234, nrm)
    assert (u > 0).all() and (u <= 1).all() and len(np.unique(u)) > 59000
    got = oracle.sample_batch(oracle.HEMI, u, nrm)
    assert np.array_equal(_bits(got), _bits(want))


def test_sample_phong_bit_equal_to_reference_utils_hpp(oracle, refcv):
    """rtx.hlsl:253-276 restated == CUDA/Utils.hpp:72-95 (Ns arrives as PWuint there: integral exponents)."""
    nrm, d = _sampler_inputs(30000, 12)
    Ns = np.random.default_rng(3).choice([2, 5, 10, 20, 50, 1000], 30000).astype(np.uint32)
    want, u = refcv.sample_phong(99, nrm, d, Ns)
    got = oracle.sample_batch(oracle.PHONG, u, nrm, d, Ns.astype(np.float32))
    assert np.array_equal(_bits(got), _bits(want))


def test_sample_fresnel_twin_bit_equal_to_reference_utils_hpp(oracle, refcv):
    """CUDA/Utils.hpp:97-137 restated (sample_fresnel_twin: normalised in the refract branches only) is bit-equal to
    the reference's code; the shader form (rtx.hlsl:213-251, normalize() on every branch, :250) takes the same
    branch for every tuple and differs from it only by that normalisation (a few ulp)."""
    n = 40000
    nrm, d = _sampler_inputs(n, 13)
    g = np.random.default_rng(4)
    Tr = g.choice([0.9, 0.5, 1.0, 0.1], n).astype(np.float32)
    Ni = g.choice([1.5, 1.33, 2.4, 1.0], n).astype(np.float32)
    want, u = refcv.sample_fresnel(7, nrm, d, Tr, Ni)
    got = oracle.sample_batch(oracle.FRESNEL_TWIN, u, nrm, d, Tr, Ni)
    assert np.array_equal(_bits(got), _bits(want))
    hlsl = oracle.sample_batch(oracle.FRESNEL, u, nrm, d, Tr, Ni)
    assert np.abs(hlsl - want).max() < 1e-6
    # every branch of :218-249 is exercised: refraction entering / leaving, total internal reflection, reflection
    c = (d * nrm).sum(1)
    refl = (d - 2 * c[:, None] * nrm)
    is_refl = np.abs(want - refl).max(1) < 1e-5
    assert (is_refl & (c > 0)).sum() > 1000 and (~is_refl & (c > 0)).sum() > 1000 and (~is_refl & (c <= 0)).sum() > 1000
    tir = (c > 0) & (1 - (1 - c * c) * Ni * Ni < 0) & (u < Tr * (1 - (1 - np.abs(c)) ** 5))
    assert tir.sum() > 500 and is_refl[tir].all()


def test_normalize_twin_bit_equal_to_reference_utils_hpp(oracle, refcv):
    """CUDA/Utils.hpp:27-34 (divide by the length unless it is below FLT_EPSILON): the twin's path loop normalises the
    shading normal and the camera ray with it (CUTracer.cu:126,210)."""
    g = np.random.default_rng(8)
    for v in list(g.normal(size=(2000, 3)).astype(np.float32)) + [np.float32([0, 0, 0]), np.float32([1e-8, 0, 0]), np.float32([3e-8, 4e-8, 0])]:
        assert np.array_equal(_bits(oracle.normalize_twin(v)), _bits(refcv.normalize(v)))


def test_sampler_api_takes_uniforms_as_they_are(oracle):
    """The reference streams are not on Philox's 2^-24 grid: curand_uniform is k 2^-32 + 2^-33, Park-Miller sd / (2^31 - 1).
    oracle_sample_* must use the float it is given (round 1 snapped it to 2^-24)."""
    n = np.float32([[0.36, 0.48, 0.8]])
    u = np.float32(np.float32(0.3) + np.float32(2.0 ** -25))   # one ulp above 0.3: off the 2^-24 grid
    a = oracle.sample_batch(oracle.HEMI, np.float32([[u, 0.25]]), n)
    b = oracle.sample_batch(oracle.HEMI, np.float32([[0.3, 0.25]]), n)
    assert not np.array_equal(_bits(a), _bits(b))
    # cos(theta) about the normal is sqrt(1 - u) for the float given
    assert abs(float(a[0] @ n[0]) - np.sqrt(1 - float(u))) < 1e-6
