"""MCPT_BUILD_GPU: the level-synchronous SAH builder on the device (csrc/kd_build_gpu.cu; SURVEY.md 8(f) rank 1, first option,
anchor MCRT/QuinEngine/Utils/KDTree.hpp:58-287).  Its tree differs from the reference's and from the host SAH builder's; what
must not differ is what any traversal of it returns: the oracle's closest hit, bit for bit."""
import hashlib

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
CAMERA_Z = {"scene01": 17.0, "scene01_lambert": 17.0, "scene02": 23.0, "scene03": 4.75}


def _bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def _random_rays(s, n, seed):
    g = np.random.Generator(np.random.Philox(seed))
    lo, hi = s.verts.min(0), s.verts.max(0)
    o = g.uniform(lo, hi, (n, 3))
    d = g.normal(size=(n, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    return np.concatenate([o, d], 1).astype(np.float32)


def _brute(oracle, s):
    from oracle.pyoracle import TreeArrays
    n = len(s.tris)
    box = np.concatenate([s.verts.min(0) - 1, s.verts.max(0) + 1]).astype(np.float32)[None]
    one = TreeArrays(np.array([0xFFFFFFFF], np.uint32), np.array([0xFFFFFFFF], np.uint32), box, np.zeros(1, np.uint32),
                     np.zeros(1, np.float32), np.zeros(1, np.uint32), np.array([n], np.uint32), np.arange(n, dtype=np.uint32), 0, 1)
    return oracle.scene(s, one)


def _check_tree(T, s):
    """Structural invariants of the layout every consumer relies on (ShaderResource.hpp:128-179 conventions)."""
    t = T.get_tree()
    n = len(t.left)
    inner = t.split_axis != 0
    assert (t.left[~inner] == 0xFFFFFFFF).all() and (t.split_axis <= 3).all()
    kids = np.sort(np.concatenate([t.left[inner], t.left[inner] + 1]))
    assert np.array_equal(kids, np.arange(1, n))                                     # breadth-first, children adjacent, every node reachable once
    assert (t.left[inner] > np.flatnonzero(inner)).all()                              # children after their parent
    # cells: children partition the parent's cell at the plane
    a = t.split_axis[inner].astype(int) - 1
    L = t.left[inner].astype(int)
    P = np.flatnonzero(inner)
    assert np.array_equal(t.aabb[L, 3 + a], t.split_value[inner]) and np.array_equal(t.aabb[L + 1, a], t.split_value[inner])
    for ax in range(3):
        m = a != ax
        assert np.array_equal(t.aabb[L[m], ax], t.aabb[P[m], ax]) and np.array_equal(t.aabb[L[m] + 1, 3 + ax], t.aabb[P[m], 3 + ax])
    # leaves: ids ascending, every triangle somewhere, every referenced triangle's box overlaps its leaf's cell
    leaves = np.flatnonzero(~inner)
    assert t.leaf_count[leaves].sum() == len(t.leaf_ids) and np.array_equal(np.unique(t.leaf_ids), np.arange(len(s.tris)))
    v = s.verts[s.tris[:, :3]]
    lo, hi = v.min(1), v.max(1)
    for k in leaves[:: max(1, len(leaves) // 2000)]:
        ids = t.leaf_ids[t.leaf_offset[k]: t.leaf_offset[k] + t.leaf_count[k]]
        assert (np.diff(ids.astype(np.int64)) > 0).all()
        assert (lo[ids] <= t.aabb[k, 3:]).all() and (hi[ids] >= t.aabb[k, :3]).all()
    return t


@pytest.mark.parametrize("name", ["scene01", "scene01_lambert", "scene02", "scene03"])
def test_gpu_built_tree_gives_the_oracles_hits(gpu, oracle, scenes, name):
    T, s = gpu, scenes[name]
    T.create_geometry(s.verts, s.normals, s.tris, s.mats, T.BUILD_GPU)
    t = _check_tree(T, s)
    T.check_stackless_layout()                                                        # leaf blocks and neighbour links built from it are consistent
    info = T.tree_info()
    assert info.num_nodes == len(t.left) and info.num_leaf_refs == len(t.leaf_ids)
    ez = CAMERA_Z[name]
    cam = oracle.primary_rays(oracle.cfg(512, 512, 1, 5, jitter=0, eye=(0, 5, ez), at=(0, 5, ez - 1)))
    osc, brute = oracle.scene(s), _brute(oracle, s)
    for rays in (cam, _random_rays(s, 1 << 18, 5)):
        tri, tt, b, g, _ = osc.intersect(rays)
        hit = tri >= 0
        for variant in (T.ISECT_REFERENCE, T.ISECT_KD, T.ISECT_ROPES):
            gtri, gt, gb, gg = T.intersect_batch(rays, variant)
            bad = np.flatnonzero((gtri != tri) | (hit & ((_bits(gt) != _bits(tt)) | (_bits(gb) != _bits(b)) | (_bits(gg) != _bits(g)))))
            assert len(bad) <= 1e-4 * len(rays), (variant, len(bad))
            if len(bad):   # adjudicated by the oracle's arithmetic over ALL triangles (tests/test_fullsize.py has the long form)
                btri, bt, bb, bg, _ = brute.intersect(rays[bad])
                ok = (btri == gtri[bad]) & (_bits(bt) == _bits(gt[bad]))
                edge = np.minimum(np.minimum(bb, bg), 1 - bb - bg)
                assert (ok | ((btri >= 0) & (edge < 4e-7))).all(), (variant, bad[~ok])


def test_gpu_builder_is_deterministic_and_renders_the_same_image(gpu, scenes):
    """The scatter's atomics leave references in arrival order; leaves are sorted afterwards, so two builds give the same tree
    byte for byte.  Same hits -> same paths -> the image of the host-built SAH tree (up to the rare ray that grazes an edge)."""
    T, s = gpu, scenes["scene01"]
    digests, imgs = [], []
    for mode in (T.BUILD_GPU, T.BUILD_GPU, T.BUILD_SAH):
        T.create_geometry(s.verts, s.normals, s.tris, s.mats, mode)
        t = T.get_tree()
        h = hashlib.sha256()
        for a in (t.left, t.aabb, t.split_axis, t.split_value, t.leaf_offset, t.leaf_count, t.leaf_ids):
            h.update(np.ascontiguousarray(a).tobytes())
        digests.append(h.hexdigest())
        imgs.append(T.render_scene(T.render_cfg(160, 120, 8, 5, seed=4)))
    assert digests[0] == digests[1] and digests[0] != digests[2]
    assert np.array_equal(imgs[0], imgs[1])
    assert np.isclose(imgs[0], imgs[2], rtol=1e-4, atol=1e-6).all(-1).mean() > 0.999


def test_gpu_builder_at_one_million_triangles(gpu, oracle, scenes):
    """BASELINE configs[3]'s scene: built on the device, checked structurally, its neighbour links verified, and 2 048 rays
    adjudicated by brute force over all 1 002 550 triangles."""
    from montecarlopathtracer_b200 import scenes as S
    from oracle.pyoracle import SceneArrays
    T, b = gpu, scenes["scene01_lambert"]
    s = SceneArrays(*S.cornell_heightfield((b.verts, b.normals, b.tris, b.mats), 708))
    T.create_geometry(s.verts, s.normals, s.tris, s.mats, T.BUILD_GPU)
    info = T.tree_info()
    assert info.build_seconds < 1.0
    _check_tree(T, s)
    T.check_stackless_layout()
    cam = T.primary_rays(T.render_cfg(1024, 1024, 1, 8, flags=T.FLAG_NO_JITTER), 0).reshape(1024, 1024, 6)
    g = np.random.default_rng(11)
    rays = np.concatenate([cam[g.integers(0, 1024, 1024), g.integers(0, 1024, 1024)], _random_rays(s, 1024, 19)])
    btri, bt, bb, bg, _ = _brute(oracle, s).intersect(rays)
    for variant in (T.ISECT_REFERENCE, T.ISECT_KD, T.ISECT_ROPES):
        tri, t, be, ga = T.intersect_batch(rays, variant)
        bad = np.flatnonzero((tri != btri) | ((btri >= 0) & (_bits(t) != _bits(bt))))
        assert len(bad) <= 2, (variant, len(bad))
        if len(bad):
            edge = np.minimum(np.minimum(bb[bad], bg[bad]), 1 - bb[bad] - bg[bad])
            assert (np.where(tri[bad] >= 0, t[bad], np.inf) > bt[bad]).all() and (edge < 4e-7).all()
    img = T.render_scene(T.render_cfg(256, 256, 16, 8, seed=2))
    assert np.isfinite(img).all() and img.mean() > 1e-4


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["scene01", "scene02", "scene03"])
def test_device_pack_equals_the_host_pack(gpu, scenes, name):
    """MCPT_BUILD_GPU makes the render-path tables (leaf blocks with neighbour links, location nodes, shading records) on the
    device; the host's pack_scene + build_ropes of the same downloaded tree must give the same words, and renders through
    either must be the same image."""
    T, s = gpu, scenes[name]
    try:
        T.set_tuning("defaults", 0)
        T.create_geometry(s.verts, s.normals, s.tris, s.mats, T.BUILD_GPU)
        assert T.compare_device_tables() == (0, 0, 0)
        T.check_stackless_layout()
        cfg = T.render_cfg(120, 90, 6, 5, seed=4)
        a = T.render_scene(cfg); ra = T.counters().rays
        rays = np.random.default_rng(3).normal(size=(4096, 6)).astype(np.float32); rays[:, :3] *= 3; rays[:, 1] += 5
        hits_a = T.intersect_batch(rays, T.ISECT_ROPES)
        T.set_tuning("verify_layout", 1)      # install-time self-check: host pack of the same tree, layout check, word-for-word comparison
        T.create_geometry(s.verts, s.normals, s.tris, s.mats, T.BUILD_GPU)
        T.set_tuning("verify_layout", 0)
        T.set_tuning("device_pack", 0)
        T.create_geometry(s.verts, s.normals, s.tris, s.mats, T.BUILD_GPU)
        assert T.compare_device_tables() == (0, 0, 0)
        b = T.render_scene(cfg); rb = T.counters().rays
        hits_b = T.intersect_batch(rays, T.ISECT_ROPES)
        assert np.array_equal(a, b) and ra == rb
        for x, y in zip(hits_a, hits_b):
            assert np.array_equal(x, y)
    finally:
        T.set_tuning("defaults", 0)


@pytest.mark.gpu
def test_device_pack_at_one_million_triangles(gpu, scenes):
    from montecarlopathtracer_b200 import scenes as S
    T = gpu
    T.create_geometry(scenes["scene01_lambert"].verts, scenes["scene01_lambert"].normals, scenes["scene01_lambert"].tris, scenes["scene01_lambert"].mats, T.BUILD_SAH)
    hf = S.cornell_heightfield(T.get_scene(), 708)
    T.create_geometry(*hf, T.BUILD_GPU)
    assert T.compare_device_tables() == (0, 0, 0)
    T.check_stackless_layout()
    # a refused scene after a device-packed one: the previous scene is untouched
    img = T.render_scene(T.render_cfg(64, 48, 2, 5, seed=1))
    bad = hf[2].copy(); bad[5, 1] = len(hf[0]) + 7      # a vertex index past the table
    with pytest.raises(RuntimeError):
        T.create_geometry(hf[0], hf[1], bad, hf[3], T.BUILD_GPU)
    assert np.array_equal(T.render_scene(T.render_cfg(64, 48, 2, 5, seed=1)), img)
    # the host copy of the tables appears when something asks for it (here: the parity tables of the stack traversal)
    rays = np.random.default_rng(5).normal(size=(2048, 6)).astype(np.float32); rays[:, 1] += 5
    a, b = T.intersect_batch(rays, T.ISECT_ROPES), T.intersect_batch(rays, T.ISECT_KD)
    assert np.array_equal(a[0], b[0]) and np.array_equal(_bits(a[1]), _bits(b[1]))
    T.destroy_geometry()
