"""TEST INFRASTRUCTURE -- ctypes/numpy front end of the CPU oracle.  NOT product code.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this module (see oracle/oracle.h).

Two libraries sit behind it:

* ``oracle/liboracle.so``       -- the restatement (kd_restate.cpp, shader_restate.cpp), built
  anywhere with ``make -C oracle liboracle.so``;
* ``oracle/_ref/libmcpt_ref.so`` -- the reference's OWN tinyobjloader + KDTree.hpp compiled
  from /root/reference (ref_bridge.cpp); exists only where it was built in the authoring
  container (it travels to the GPU box as a prebuilt file).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from dataclasses import dataclass

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_SO = os.path.join(HERE, "liboracle.so")
REF_SO = os.path.join(HERE, "_ref", "libmcpt_ref.so")
REFCV_SO = os.path.join(HERE, "_ref", "libmcpt_refcv.so")
REFERENCE_ROOT = "/root/reference"

_vp = C.c_void_p


def _ptr(a):
    return None if a is None else a.ctypes.data_as(_vp)


def build(force: bool = False) -> None:
    """Compile liboracle.so (always possible) and _ref (only where the reference is mounted)."""
    if force or not os.path.exists(ORACLE_SO) or _stale(ORACLE_SO):
        subprocess.check_call(["make", "-C", HERE, "liboracle.so"], stdout=subprocess.DEVNULL)
    if os.path.isdir(REFERENCE_ROOT) and (force or not os.path.exists(REF_SO) or not os.path.exists(REFCV_SO)
                                          or os.path.getmtime(REF_SO) < os.path.getmtime(os.path.join(HERE, "ref_bridge.cpp"))
                                          or os.path.getmtime(REFCV_SO) < os.path.getmtime(os.path.join(HERE, "ref_bridge_cv.cpp"))):
        subprocess.check_call(["make", "-C", HERE, "ref"], stdout=subprocess.DEVNULL)
    if os.path.isdir(REFERENCE_ROOT):
        # the compiled drop-in proof (needs libmcpt.so) and the reference's own CUDA tracer for sm_100a (tests/test_dropin.py)
        mcpt = os.path.join(HERE, "..", "montecarlopathtracer_b200", "lib", "libmcpt.so")
        shim = [os.path.join(HERE, "..", "tests", "dropin", f) for f in ("CUTracerMcpt.cpp", "main_dropin.cpp")]
        dropin, refcu = os.path.join(HERE, "_ref", "dropin_main"), os.path.join(HERE, "_ref", "cutracer_ref")
        if os.path.exists(mcpt) and (force or not os.path.exists(dropin) or any(os.path.getmtime(f) > os.path.getmtime(dropin) for f in shim)):
            subprocess.check_call(["make", "-C", HERE, "dropin"], stdout=subprocess.DEVNULL)
        if force or not os.path.exists(refcu) or os.path.getmtime(shim[1]) > os.path.getmtime(refcu):
            subprocess.check_call(["make", "-C", HERE, "refcu"], stdout=subprocess.DEVNULL)


def _stale(so: str) -> bool:
    t = os.path.getmtime(so)
    for f in ("kd_restate.cpp", "shader_restate.cpp", "oracle_api.cpp", "oracle_internal.h", "oracle.h"):
        if os.path.getmtime(os.path.join(HERE, f)) > t:
            return True
    return False


class Cfg(C.Structure):
    """Mirror of ``oracle_cfg`` (oracle.h)."""
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("spp", C.c_int32), ("depth", C.c_int32),
                ("rng_mode", C.c_int32), ("seed", C.c_uint32), ("sample_offset", C.c_int32),
                ("jitter", C.c_int32), ("eye", C.c_float * 3), ("at", C.c_float * 3), ("up", C.c_float * 3),
                ("fovy", C.c_float), ("x0", C.c_int32), ("y0", C.c_int32), ("x1", C.c_int32),
                ("y1", C.c_int32), ("threads", C.c_int32), ("tracer_mode", C.c_int32), ("emission_scale", C.c_float)]


class Counters(C.Structure):
    _fields_ = [("paths", C.c_uint64), ("rays", C.c_uint64), ("nodes_popped", C.c_uint64),
                ("nodes_entered", C.c_uint64), ("tris_tested", C.c_uint64), ("seconds", C.c_double)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


@dataclass
class SceneArrays:
    """Flat scene tables: the fields of CSTriangle / CSMaterial (Structure.hpp:196-211)."""
    verts: np.ndarray    # (V,3) float32
    normals: np.ndarray  # (N,3) float32
    tris: np.ndarray     # (T,7) uint32  v0 v1 v2 n0 n1 n2 matId
    mats: np.ndarray     # (M,12) float32 Ka Kd Ks Ns Tr Ni

    MAGIC = b"MCPTSCN1"

    def save(self, path: str) -> None:
        with open(path, "wb") as f:
            f.write(self.MAGIC)
            np.array([len(self.verts), len(self.normals), len(self.tris), len(self.mats)], dtype="<u4").tofile(f)
            self.verts.astype("<f4").tofile(f)
            self.normals.astype("<f4").tofile(f)
            self.tris.astype("<u4").tofile(f)
            self.mats.astype("<f4").tofile(f)

    @classmethod
    def load(cls, path: str) -> "SceneArrays":
        with open(path, "rb") as f:
            assert f.read(8) == cls.MAGIC, "not an MCPTSCN1 file"
            nV, nN, nT, nM = np.fromfile(f, dtype="<u4", count=4).tolist()
            verts = np.fromfile(f, dtype="<f4", count=3 * nV).reshape(nV, 3)
            normals = np.fromfile(f, dtype="<f4", count=3 * nN).reshape(nN, 3)
            tris = np.fromfile(f, dtype="<u4", count=7 * nT).reshape(nT, 7)
            mats = np.fromfile(f, dtype="<f4", count=12 * nM).reshape(nM, 12)
        return cls(verts, normals, tris, mats)


@dataclass
class TreeArrays:
    """Flattened tree in the reference's BFS order (ShaderResource.hpp:128-179)."""
    left: np.ndarray
    right: np.ndarray
    aabb: np.ndarray         # (n,6)
    split_axis: np.ndarray   # 0 = leaf, 1..3
    split_value: np.ndarray
    leaf_offset: np.ndarray
    leaf_count: np.ndarray
    leaf_ids: np.ndarray
    max_depth: int
    num_leaves: int

    def same_as(self, o: "TreeArrays") -> bool:
        return (np.array_equal(self.left, o.left) and np.array_equal(self.right, o.right)
                and np.array_equal(self.aabb.view(np.uint32), o.aabb.view(np.uint32))
                and np.array_equal(self.split_axis, o.split_axis)
                and np.array_equal(self.split_value.view(np.uint32), o.split_value.view(np.uint32))
                and np.array_equal(self.leaf_offset, o.leaf_offset)
                and np.array_equal(self.leaf_count, o.leaf_count)
                and np.array_equal(self.leaf_ids, o.leaf_ids))


def _copy_tree(lib, prefix, handle) -> TreeArrays:
    cnt = (C.c_int * 4)()
    getattr(lib, prefix + "_tree_counts")(handle, cnt)
    n, nleaf, nref, depth = list(cnt)
    t = TreeArrays(np.empty(n, np.uint32), np.empty(n, np.uint32), np.empty((n, 6), np.float32),
                   np.empty(n, np.uint32), np.empty(n, np.float32), np.empty(n, np.uint32),
                   np.empty(n, np.uint32), np.empty(max(nref, 1), np.uint32), depth, nleaf)
    getattr(lib, prefix + "_tree_copy")(handle, _ptr(t.left), _ptr(t.right), _ptr(t.aabb), _ptr(t.split_axis),
                                        _ptr(t.split_value), _ptr(t.leaf_offset), _ptr(t.leaf_count),
                                        _ptr(t.leaf_ids))
    t.leaf_ids = t.leaf_ids[:nref]
    return t


# --------------------------------------------------------------------------------------
class Reference:
    """The reference's own host code (tinyobjloader 1.1.1 + KDTree.hpp) compiled in oracle/_ref."""

    def __init__(self):
        if not os.path.exists(REF_SO):
            raise FileNotFoundError(REF_SO)
        L = C.CDLL(REF_SO)
        L.ref_scene_load_obj.restype = _vp
        L.ref_scene_load_obj.argtypes = [C.c_char_p, C.c_char_p]
        L.ref_scene_from_arrays.restype = _vp
        L.ref_scene_from_arrays.argtypes = [_vp, C.c_int, _vp, C.c_int]
        L.ref_scene_free.argtypes = [_vp]
        L.ref_scene_counts.argtypes = [_vp, _vp]
        L.ref_scene_copy.argtypes = [_vp] * 5
        L.ref_scene_num_shapes.argtypes = [_vp]
        L.ref_scene_shape.argtypes = [_vp, C.c_int, C.c_char_p, C.c_int]
        L.ref_tree_build.argtypes = [_vp]
        L.ref_tree_build_seconds.argtypes = [_vp]
        L.ref_tree_build_seconds.restype = C.c_double
        L.ref_tree_counts.argtypes = [_vp, _vp]
        L.ref_tree_copy.argtypes = [_vp] * 9
        self.L = L

    @staticmethod
    def available() -> bool:
        return os.path.exists(REF_SO)

    def load_obj(self, obj_path: str, mtl_base: str | None = None):
        """QuinModel(objPath, mtlBase) -> (handle, SceneArrays, shapes[(name, ntris)])."""
        if mtl_base is None:
            mtl_base = os.path.dirname(obj_path) + "/"
        h = self.L.ref_scene_load_obj(obj_path.encode(), mtl_base.encode())
        if not h:
            raise IOError("tinyobj::LoadObj failed for " + obj_path)
        cnt = (C.c_int * 4)()
        self.L.ref_scene_counts(h, cnt)
        nV, nN, nT, nM = list(cnt)
        s = SceneArrays(np.empty((nV, 3), np.float32), np.empty((nN, 3), np.float32),
                        np.empty((nT, 7), np.uint32), np.empty((nM, 12), np.float32))
        self.L.ref_scene_copy(h, _ptr(s.verts), _ptr(s.normals), _ptr(s.tris), _ptr(s.mats))
        shapes = []
        buf = C.create_string_buffer(256)
        for i in range(self.L.ref_scene_num_shapes(h)):
            n = self.L.ref_scene_shape(h, i, buf, 256)
            shapes.append((buf.value.decode(), n))
        return h, s, shapes

    def from_arrays(self, verts: np.ndarray, tri_v: np.ndarray):
        verts = np.ascontiguousarray(verts, np.float32)
        tri_v = np.ascontiguousarray(tri_v, np.uint32)
        return self.L.ref_scene_from_arrays(_ptr(verts), len(verts), _ptr(tri_v), len(tri_v))

    def build_tree(self, handle) -> tuple[TreeArrays, float]:
        """Runs KDTree::KDTree (KDTree.hpp:58-287) and the BFS flatten."""
        self.L.ref_tree_build(handle)
        return _copy_tree(self.L, "ref", handle), self.L.ref_tree_build_seconds(handle)

    def free(self, handle):
        self.L.ref_scene_free(handle)


class ReferenceCV:
    """The CUDA tracer's own OBJ/MTL reader (CVMCTracer/.../Framework/ObjReader.cpp) and Matrix33f::det,
    compiled in oracle/_ref/libmcpt_refcv.so (ref_bridge_cv.cpp)."""

    def __init__(self):
        if not os.path.exists(REFCV_SO):
            raise FileNotFoundError(REFCV_SO)
        L = C.CDLL(REFCV_SO)
        L.refcv_load_obj.restype = _vp
        L.refcv_load_obj.argtypes = [C.c_char_p]
        L.refcv_free.argtypes = [_vp]
        L.refcv_counts.argtypes = [_vp, _vp]
        L.refcv_copy.argtypes = [_vp] * 5
        L.refcv_det33.argtypes = [_vp]
        L.refcv_det33.restype = C.c_float
        L.refcv_sample_hemi.argtypes = [C.c_ulonglong, C.c_int, _vp, _vp, _vp]
        L.refcv_sample_phong.argtypes = [C.c_ulonglong, C.c_int, _vp, _vp, _vp, _vp, _vp]
        L.refcv_sample_fresnel.argtypes = [C.c_ulonglong, C.c_int, _vp, _vp, _vp, _vp, _vp, _vp]
        L.refcv_normalize.argtypes = [_vp]
        self.L = L

    @staticmethod
    def available() -> bool:
        return os.path.exists(REFCV_SO)

    def load_obj(self, obj_path: str) -> SceneArrays:
        """readObj + the table logic of CreateGeometry (CUTracer.cu:259-311)."""
        h = self.L.refcv_load_obj(obj_path.encode())
        if not h:
            raise IOError("ObjModel::readObj failed for " + obj_path)
        cnt = (C.c_int * 4)()
        self.L.refcv_counts(h, cnt)
        nV, nN, nT, nM = list(cnt)
        s = SceneArrays(np.empty((nV, 3), np.float32), np.empty((nN, 3), np.float32),
                        np.empty((nT, 7), np.uint32), np.empty((nM, 12), np.float32))
        self.L.refcv_copy(h, _ptr(s.verts), _ptr(s.normals), _ptr(s.tris), _ptr(s.mats))
        self.L.refcv_free(h)
        return s

    # ---- CUDA/Utils.hpp:46-137, the reference's own samplers compiled on the host -------------------
    # Each returns (directions (n,3), the uniforms curand_uniform handed the sampler in each call).
    def sample_hemi(self, seed: int, normals: np.ndarray):
        normals = np.ascontiguousarray(normals, np.float32)
        n = len(normals)
        out, u = np.empty((n, 3), np.float32), np.empty((n, 2), np.float32)
        self.L.refcv_sample_hemi(seed, n, _ptr(normals), _ptr(out), _ptr(u))
        return out, u

    def sample_phong(self, seed: int, normals: np.ndarray, dirs: np.ndarray, Ns: np.ndarray):
        normals = np.ascontiguousarray(normals, np.float32)
        dirs = np.ascontiguousarray(dirs, np.float32)
        Ns = np.ascontiguousarray(Ns, np.uint32)
        n = len(normals)
        out, u = np.empty((n, 3), np.float32), np.empty((n, 2), np.float32)
        self.L.refcv_sample_phong(seed, n, _ptr(normals), _ptr(dirs), _ptr(Ns), _ptr(out), _ptr(u))
        return out, u

    def sample_fresnel(self, seed: int, normals: np.ndarray, dirs: np.ndarray, Tr: np.ndarray, Ni: np.ndarray):
        normals = np.ascontiguousarray(normals, np.float32)
        dirs = np.ascontiguousarray(dirs, np.float32)
        Tr = np.ascontiguousarray(Tr, np.float32)
        Ni = np.ascontiguousarray(Ni, np.float32)
        n = len(normals)
        out, u = np.empty((n, 3), np.float32), np.empty(n, np.float32)
        self.L.refcv_sample_fresnel(seed, n, _ptr(normals), _ptr(dirs), _ptr(Tr), _ptr(Ni), _ptr(out), _ptr(u))
        return out, u

    def normalize(self, v) -> np.ndarray:
        v = np.array(v, np.float32).reshape(3)
        self.L.refcv_normalize(_ptr(v))
        return v

    def det33(self, m) -> np.float32:
        m = np.ascontiguousarray(m, np.float32).reshape(9)
        return np.float32(self.L.refcv_det33(_ptr(m)))


# --------------------------------------------------------------------------------------
class Oracle:
    """The CPU restatement (liboracle.so)."""

    def __init__(self):
        build()
        L = C.CDLL(ORACLE_SO)
        L.oracle_default_cfg.argtypes = [C.POINTER(Cfg), C.c_int, C.c_int, C.c_int, C.c_int]
        L.oracle_scene_create.restype = _vp
        L.oracle_scene_create.argtypes = [_vp, C.c_int, _vp, C.c_int, _vp, C.c_int, _vp, C.c_int]
        L.oracle_scene_free.argtypes = [_vp]
        L.oracle_tree_build.argtypes = [_vp]
        L.oracle_tree_set.argtypes = [_vp, C.c_int] + [_vp] * 8 + [C.c_int]
        L.oracle_tree_counts.argtypes = [_vp, _vp]
        L.oracle_tree_copy.argtypes = [_vp] * 9
        L.oracle_primary_rays.argtypes = [C.POINTER(Cfg), C.c_int, _vp]
        L.oracle_intersect.argtypes = [_vp, _vp, C.c_int64, _vp, _vp, _vp, _vp, C.POINTER(Counters), C.c_int]
        L.oracle_render.argtypes = [_vp, C.POINTER(Cfg), _vp, _vp, _vp, C.POINTER(Counters)]
        L.oracle_rand_seed.restype = C.c_uint32
        L.oracle_rand_seed.argtypes = [C.c_uint32, C.c_uint32]
        L.oracle_rand_gen.restype = C.c_float
        L.oracle_rand_gen.argtypes = [C.POINTER(C.c_uint32)]
        L.oracle_frame_seeds.argtypes = [C.c_int, C.c_int, _vp]
        L.oracle_philox4x32_10.argtypes = [_vp, _vp, _vp]
        L.oracle_sample_hemi.argtypes = [C.c_float, C.c_float, _vp, _vp]
        L.oracle_sample_phong.argtypes = [C.c_float, C.c_float, _vp, _vp, C.c_float, _vp]
        L.oracle_sample_fresnel.argtypes = [C.c_float, _vp, _vp, C.c_float, C.c_float, _vp]
        L.oracle_sample_fresnel_twin.argtypes = [C.c_float, _vp, _vp, C.c_float, C.c_float, _vp]
        L.oracle_normalize_twin.argtypes = [_vp]
        L.oracle_sample_batch.argtypes = [C.c_int, C.c_int64] + [_vp] * 6
        self.L = L

    def cfg(self, width, height, spp, depth, **kw) -> Cfg:
        c = Cfg()
        self.L.oracle_default_cfg(C.byref(c), width, height, spp, depth)
        for k, v in kw.items():
            if k in ("eye", "at", "up"):
                getattr(c, k)[:] = list(v)
            else:
                setattr(c, k, v)
        return c

    def scene(self, s: SceneArrays, tree: TreeArrays | None = None) -> "OracleScene":
        return OracleScene(self, s, tree)

    def frame_seeds(self, first: int, count: int) -> np.ndarray:
        out = np.empty(count, np.uint32)
        self.L.oracle_frame_seeds(first, count, _ptr(out))
        return out

    def philox(self, ctr, key) -> np.ndarray:
        ctr = np.asarray(ctr, np.uint32)
        key = np.asarray(key, np.uint32)
        out = np.empty(4, np.uint32)
        self.L.oracle_philox4x32_10(_ptr(ctr), _ptr(key), _ptr(out))
        return out

    HEMI, PHONG, FRESNEL, FRESNEL_TWIN = 0, 1, 2, 3

    def sample_batch(self, kind: int, u, normals, dirs=None, par0=None, par1=None) -> np.ndarray:
        """The restated samplers (rtx.hlsl:213-302; kind 3: the twin's Utils.hpp:97-137) on n tuples; the uniforms
        `u` ((n,2) for hemi/phong, (n,) for fresnel) are consumed exactly as given."""
        c = lambda a: None if a is None else np.ascontiguousarray(a, np.float32)
        u, normals, dirs, par0, par1 = c(u), c(normals), c(dirs), c(par0), c(par1)
        n = len(normals)
        out = np.empty((n, 3), np.float32)
        self.L.oracle_sample_batch(kind, n, _ptr(u), _ptr(normals), _ptr(dirs), _ptr(par0), _ptr(par1), _ptr(out))
        return out

    def normalize_twin(self, v) -> np.ndarray:
        v = np.array(v, np.float32).reshape(3)
        self.L.oracle_normalize_twin(_ptr(v))
        return v

    def primary_rays(self, cfg: Cfg, frame: int = 0) -> np.ndarray:
        w = (cfg.x1 - cfg.x0) if cfg.x1 > cfg.x0 else cfg.width
        h = (cfg.y1 - cfg.y0) if cfg.y1 > cfg.y0 else cfg.height
        rays = np.empty((h * w, 6), np.float32)
        self.L.oracle_primary_rays(C.byref(cfg), frame, _ptr(rays))
        return rays


class OracleScene:
    def __init__(self, o: Oracle, s: SceneArrays, tree: TreeArrays | None):
        self.o = o
        self.arrays = s
        v = np.ascontiguousarray(s.verts, np.float32)
        n = np.ascontiguousarray(s.normals, np.float32)
        t = np.ascontiguousarray(s.tris, np.uint32)
        m = np.ascontiguousarray(s.mats, np.float32)
        self.h = o.L.oracle_scene_create(_ptr(v), len(v), _ptr(n), len(n), _ptr(t), len(t), _ptr(m), len(m))
        if tree is None:
            o.L.oracle_tree_build(self.h)
        else:
            self.set_tree(tree)

    def set_tree(self, t: TreeArrays):
        arrs = [np.ascontiguousarray(a) for a in (t.left, t.right, t.aabb, t.split_axis, t.split_value,
                                                  t.leaf_offset, t.leaf_count, t.leaf_ids)]
        self.o.L.oracle_tree_set(self.h, len(t.left), *[_ptr(a) for a in arrs], len(t.leaf_ids))

    def tree(self) -> TreeArrays:
        return _copy_tree(self.o.L, "oracle", self.h)

    def intersect(self, rays: np.ndarray, threads: int = 0):
        """rtx.hlsl:144-211 on a ray batch -> (tri int32, t, beta, gamma, counters)."""
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 6)
        n = len(rays)
        tri = np.empty(n, np.int32)
        t = np.zeros(n, np.float32)
        b = np.zeros(n, np.float32)
        g = np.zeros(n, np.float32)
        cnt = Counters()
        self.o.L.oracle_intersect(self.h, _ptr(rays), n, _ptr(tri), _ptr(t), _ptr(b), _ptr(g), C.byref(cnt), threads)
        return tri, t, b, g, cnt

    def render(self, cfg: Cfg, display: bool = False):
        """-> (sum (h,w,3) f64, sumsq (h,w,3) f64, display (h,w,4) f32 | None, counters)."""
        w = (cfg.x1 - cfg.x0) if cfg.x1 > cfg.x0 else cfg.width
        h = (cfg.y1 - cfg.y0) if cfg.y1 > cfg.y0 else cfg.height
        s = np.zeros((h, w, 3), np.float64)
        s2 = np.zeros((h, w, 3), np.float64)
        d = np.zeros((h, w, 4), np.float32) if display else None
        cnt = Counters()
        self.o.L.oracle_render(self.h, C.byref(cfg), _ptr(s), _ptr(s2), _ptr(d), C.byref(cnt))
        return s, s2, d, cnt

    def close(self):
        if self.h:
            self.o.L.oracle_scene_free(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
