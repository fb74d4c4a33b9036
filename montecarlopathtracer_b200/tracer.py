"""Host-side mirror of the reference's tracer interface over libmcpt.so's C ABI.

The reference exposes the hot path as ``PW::Tracer::{Initialize, CreateGeometry, DestroyGeometry,
RenderScene}`` (CVMCTracer/CVMCTracer/CUDA/CUTracer.h:9-12); the functions below keep those names
(snake-cased), argument meaning and call order (``main.cpp:16-18``: read scene -> create geometry
-> render).  Everything is a thin ctypes call into ``include/mcpt.h``; there is NO CPU fallback:
if the CUDA library is missing or no GPU is visible the calls raise.
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("MCPT_LIB") or os.path.join(HERE, "lib", "libmcpt.so")   # MCPT_LIB: an A/B build (build.py)

# enums of include/mcpt.h
BUILD_REFERENCE, BUILD_SAH, BUILD_GPU = 0, 1, 2
LOADER_TINYOBJ, LOADER_OBJREADER = 0, 1
RNG_PHILOX, RNG_REFERENCE = 0, 1
FLAG_NEE, FLAG_MOMENTS, FLAG_NO_JITTER, FLAG_REF_TRAVERSAL, FLAG_SIMPLE_KERNEL, FLAG_RENDER_ARITHMETIC = 1, 2, 4, 8, 16, 32
SAMPLE_HEMI, SAMPLE_PHONG, SAMPLE_FRESNEL, SAMPLE_FRESNEL_TWIN = 0, 1, 2, 3
ISECT_REFERENCE, ISECT_KD, ISECT_ROPES = 0, 1, 2
TRACER_MCRT, TRACER_CUDA_TWIN = 0, 1
ERR_NO_DEVICE = 4
ERR_NCCL = 5
UNIQUE_ID_BYTES = 128
ERR_INTERNAL = 6

_vp = C.c_void_p


class McptError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"libmcpt error {code}: {msg}")
        self.code = code


class SceneDesc(C.Structure):
    _fields_ = [("vertices", _vp), ("num_vertices", C.c_int32), ("normals", _vp), ("num_normals", C.c_int32),
                ("triangles", _vp), ("num_triangles", C.c_int32), ("materials", _vp), ("num_materials", C.c_int32),
                ("build_mode", C.c_int32)]


class RenderCfg(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("spp", C.c_int32), ("max_depth", C.c_int32),
                ("sample_offset", C.c_int32), ("seed", C.c_uint32), ("rng_mode", C.c_int32), ("flags", C.c_uint32),
                ("eye", C.c_float * 3), ("at", C.c_float * 3), ("up", C.c_float * 3), ("fovy", C.c_float),
                ("emission_scale", C.c_float), ("tracer_mode", C.c_int32)]


class Counters(C.Structure):
    _fields_ = [("paths", C.c_uint64), ("rays", C.c_uint64), ("kernel_ms", C.c_double), ("total_ms", C.c_double),
                ("launches", C.c_uint32), ("reserved", C.c_uint32)]


class TreeInfo(C.Structure):
    _fields_ = [("num_nodes", C.c_int32), ("num_leaves", C.c_int32), ("num_leaf_refs", C.c_int32),
                ("max_depth", C.c_int32), ("num_emitter_triangles", C.c_int32), ("build_seconds", C.c_double),
                ("device_bytes", C.c_uint64)]


_lib = None


def lib() -> C.CDLL:
    """Loads libmcpt.so; raises if it has not been built (``python -m montecarlopathtracer_b200.build``)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise FileNotFoundError(f"{LIB_PATH} is missing: build it with `python -m montecarlopathtracer_b200.build` "
                                "(the tracer is CUDA-only, there is no CPU path)")
    L = C.CDLL(LIB_PATH)
    L.mcpt_initialize.argtypes = [C.c_int]
    L.mcpt_create_geometry.argtypes = [C.POINTER(SceneDesc)]
    L.mcpt_load_obj.argtypes = [C.c_char_p, C.c_char_p, C.c_int, C.c_int]
    L.mcpt_load_scene_file.argtypes = [C.c_char_p, C.c_int]
    L.mcpt_get_tree_info.argtypes = [C.POINTER(TreeInfo)]
    L.mcpt_debug_stackless_walk.argtypes = [_vp, C.c_int, C.POINTER(C.c_int), _vp, _vp, _vp, _vp, C.c_int, C.POINTER(C.c_int)]
    L.mcpt_check_stackless_layout.argtypes = []
    L.mcpt_debug_compare_device_tables.argtypes = [_vp]
    L.mcpt_get_tree.argtypes = [_vp] * 8
    L.mcpt_get_scene.argtypes = [_vp] * 5
    L.mcpt_default_render_cfg.argtypes = [C.POINTER(RenderCfg), C.c_int, C.c_int, C.c_int, C.c_int]
    L.mcpt_default_render_cfg.restype = None
    L.mcpt_cuda_twin_render_cfg.argtypes = [C.POINTER(RenderCfg), C.c_int, C.c_int, C.c_int, C.c_int]
    L.mcpt_cuda_twin_render_cfg.restype = None
    L.mcpt_render_scene.argtypes = [C.POINTER(RenderCfg), _vp, _vp]
    L.mcpt_render_accumulate_device.argtypes = [C.POINTER(RenderCfg), _vp, _vp, _vp]
    L.mcpt_render_accumulate_device_noalloc.argtypes = [C.POINTER(RenderCfg), _vp, _vp]
    L.mcpt_intersect_batch.argtypes = [_vp, C.c_int64, C.c_int, _vp, _vp, _vp, _vp]
    L.mcpt_intersect_batch_device.argtypes = [_vp, C.c_int64, C.c_int, _vp, _vp, _vp, _vp, _vp]
    L.mcpt_primary_rays.argtypes = [C.POINTER(RenderCfg), C.c_int, _vp]
    L.mcpt_sample_batch.argtypes = [C.c_int, C.c_int64, _vp, _vp, _vp, _vp, _vp, C.c_int, _vp]
    L.mcpt_get_counters.argtypes = [C.POINTER(Counters)]
    L.mcpt_traversal_stats.argtypes = [_vp, C.c_int64, C.c_int, _vp]
    L.mcpt_measure_fp32_peak.argtypes = [C.POINTER(C.c_double)]
    L.mcpt_selfcheck_div3.argtypes = [C.c_int64, C.c_uint32, C.c_int, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
    L.mcpt_write_png.argtypes = [C.c_char_p, _vp, C.c_int, C.c_int, C.c_float]
    L.mcpt_write_pfm.argtypes = [C.c_char_p, _vp, C.c_int, C.c_int]
    L.mcpt_rank_share.argtypes = [C.c_int32, C.c_int32, C.c_int, C.c_int, C.POINTER(C.c_int32), C.POINTER(C.c_int32)]
    L.mcpt_rank_share.restype = None
    L.mcpt_comm_get_unique_id.argtypes = [_vp]
    L.mcpt_comm_init.argtypes = [C.c_int, C.c_int, _vp]
    L.mcpt_comm_info.argtypes = [C.POINTER(C.c_int)] * 3
    L.mcpt_render_scene_distributed.argtypes = [C.POINTER(RenderCfg), _vp]
    L.mcpt_render_accumulate_distributed_device.argtypes = [C.POINTER(RenderCfg), _vp, _vp]
    L.mcpt_set_tuning.argtypes = [C.c_char_p, C.c_double]
    L.mcpt_get_tuning.argtypes = [C.c_char_p, C.POINTER(C.c_double)]
    L.mcpt_last_error.restype = C.c_char_p
    L.mcpt_version.restype = C.c_char_p
    _lib = L
    return L


def _check(rc: int) -> None:
    if rc != 0:
        raise McptError(rc, lib().mcpt_last_error().decode(errors="replace"))


def _p(a):
    return None if a is None else a.ctypes.data_as(_vp)


# ---- PW::Tracer::Initialize ---------------------------------------------------------------
def initialize(device: int = 0) -> None:
    _check(lib().mcpt_initialize(device))


def shutdown() -> None:
    _check(lib().mcpt_shutdown())


# ---- PW::Tracer::CreateGeometry ------------------------------------------------------------
def create_geometry(vertices, normals, triangles, materials, build_mode: int = BUILD_REFERENCE) -> None:
    """Flat tables: vertices (V,3) f32, normals (N,3) f32, triangles (T,7) u32 {v0 v1 v2 n0 n1 n2 mat},
    materials (M,12) f32 {Ka Kd Ks Ns Tr Ni} -- the fields of CSTriangle / CSMaterial."""
    v = np.ascontiguousarray(vertices, np.float32).reshape(-1, 3)
    n = np.ascontiguousarray(normals, np.float32).reshape(-1, 3)
    t = np.ascontiguousarray(triangles, np.uint32).reshape(-1, 7)
    m = np.ascontiguousarray(materials, np.float32).reshape(-1, 12)
    d = SceneDesc(_p(v), len(v), _p(n), len(n), _p(t), len(t), _p(m), len(m), build_mode)
    _check(lib().mcpt_create_geometry(C.byref(d)))


def load_obj(obj_path: str, mtl_base: str | None = None, loader: int = LOADER_TINYOBJ,
             build_mode: int = BUILD_REFERENCE) -> None:
    if mtl_base is None:
        mtl_base = os.path.dirname(obj_path) + "/"
    _check(lib().mcpt_load_obj(obj_path.encode(), mtl_base.encode(), loader, build_mode))


def load_scene_file(path: str, build_mode: int = BUILD_REFERENCE) -> None:
    _check(lib().mcpt_load_scene_file(path.encode(), build_mode))


def destroy_geometry() -> None:
    _check(lib().mcpt_destroy_geometry())


def compare_device_tables() -> tuple[int, int, int]:
    """(leaf-block float4s, nodes, shading float4s) in which the device's render-path tables differ from the host pack of the same tree."""
    bad = (C.c_uint64 * 3)()
    _check(lib().mcpt_debug_compare_device_tables(bad))
    return int(bad[0]), int(bad[1]), int(bad[2])


def check_stackless_layout() -> None:
    """Host-side structural check of the leaf blocks and neighbour links built for the current geometry (no device)."""
    _check(lib().mcpt_check_stackless_layout())


def debug_stackless_walk(ray, max_hops: int = 4096):
    """Host-side diagnostic: (leaf block offsets, exit distances, exit axes, list of triangle-id lists per leaf) of the
    stackless walk of one ray {origin, direction}; no triangle is tested, nothing runs on the device."""
    ray = np.ascontiguousarray(ray, np.float32).reshape(6)
    off = np.empty(max_hops, np.uint32); te = np.empty(max_hops, np.float32); ax = np.empty(max_hops, np.uint32)
    ids = np.empty(64 * max_hops, np.uint32)
    hops, nids = C.c_int(), C.c_int()
    _check(lib().mcpt_debug_stackless_walk(_p(ray), max_hops, C.byref(hops), _p(off), _p(te), _p(ax), _p(ids), len(ids), C.byref(nids)))
    leaves, cur = [], []
    for v in ids[:nids.value]:
        if v == 0xFFFFFFFF:
            leaves.append(cur); cur = []
        else:
            cur.append(int(v))
    return off[:hops.value], te[:hops.value], ax[:hops.value], leaves


def tree_info() -> TreeInfo:
    info = TreeInfo()
    _check(lib().mcpt_get_tree_info(C.byref(info)))
    return info


@dataclass
class Tree:
    left: np.ndarray
    right: np.ndarray
    aabb: np.ndarray
    split_axis: np.ndarray
    split_value: np.ndarray
    leaf_offset: np.ndarray
    leaf_count: np.ndarray
    leaf_ids: np.ndarray


def get_tree() -> Tree:
    """The built tree in the reference's BFS layout (ShaderResource.hpp:128-179)."""
    info = tree_info()
    n, r = info.num_nodes, max(info.num_leaf_refs, 1)
    t = Tree(np.empty(n, np.uint32), np.empty(n, np.uint32), np.empty((n, 6), np.float32), np.empty(n, np.uint32),
             np.empty(n, np.float32), np.empty(n, np.uint32), np.empty(n, np.uint32), np.empty(r, np.uint32))
    _check(lib().mcpt_get_tree(_p(t.left), _p(t.right), _p(t.aabb), _p(t.split_axis), _p(t.split_value),
                               _p(t.leaf_offset), _p(t.leaf_count), _p(t.leaf_ids)))
    t.leaf_ids = t.leaf_ids[:info.num_leaf_refs]
    return t


def get_scene():
    """-> (vertices, normals, triangles, materials) as loaded."""
    cnt = (C.c_int32 * 4)()
    _check(lib().mcpt_get_scene(cnt, None, None, None, None))
    v = np.empty((cnt[0], 3), np.float32)
    n = np.empty((cnt[1], 3), np.float32)
    t = np.empty((cnt[2], 7), np.uint32)
    m = np.empty((cnt[3], 12), np.float32)
    _check(lib().mcpt_get_scene(cnt, _p(v), _p(n), _p(t), _p(m)))
    return v, n, t, m


# ---- PW::Tracer::RenderScene ---------------------------------------------------------------
def render_cfg(width: int, height: int, spp: int, max_depth: int, **kw) -> RenderCfg:
    cfg = RenderCfg()
    lib().mcpt_default_render_cfg(C.byref(cfg), width, height, spp, max_depth)
    for k, v in kw.items():
        if k in ("eye", "at", "up"):
            getattr(cfg, k)[:] = list(v)
        else:
            setattr(cfg, k, v)
    return cfg


def cuda_twin_render_cfg(width: int, height: int, spp: int, scene_id: int = 1, **kw) -> RenderCfg:
    """The cfg of PW::Tracer::RenderScene(sceneID, ...) (CUTracer.cu:340-404)."""
    cfg = RenderCfg()
    lib().mcpt_cuda_twin_render_cfg(C.byref(cfg), width, height, spp, scene_id)
    for k, v in kw.items():
        if k in ("eye", "at", "up"):
            getattr(cfg, k)[:] = list(v)
        else:
            setattr(cfg, k, v)
    return cfg


def render_scene(cfg: RenderCfg, out: np.ndarray | None = None, moments: bool = False):
    """-> hostcolor (H,W,3) float32 linear mean radiance, top row first (and E[x^2] if moments)."""
    if out is None:
        out = np.empty((cfg.height, cfg.width, 3), np.float32)
    m2 = None
    if moments:
        cfg.flags |= FLAG_MOMENTS
        m2 = np.empty((cfg.height, cfg.width, 3), np.float32)
    _check(lib().mcpt_render_scene(C.byref(cfg), _p(out), _p(m2)))
    return (out, m2) if moments else out


def render_accumulate_device(cfg: RenderCfg, dev_sum_ptr: int, dev_sumsq_ptr: int = 0, stream: int = 0) -> None:
    """Adds cfg.spp samples into a caller-owned device float4 buffer (e.g. a torch tensor's data_ptr())."""
    _check(lib().mcpt_render_accumulate_device(C.byref(cfg), _vp(dev_sum_ptr), _vp(dev_sumsq_ptr or None),
                                               _vp(stream or None)))


NOALLOC_MAX_SLICE_PIXELS = 1 << 20   # MCPT_NOALLOC_MAX_SLICE_PIXELS


def render_accumulate_device_noalloc(cfg: RenderCfg, dev_sum_ptr: int, stream: int = 0) -> None:
    """render_accumulate_device without any device allocation: fixed 16 MiB static scratch, width*height <= 2**20
    (include/mcpt.h: mcpt_render_accumulate_device_noalloc)."""
    _check(lib().mcpt_render_accumulate_device_noalloc(C.byref(cfg), _vp(dev_sum_ptr), _vp(stream or None)))


# ---- multi-GPU: spp split + one NCCL reduce behind the C ABI -------------------------------------
def rank_share(spp: int, sample_offset: int, rank: int, nranks: int) -> tuple[int, int]:
    """(sample_offset, spp) of rank `rank` when a job of `spp` samples starting at `sample_offset` is split over `nranks`."""
    off, n = C.c_int32(), C.c_int32()
    lib().mcpt_rank_share(spp, sample_offset, rank, nranks, C.byref(off), C.byref(n))
    return off.value, n.value


def comm_get_unique_id() -> bytes:
    """Rank 0: the 128-byte NCCL id the other ranks need for comm_init (hand it over by any side channel)."""
    buf = C.create_string_buffer(UNIQUE_ID_BYTES)
    _check(lib().mcpt_comm_get_unique_id(buf))
    return buf.raw


def comm_init(rank: int, nranks: int, unique_id: bytes | None = None) -> None:
    buf = C.create_string_buffer(unique_id, UNIQUE_ID_BYTES) if unique_id is not None else None
    _check(lib().mcpt_comm_init(rank, nranks, buf))


def comm_destroy() -> None:
    _check(lib().mcpt_comm_destroy())


def comm_info() -> tuple[int, int, int]:
    r, n, v = C.c_int(), C.c_int(), C.c_int()
    lib().mcpt_comm_info(C.byref(r), C.byref(n), C.byref(v))
    return r.value, n.value, v.value


def render_scene_distributed(whole_job: RenderCfg, out: np.ndarray | None = None):
    """All ranks call this with the cfg of the WHOLE job; rank 0 gets hostcolor (H,W,3), the others None."""
    rank = comm_info()[0]
    if rank == 0 and out is None:
        out = np.empty((whole_job.height, whole_job.width, 3), np.float32)
    _check(lib().mcpt_render_scene_distributed(C.byref(whole_job), _p(out) if rank == 0 else None))
    return out if rank == 0 else None


def render_accumulate_distributed_device(whole_job: RenderCfg, dev_sum_ptr: int, stream: int = 0) -> None:
    _check(lib().mcpt_render_accumulate_distributed_device(C.byref(whole_job), _vp(dev_sum_ptr), _vp(stream or None)))


def intersect_batch(rays: np.ndarray, variant: int = ISECT_REFERENCE):
    """rays (n,6) -> (tri int32 [-1 = miss], t, beta, gamma)."""
    rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 6)
    n = len(rays)
    tri = np.empty(n, np.int32)
    t = np.empty(n, np.float32)
    b = np.empty(n, np.float32)
    g = np.empty(n, np.float32)
    _check(lib().mcpt_intersect_batch(_p(rays), n, variant, _p(tri), _p(t), _p(b), _p(g)))
    return tri, t, b, g


def intersect_batch_device(rays_ptr: int, n: int, variant: int, tri_ptr: int, t_ptr: int = 0, beta_ptr: int = 0,
                           gamma_ptr: int = 0, stream: int = 0) -> None:
    _check(lib().mcpt_intersect_batch_device(_vp(rays_ptr), n, variant, _vp(tri_ptr), _vp(t_ptr or None),
                                             _vp(beta_ptr or None), _vp(gamma_ptr or None), _vp(stream or None)))


def primary_rays(cfg: RenderCfg, sample: int = 0) -> np.ndarray:
    rays = np.empty((cfg.height * cfg.width, 6), np.float32)
    _check(lib().mcpt_primary_rays(C.byref(cfg), sample, _p(rays)))
    return rays


def sample_batch(kind: int, u, normals, dirs=None, par0=None, par1=None, render_arithmetic: bool = False) -> np.ndarray:
    """The device BSDF samplers (rtx.hlsl:213-302; SAMPLE_FRESNEL_TWIN: Utils.hpp:97-137) on n tuples.
    u (n,2) uniforms used as given; render_arithmetic: the hardware-unit forms the render kernels run."""
    c = lambda a: None if a is None else np.ascontiguousarray(a, np.float32)
    normals = c(normals).reshape(-1, 3)
    n = len(normals)
    u = c(u)
    if u.ndim == 1:   # fresnel draws one number
        u = np.ascontiguousarray(np.stack([u, u], 1))
    out = np.empty((n, 3), np.float32)
    _check(lib().mcpt_sample_batch(kind, n, _p(u), _p(normals), _p(c(dirs)), _p(c(par0)), _p(c(par1)), int(render_arithmetic), _p(out)))
    return out


def counters() -> Counters:
    c = Counters()
    _check(lib().mcpt_get_counters(C.byref(c)))
    return c


def traversal_stats(rays: np.ndarray, variant: int = ISECT_KD):
    """-> (nodes popped, leaves/boxes entered, triangle tests) summed over the batch."""
    rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 6)
    out = np.zeros(3, np.uint64)
    _check(lib().mcpt_traversal_stats(_p(rays), len(rays), variant, _p(out)))
    return tuple(int(x) for x in out)


def measure_fp32_peak() -> float:
    """Measured FP32 FMA rate of the bound device, GFLOP/s."""
    g = C.c_double()
    _check(lib().mcpt_measure_fp32_peak(C.byref(g)))
    return g.value


def selfcheck_div3(n: int, seed: int = 1, spread: int = 30) -> tuple:
    """div3_rn (the kernels' shared-reciprocal division) against __fdiv_rn on n random quadruples with exponents
    in 2^-spread..2^spread -> (mismatching quotients, quadruples that took the shared-reciprocal branch)."""
    bad, fast = C.c_uint64(), C.c_uint64()
    _check(lib().mcpt_selfcheck_div3(n, seed, spread, C.byref(bad), C.byref(fast)))
    return bad.value, fast.value


def write_png(path: str, rgb: np.ndarray, gamma: float = 2.2) -> None:
    """8-bit PNG of clamp(rgb ** (1/gamma)); gamma 2.2 is the reference's display value (rtx.hlsl:402)."""
    rgb = np.ascontiguousarray(rgb, np.float32)
    h, w, _ = rgb.shape
    _check(lib().mcpt_write_png(path.encode(), _p(rgb), w, h, gamma))


def write_pfm(path: str, rgb: np.ndarray) -> None:
    rgb = np.ascontiguousarray(rgb, np.float32)
    h, w, _ = rgb.shape
    _check(lib().mcpt_write_pfm(path.encode(), _p(rgb), w, h))


def set_tuning(name: str, value: float) -> None:
    """A/B and tuning knobs of include/mcpt.h (clamped by the library; "defaults" restores the shipped values).  The
    library never reads the environment: scripts that sweep knobs call this explicitly."""
    _check(lib().mcpt_set_tuning(name.encode(), float(value)))


def get_tuning(name: str) -> float:
    v = C.c_double()
    _check(lib().mcpt_get_tuning(name.encode(), C.byref(v)))
    return v.value


_ENV_KNOBS = {"MCPT_SM_SHADE": "shade_threshold", "MCPT_SM_CAMR": "cam_threshold", "MCPT_SM_TRICAP": "tri_cap", "MCPT_CHUNKS": "chunks",
              "MCPT_SM_MINBLOCKS": "min_blocks", "MCPT_SM_VARIANT": "variant", "MCPT_SAH_CI": "sah_ci", "MCPT_SAH_EB": "sah_eb",
              "MCPT_LEAF_BFS": "leaf_bfs", "MCPT_SAH_RESORT": "sah_resort", "MCPT_BUILD_TIMING": "build_timing", "MCPT_SM_STATS": "sm_stats", "MCPT_POOL": "pool"}


def tuning_from_env() -> dict:
    """For the sweep scripts under tools/: forwards the round-1 MCPT_* environment knobs to set_tuning.  Called by
    those scripts only -- never by the library, the tests or the default bench."""
    applied = {}
    for env, knob in _ENV_KNOBS.items():
        v = os.environ.get(env)
        if v:
            set_tuning(knob, float(v))
            applied[knob] = get_tuning(knob)
    return applied


def version() -> str:
    return lib().mcpt_version().decode()
