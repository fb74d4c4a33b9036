"""The drop-in, compiled (VERDICT r1 missing #5/#6).

oracle/_ref/dropin_main = tests/dropin/CUTracerMcpt.cpp (PW::Tracer's four functions over libmcpt.so) compiled against
the reference's REAL CUDA/CUTracer.h and Framework/ObjReader.cpp and driven by the call sequence of the reference's own
main (CVMCTracer/CVMCTracer/main.cpp:16-18: readObj -> CreateGeometry -> RenderScene) -- `make -C oracle dropin`.
oracle/_ref/cutracer_ref = the same main linked with the reference's OWN CUDA/CUTracer.cu, compiled unmodified for sm_100a
(`make -C oracle refcu`); its 800x600 x 10 000 spp images, run once on a B200 by tools/reference_twin.py, are frozen as 4x4
tile means in tests/golden/twin_reference_scene*.npz: the statistical witness of MCPT_TRACER_CUDA_TWIN produced by the
reference's own sampleMC (its cuRAND seeds come from std::random_device, CUTracer.cu:375-376: no stream to reproduce)."""
import os
import subprocess
import sys

import numpy as np
import pytest

from conftest import ROOT

sys.path.insert(0, os.path.join(ROOT, "tests", "dropin"))
from objwriter import write_obj  # noqa: E402

DROPIN = os.path.join(ROOT, "oracle", "_ref", "dropin_main")
GOLD = os.path.join(ROOT, "tests", "golden")


def _load_f32(path):
    with open(path, "rb") as f:
        w, h = np.fromfile(f, "<u4", 2)
        return np.fromfile(f, "<f4").reshape(h, w, 3)


def test_obj_writer_round_trips_through_the_objreader_convention(tracer, scenes, tmp_path):
    """The OBJ the GPU tests hand to the reference's reader carries the repo's scene tables unchanged."""
    T, s = tracer, scenes["scene01"]
    obj = str(tmp_path / "scene01.obj")
    write_obj(obj, s.verts, s.normals, s.tris, s.mats)
    T.load_obj(obj, loader=T.LOADER_OBJREADER, build_mode=T.BUILD_SAH)
    v, n, t, m = T.get_scene()
    off_v, off_n, off_m = len(v) - len(s.verts), len(n) - len(s.normals), len(m) - len(s.mats)   # the reader's dummy slots
    assert np.array_equal(v[off_v:], s.verts) and np.array_equal(n[off_n:], s.normals)
    assert np.array_equal(t[:, :3], s.tris[:, :3] + off_v) and np.array_equal(t[:, 3:6], s.tris[:, 3:6] + off_n)
    assert np.array_equal(m[t[:, 6]], s.mats[s.tris[:, 6]])
    T.destroy_geometry()


def test_dropin_binary_links_and_refuses_without_a_gpu(scenes, tmp_path):
    import torch
    if not os.path.exists(DROPIN):
        pytest.skip("oracle/_ref/dropin_main not built (needs /root/reference at build time: make -C oracle dropin)")
    if torch.cuda.is_available():
        pytest.skip("a GPU is visible here")
    s = scenes["scene01"]
    obj = str(tmp_path / "scene01.obj")
    write_obj(obj, s.verts, s.normals, s.tris, s.mats)
    r = subprocess.run([DROPIN, obj, "1", str(tmp_path / "out.f32")], capture_output=True, text=True, timeout=120)
    assert r.returncode == 1 and "CreateGeometry failed" in r.stderr      # no device -> the shim reports, nothing falls back to a CPU path


@pytest.mark.gpu
@pytest.mark.parametrize("scene_id,name", [(1, "scene01"), (2, "scene02")])
def test_dropin_main_renders_the_reference_workload(gpu, scenes, tmp_path, scene_id, name):
    """main.cpp's sequence through the compiled shim: 800x600, 10 000 spp, depth 7 + 1 (stdafx.h:41-45) -- the workload of
    MC.docx's table.  (i) its hostcolor equals tracer.py's render of the same file bit for bit (same library, same cfg);
    (ii) it agrees statistically with the image the reference's own CUDA/CUTracer.cu rendered on a B200."""
    T, s = gpu, scenes[name]
    if not os.path.exists(DROPIN):
        pytest.skip("oracle/_ref/dropin_main not built")
    obj = str(tmp_path / (name + ".obj"))
    write_obj(obj, s.verts, s.normals, s.tris, s.mats)
    out = str(tmp_path / "hostcolor.f32")
    r = subprocess.run([DROPIN, obj, str(scene_id), out], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr
    assert r.stdout.split()[:3] == ["800", "x", "600,"] and "10000 spp" in r.stdout
    img = _load_f32(out)
    assert img.shape == (600, 800, 3) and np.isfinite(img).all() and (img >= 0).all() and img.max() > 0
    T.load_obj(obj, loader=T.LOADER_OBJREADER, build_mode=T.BUILD_SAH)
    mine = T.render_scene(T.cuda_twin_render_cfg(800, 600, 10000, scene_id))
    assert np.array_equal(mine, img)
    fx = os.path.join(GOLD, f"twin_reference_scene{scene_id}.npz")
    if not os.path.exists(fx):
        pytest.skip("tests/golden/twin_reference_scene*.npz not generated (tools/reference_twin.py on a GPU box)")
    # Two independent 10 000-spp estimates (the reference's cuRAND stream is seeded from random_device): the image means agreed
    # to 1e-4 when the fixture was made; a 40x40-pixel block mean carries ~0.3 % of Monte Carlo noise from each side.
    ref = np.load(fx)["tile_mean"].astype(np.float64)
    got = img.astype(np.float64).reshape(150, 4, 200, 4, 3).mean((1, 3))
    assert abs(got.mean() - ref.mean()) / ref.mean() < 0.002                # measured 8e-5 / 1.1e-4; north_star's bar is 1 %
    blk = lambda a: a.reshape(15, 10, 20, 10, 3).mean((1, 3))
    bo, bg = blk(ref), blk(got)
    bright = bo > 0.25 * ref.mean()
    rel = np.abs(bg - bo)[bright] / bo[bright]
    assert rel.mean() < 0.01 and rel.max() < 0.04, (rel.mean(), rel.max())
