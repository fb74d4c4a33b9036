"""CPU tests of the product's host side: the C-ABI library loads and exports every declared symbol,
the KD builder reproduces the oracle's (= the reference's) tree, and GPU entry points fail loudly
without a device.  No compute call is made here."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from conftest import ROOT, heightfield


def test_abi_exports_every_declared_symbol(tracer):
    hdr = open(os.path.join(ROOT, "include", "mcpt.h")).read()
    names = set(re.findall(r"\b(mcpt_[a-z_0-9]+)\s*\(", hdr))
    assert len(names) >= 18
    L = tracer.lib()
    for n in sorted(names):
        assert hasattr(L, n), f"libmcpt.so does not export {n}"
    assert "sm_100a" in tracer.version()


def test_struct_layouts_match_header(tracer):
    # sizes the C compiler gives the structs of include/mcpt.h (checked with a tiny C program)
    import subprocess, tempfile
    src = '#include <stdio.h>\n#include "mcpt.h"\nint main(){printf("%zu %zu %zu %zu",sizeof(mcpt_scene_desc),sizeof(mcpt_render_cfg),sizeof(mcpt_counters),sizeof(mcpt_tree_info));return 0;}'
    with tempfile.TemporaryDirectory() as d:
        open(os.path.join(d, "s.c"), "w").write(src)
        subprocess.check_call(["/usr/bin/gcc", "-I" + os.path.join(ROOT, "include"), os.path.join(d, "s.c"), "-o", os.path.join(d, "s")])
        sizes = list(map(int, subprocess.check_output([os.path.join(d, "s")]).split()))
    assert sizes == [C.sizeof(tracer.SceneDesc), C.sizeof(tracer.RenderCfg), C.sizeof(tracer.Counters), C.sizeof(tracer.TreeInfo)]


@pytest.mark.parametrize("name", ["scene01", "scene01_lambert", "scene02", "scene03"])
def test_host_builder_reproduces_reference_tree(tracer, oracle, scenes, name):
    s = scenes[name]
    tracer.create_geometry(s.verts, s.normals, s.tris, s.mats, tracer.BUILD_REFERENCE)
    t, o = tracer.get_tree(), oracle.scene(s).tree()
    assert np.array_equal(t.left, o.left) and np.array_equal(t.right, o.right)
    assert np.array_equal(t.aabb.view(np.uint32), o.aabb.view(np.uint32))
    assert np.array_equal(t.split_axis, o.split_axis)
    assert np.array_equal(t.split_value.view(np.uint32), o.split_value.view(np.uint32))
    assert np.array_equal(t.leaf_offset, o.leaf_offset) and np.array_equal(t.leaf_count, o.leaf_count)
    assert np.array_equal(t.leaf_ids, o.leaf_ids)
    info = tracer.tree_info()
    assert info.num_nodes == len(o.left) and info.max_depth == o.max_depth
    v, n, tr, m = tracer.get_scene()
    assert np.array_equal(tr, s.tris) and np.array_equal(v, s.verts)
    tracer.destroy_geometry()


def test_host_builder_on_tessellated_surface(tracer, oracle):
    from oracle.pyoracle import SceneArrays
    verts, tri_v = heightfield(48)
    tris = np.zeros((len(tri_v), 7), np.uint32)
    tris[:, :3] = tri_v
    s = SceneArrays(verts, np.array([[0, 1, 0]], np.float32), tris, np.zeros((1, 12), np.float32))
    tracer.create_geometry(s.verts, s.normals, s.tris, s.mats)
    t, o = tracer.get_tree(), oracle.scene(s).tree()
    assert np.array_equal(t.left, o.left) and np.array_equal(t.leaf_ids, o.leaf_ids)
    assert np.array_equal(t.aabb.view(np.uint32), o.aabb.view(np.uint32))
    tracer.destroy_geometry()


@pytest.mark.parametrize("build_mode", [0, 1], ids=["reference_rule", "sah"])
@pytest.mark.parametrize("name", ["scene01", "scene01_lambert", "scene02", "scene03", "heightfield"])
def test_stackless_layout_is_consistent(tracer, scenes, name, build_mode):
    """Leaf blocks, leaf handles and neighbour links of the render path's tables (csrc/kd_ropes.cpp), checked on the host for
    both builders: blocks contiguous in leaf order, handles = (offset, min(count, 7)), records = the leaf's triangle records,
    and every rope touches and covers the face it belongs to (or the face lies on the scene box)."""
    if name == "heightfield":
        from montecarlopathtracer_b200 import scenes as gen
        s0 = scenes["scene01_lambert"]
        v, n, t, m = gen.cornell_heightfield((s0.verts, s0.normals, s0.tris, s0.mats), 60)
    else:
        s = scenes[name]
        v, n, t, m = s.verts, s.normals, s.tris, s.mats
    tracer.create_geometry(v, n, t, m, build_mode)
    tracer.check_stackless_layout()
    tracer.destroy_geometry()
    with pytest.raises(tracer.McptError):
        tracer.check_stackless_layout()  # no geometry


SAH_TREES = {"scene01": (1993, "bc608806e4beca86"), "scene01_lambert": (1993, "bc608806e4beca86"),
             "scene02": (4041, "1d590ad416dbdf81"), "scene03": (12233, "d73efaab7c56034d")}


@pytest.mark.parametrize("name", sorted(SAH_TREES))
def test_sah_tree_is_frozen(tracer, scenes, name):
    """The render path's own SAH tree (csrc/kd_build_sah.cpp), flattened: frozen so that work on the builder's speed
    (radix passes, clipping shortcuts, task structure) cannot change the tree, which every measured number depends on."""
    import hashlib
    s = scenes[name]
    tracer.create_geometry(s.verts, s.normals, s.tris, s.mats, tracer.BUILD_SAH)
    t = tracer.get_tree()
    h = hashlib.sha1()
    for a in (t.left, t.right, t.aabb, t.split_axis, t.split_value, t.leaf_offset, t.leaf_count, t.leaf_ids):
        h.update(np.ascontiguousarray(a).tobytes())
    assert (len(t.left), h.hexdigest()[:16]) == SAH_TREES[name]
    tracer.destroy_geometry()


def test_scene_file_round_trip(tracer, scenes, tmp_path):
    p = str(tmp_path / "s.mcpt")
    scenes["scene02"].save(p)
    tracer.load_scene_file(p)
    v, n, t, m = tracer.get_scene()
    assert np.array_equal(t, scenes["scene02"].tris) and np.array_equal(m, scenes["scene02"].mats)
    assert tracer.tree_info().num_emitter_triangles == 4 * 420   # the four spherical lights
    tracer.destroy_geometry()


def test_errors_are_reported_not_thrown_across_the_abi(tracer, scenes, tmp_path):
    s = scenes["scene01"]
    with pytest.raises(tracer.McptError) as e:
        tracer.create_geometry(s.verts, s.normals, np.zeros((0, 7), np.uint32), s.mats)
    assert e.value.code == 1
    bad = s.tris.copy()
    bad[5, 1] = 10 ** 6
    with pytest.raises(tracer.McptError, match="vertex index out of range"):
        tracer.create_geometry(s.verts, s.normals, bad, s.mats)
    with pytest.raises(tracer.McptError) as e:
        tracer.load_scene_file(str(tmp_path / "missing.mcpt"))
    assert e.value.code == 2
    (tmp_path / "junk.mcpt").write_bytes(b"not a scene")
    with pytest.raises(tracer.McptError):
        tracer.load_scene_file(str(tmp_path / "junk.mcpt"))
    with pytest.raises(tracer.McptError, match="no geometry"):
        tracer.get_tree()


def test_hostile_scene_file_headers_are_refused_before_anything_is_sized(tracer, scenes, tmp_path):
    """The counts of a .mcpt header come from an untrusted file: 7 * 2^32 uint32 of triangles must come
    back as MCPT_ERR_IO, not as a bad_alloc / length_error that terminates the host application."""
    import struct
    p = tmp_path / "huge.mcpt"
    p.write_bytes(b"MCPTSCN1" + struct.pack("<4I", 0xFFFFFFFF, 0xFFFFFFFF, 0xFFFFFFFF, 0xFFFFFFFF) + b"\0" * 64)
    with pytest.raises(tracer.McptError, match="header counts") as e:
        tracer.load_scene_file(str(p))
    assert e.value.code == 2
    good = tmp_path / "good.mcpt"
    scenes["scene01"].save(str(good))
    raw = good.read_bytes()
    (tmp_path / "short.mcpt").write_bytes(raw[:-100])          # truncated
    (tmp_path / "long.mcpt").write_bytes(raw + b"\0" * 4)      # trailing bytes
    for name in ("short.mcpt", "long.mcpt"):
        with pytest.raises(tracer.McptError, match="header counts"):
            tracer.load_scene_file(str(tmp_path / name))


def test_failed_load_keeps_the_previous_geometry(tracer, scenes):
    """install_scene commits only on success: after a refused scene the context still answers with the
    previous scene's host tables AND tree, never a mix."""
    s = scenes["scene02"]
    tracer.create_geometry(s.verts, s.normals, s.tris, s.mats, tracer.BUILD_SAH)
    before = (tracer.tree_info().num_nodes, tracer.get_scene()[2].copy())
    bad = scenes["scene01"].tris.copy()
    bad[0, 6] = 999
    with pytest.raises(tracer.McptError, match="material index out of range"):
        tracer.create_geometry(scenes["scene01"].verts, scenes["scene01"].normals, bad, scenes["scene01"].mats)
    assert tracer.tree_info().num_nodes == before[0] and np.array_equal(tracer.get_scene()[2], before[1])
    tracer.check_stackless_layout()
    tracer.destroy_geometry()


def test_tuning_knobs_are_explicit_clamped_and_not_environmental(tracer, scenes, monkeypatch):
    """The shipped library reads no tuning variable from the caller's environment; knobs go through
    mcpt_set_tuning and are clamped: a vote threshold above the warp size would make render_kernel_sm spin for ever."""
    T = tracer
    monkeypatch.setenv("MCPT_SM_SHADE", "99")
    monkeypatch.setenv("MCPT_SAH_CI", "5.0")
    monkeypatch.setenv("MCPT_CHUNKS", "3")
    s = scenes["scene01"]
    T.set_tuning("defaults", 0)
    T.create_geometry(s.verts, s.normals, s.tris, s.mats, T.BUILD_SAH)
    nodes = T.tree_info().num_nodes
    assert (T.get_tuning("shade_threshold"), T.get_tuning("sah_ci"), T.get_tuning("chunks")) == (18, 1.0, 0)
    for knob, value, want in (("shade_threshold", 99, 32), ("shade_threshold", -4, 1), ("cam_threshold", 1000, 32), ("tri_cap", 0, 1),
                              ("chunks", 1 << 20, 64), ("sah_ci", 0.0, 0.05)):
        T.set_tuning(knob, value)
        assert T.get_tuning(knob) == pytest.approx(want)
    with pytest.raises(T.McptError, match="unknown tuning knob"):
        T.set_tuning("no_such_knob", 1)
    with pytest.raises(T.McptError):
        T.set_tuning("chunks", float("nan"))
    # a knob that changes the tree takes effect at the next build and "defaults" restores the shipped tree
    T.set_tuning("sah_ci", 5.0)
    T.create_geometry(s.verts, s.normals, s.tris, s.mats, T.BUILD_SAH)
    assert T.tree_info().num_nodes != nodes
    T.set_tuning("defaults", 0)
    T.create_geometry(s.verts, s.normals, s.tris, s.mats, T.BUILD_SAH)
    assert T.tree_info().num_nodes == nodes
    T.destroy_geometry()
    src = "".join(open(os.path.join(ROOT, "montecarlopathtracer_b200", "csrc", f)).read()
                  for f in os.listdir(os.path.join(ROOT, "montecarlopathtracer_b200", "csrc")) if f.endswith((".cpp", ".cu", ".cuh", ".h")))
    assert "getenv" not in src


def test_no_cpu_fallback(tracer, scenes):
    """Without a CUDA device every compute entry point must refuse: the product has no CPU path."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is visible here")
    s = scenes["scene01"]
    tracer.create_geometry(s.verts, s.normals, s.tris, s.mats)
    with pytest.raises(tracer.McptError) as e:
        tracer.initialize(0)
    assert e.value.code == tracer.ERR_NO_DEVICE
    cfg = tracer.render_cfg(16, 16, 1, 5)
    for call in (lambda: tracer.render_scene(cfg), lambda: tracer.intersect_batch(np.zeros((4, 6), np.float32)),
                 lambda: tracer.primary_rays(cfg)):
        with pytest.raises(tracer.McptError) as e:
            call()
        assert e.value.code == tracer.ERR_NO_DEVICE
    tracer.destroy_geometry()


def test_image_writers(tracer, tmp_path):
    """PNG = 8-bit clamp(v^(1/2.2)) (rtx.hlsl:402), decodable by any reader; PFM = raw floats, bottom row first."""
    import struct, zlib
    g = np.random.default_rng(0)
    img = g.uniform(-0.1, 1.3, (37, 53, 3)).astype(np.float32)
    p = str(tmp_path / "a.png")
    tracer.write_png(p, img)
    raw = open(p, "rb").read()
    assert raw[:8] == b"\x89PNG\r\n\x1a\n"
    pos, idat, hdr = 8, b"", None
    while pos < len(raw):
        n, typ = struct.unpack(">I4s", raw[pos:pos + 8])
        body = raw[pos + 8:pos + 8 + n]
        assert zlib.crc32(typ + body) == struct.unpack(">I", raw[pos + 8 + n:pos + 12 + n])[0]
        if typ == b"IHDR":
            hdr = struct.unpack(">IIBBBBB", body)
        if typ == b"IDAT":
            idat += body
        pos += 12 + n
    assert hdr == (53, 37, 8, 2, 0, 0, 0)
    px = np.frombuffer(zlib.decompress(idat), np.uint8).reshape(37, 1 + 53 * 3)
    assert (px[:, 0] == 0).all()
    want = np.floor(np.clip(np.maximum(img, 0) ** (1 / 2.2), 0, 1) * 255 + 0.5)
    assert np.abs(px[:, 1:].reshape(37, 53, 3).astype(int) - want.astype(int)).max() <= 1
    q = str(tmp_path / "a.pfm")
    tracer.write_pfm(q, img)
    data = open(q, "rb").read()
    head = b"PF\n53 37\n-1.0\n"
    assert data.startswith(head)
    assert np.array_equal(np.frombuffer(data[len(head):], np.float32).reshape(37, 53, 3)[::-1], img)
    with pytest.raises(tracer.McptError):
        tracer.write_png(str(tmp_path / "no_such_dir" / "x.png"), img)
