"""CPU tests of the product's OBJ/MTL ingest (csrc/obj_loader.cpp) against the reference's OWN readers
compiled in oracle/_ref: tinyobjloader 1.1.1 as QuinModel calls it, and PW::FileReader::ObjModel."""
import os

import numpy as np
import pytest

TRICKY_OBJ = """# quads, a concave pentagon, an n-gon, relative indices, v//n and v/t/n corners, CRLF, usemtl changes
mtllib tricky.mtl
v 0 0 0
v 1 0 0
v 1 1 0
v 0 1 0
v 0.5 0.25 0.0
v 2.0e0 0 1.5E-1
v 2 1 -0.125
v 3.14159265358979 2.718281828 -1.0000001
v +0.333333333333 1e-3 12345.6789
v -7.25 .5 0.1234567891234
vt 0 0
vt 1 1
vn 0 0 1
vn 0 1 0
vn 0.577350269 0.577350269 0.577350269
g first
usemtl red
f 1//1 2//1 3//1 4//1
f 1/1/1 2/2/1 5/1/1 3/2/2 4/1/3
usemtl glass
f -10//-3 -9//-3 -4//-2
g second
f 2//2 6//2 7//2 3//2\r
usemtl missing_name
f 6//3 8//3 9//3 10//3 7//3
o third
usemtl red
f 1//1 3//1 2//1
s off
f 4//1 3//1 2//1 1//1
"""

TRICKY_MTL = """newmtl red
Ka 0 0 0
Kd 0.75 0.25 0.25   
Ks 0.1 0.1 0.1
illum 2

newmtl glass
Kd 0 0 0
Ns 96.078431
Tr 0.9
Ni 1.5
newmtl light
Ka 0.8 0.8 0.8
d 0.25
Tr 0.5
"""


@pytest.fixture()
def tricky(tmp_path):
    (tmp_path / "tricky.obj").write_bytes(TRICKY_OBJ.encode())
    (tmp_path / "tricky.mtl").write_bytes(TRICKY_MTL.encode())
    return str(tmp_path / "tricky.obj")


def _same(a, b):
    for x, y in ((a.verts, b[0]), (a.normals, b[1]), (a.tris, b[2]), (a.mats, b[3])):
        assert x.shape == y.shape, (x.shape, y.shape)
        assert np.array_equal(np.ascontiguousarray(x).view(np.uint32), np.ascontiguousarray(y).view(np.uint32))


def test_tinyobj_convention_matches_reference_tinyobj_on_tricky_file(tracer, reference, tricky):
    h, s, shapes = reference.load_obj(tricky)
    reference.free(h)
    tracer.load_obj(tricky, loader=tracer.LOADER_TINYOBJ)
    v, n, t, m = tracer.get_scene()
    # faces without a known material carry id -1 in tinyobj; the product appends a default material for them
    unknown = s.tris[:, 6] == 0xFFFFFFFF
    assert unknown.any()
    assert (t[unknown, 6] == len(s.mats)).all() and len(m) == len(s.mats) + 1
    t2 = t.copy()
    t2[unknown, 6] = 0xFFFFFFFF
    _same(s, (v, n, t2, m[:len(s.mats)]))
    assert len(t) == 2 + 3 + 1 + 2 + 3 + 1 + 2
    tracer.destroy_geometry()


def test_objreader_convention_matches_reference_objreader_on_tricky_file(tracer, tmp_path):
    from oracle.pyoracle import ReferenceCV
    if not ReferenceCV.available():
        pytest.skip("oracle/_ref/libmcpt_refcv.so not built")
    # this reader has no relative indices and no `o`; keep what it supports, add a continuation line
    obj = TRICKY_OBJ.replace("f -10//-3 -9//-3 -4//-2", "f 1//1 2//1 \\\n 7//2").replace("\r", "")
    (tmp_path / "t.obj").write_text(obj.replace("tricky.mtl", "t.mtl"))
    (tmp_path / "t.mtl").write_text(TRICKY_MTL)
    s = ReferenceCV().load_obj(str(tmp_path / "t.obj"))
    tracer.load_obj(str(tmp_path / "t.obj"), loader=tracer.LOADER_OBJREADER)
    _same(s, tracer.get_scene())
    v, n, t, m = tracer.get_scene()
    assert (v[0] == 0).all() and t[:, :3].min() >= 1          # 1-based indices, dummy slot 0
    assert m[1, 9] == 2.0                                     # `Ks` forces Ns = 2 (ObjReader.cpp:232)
    assert len(np.unique(t[:6, 6])) == 1 and t[6, 6] != t[0, 6]                      # group `first`: every triangle takes the first one's material
    tracer.destroy_geometry()


REF_SCENES = [("scene01", "/root/reference/MCRT/QuinEngine/Res/scene01.obj"),
              ("scene02", "/root/reference/CVMCTracer/CVMCTracer/Resources/scene02.obj"),
              ("scene03", "/root/reference/CVMCTracer/CVMCTracer/Resources/scene03.obj")]


@pytest.mark.parametrize("name,path", REF_SCENES)
def test_tinyobj_convention_reproduces_the_baked_reference_scenes(tracer, scenes, name, path):
    """scenes/*.mcpt were written by the reference's own tinyobj (tests/golden/bake_scenes.py)."""
    if not os.path.exists(path):
        pytest.skip("reference not mounted")
    tracer.load_obj(path, loader=tracer.LOADER_TINYOBJ)
    _same(scenes[name], tracer.get_scene())
    tracer.destroy_geometry()


@pytest.mark.parametrize("name,path", REF_SCENES)
def test_objreader_convention_reproduces_the_reference_reader_on_its_scenes(tracer, name, path):
    from oracle.pyoracle import ReferenceCV
    if not os.path.exists(path) or not ReferenceCV.available():
        pytest.skip("reference not mounted")
    s = ReferenceCV().load_obj(path)
    tracer.load_obj(path, loader=tracer.LOADER_OBJREADER)
    _same(s, tracer.get_scene())
    tracer.destroy_geometry()


def test_loader_errors(tracer, tmp_path):
    with pytest.raises(tracer.McptError) as e:
        tracer.load_obj(str(tmp_path / "nope.obj"))
    assert e.value.code == 2
    (tmp_path / "zero.obj").write_text("v 0 0 0\nv 1 0 0\nv 0 1 0\nf 0 1 2\n")
    with pytest.raises(tracer.McptError, match="bad face index"):
        tracer.load_obj(str(tmp_path / "zero.obj"))
    (tmp_path / "empty.obj").write_text("v 0 0 0\n")
    with pytest.raises(tracer.McptError, match="no faces"):
        tracer.load_obj(str(tmp_path / "empty.obj"))
    # normals missing -> geometric normal synthesised, default material appended
    (tmp_path / "bare.obj").write_text("v 0 0 0\nv 1 0 0\nv 0 1 0\nf 1 2 3\n")
    tracer.load_obj(str(tmp_path / "bare.obj"))
    v, n, t, m = tracer.get_scene()
    assert np.allclose(n[t[0, 3]], [0, 0, 1]) and len(m) == 1
    tracer.destroy_geometry()
