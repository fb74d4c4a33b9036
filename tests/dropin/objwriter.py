"""Writes flat scene tables (the fields of mcpt_scene_desc) as an OBJ + MTL pair that the reference's own reader
(CVMCTracer/CVMCTracer/Framework/ObjReader.cpp) parses: one `g` group per run of triangles with the same material
(CUTracer.cu:300-308 binds ONE material per group, that of its first triangle), `f v//n` faces, 1-based indices."""
import os


def write_obj(path: str, verts, normals, tris, mats) -> None:
    base = os.path.splitext(os.path.basename(path))[0]
    with open(os.path.splitext(path)[0] + ".mtl", "w") as f:
        for i, m in enumerate(mats):
            f.write(f"newmtl m{i}\n")
            f.write("Ka %.9g %.9g %.9g\nKd %.9g %.9g %.9g\nKs %.9g %.9g %.9g\n" % tuple(float(x) for x in m[:9]))
            f.write("Ns %.9g\nTr %.9g\nNi %.9g\n" % (float(m[9]), float(m[10]), float(m[11])))   # Ns AFTER Ks: `Ks` forces Ns = 2 (ObjReader.cpp:225-233)
    with open(path, "w") as f:
        f.write(f"mtllib {base}.mtl\n")
        for v in verts:
            f.write("v %.9g %.9g %.9g\n" % (float(v[0]), float(v[1]), float(v[2])))
        for n in normals:
            f.write("vn %.9g %.9g %.9g\n" % (float(n[0]), float(n[1]), float(n[2])))
        last, group = None, 0
        for t in tris:
            if int(t[6]) != last:
                last = int(t[6])
                f.write(f"g group{group}\nusemtl m{last}\n")
                group += 1
            f.write("f %d//%d %d//%d %d//%d\n" % (t[0] + 1, t[3] + 1, t[1] + 1, t[4] + 1, t[2] + 1, t[5] + 1))
