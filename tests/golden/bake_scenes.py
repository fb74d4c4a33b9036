"""Generates the scene fixtures under scenes/ from the reference's own assets.

Run HERE (authoring container, /root/reference mounted):  python tests/golden/bake_scenes.py

Each OBJ/MTL pair is read by the REFERENCE'S OWN loader -- tinyobjloader 1.1.1 exactly as
Quin::Utils::QuinModel calls it (Structure.hpp:9-12), compiled in oracle/_ref -- and the resulting
flat tables (the CSTriangle / CSMaterial contents of ShaderResource.hpp:88-104,204-215) are written
as MCPTSCN1 containers.  They are the golden vectors the product's OBJ loader is checked against and
the inputs of the GPU tests/bench (the GPU box has no /root/reference).

  scene01.mcpt          MCRT/QuinEngine/Res/scene01.{obj,mtl}            K1, K3a  (862 tris)
  scene01_lambert.mcpt  same geometry; sphere_mirror / sphere_transparent -> Kd 0.5
  scene02.mcpt          CVMCTracer/.../Resources/scene02.{obj,mtl}       K3b      (1732 tris)
  scene03.mcpt          CVMCTracer/.../Resources/scene03.{obj,mtl}       K3c      (3024 tris)
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle.pyoracle import Reference, SceneArrays, build  # noqa: E402

QE = "/root/reference/MCRT/QuinEngine/Res/"
CV = "/root/reference/CVMCTracer/CVMCTracer/Resources/"


def main():
    build()
    ref = Reference()
    out = os.path.join(ROOT, "scenes")
    os.makedirs(out, exist_ok=True)
    for name, path in (("scene01", QE + "scene01.obj"), ("scene02", CV + "scene02.obj"), ("scene03", CV + "scene03.obj")):
        h, s, shapes = ref.load_obj(path)
        s.save(os.path.join(out, name + ".mcpt"))
        print(name, len(s.verts), len(s.normals), len(s.tris), len(s.mats), shapes)
        if name == "scene01":
            # K2: all-Lambert + the area light.  Materials in MTL order: blinn2SG, initialShadingGroup,
            # lambert2SG, lambert3SG, sphere_mirror, sphere_transparent (QE/Res/scene01.mtl:1-19).
            m = s.mats.copy()
            for k in (4, 5):
                m[k] = 0.0
                m[k, 3:6] = 0.5          # Kd 0.5 0.5 0.5
                m[k, 9], m[k, 10], m[k, 11] = 1.0, 0.0, 1.0   # tinyobj defaults Ns=1, Tr=1-d=0, Ni=1
            SceneArrays(s.verts, s.normals, s.tris, m).save(os.path.join(out, "scene01_lambert.mcpt"))
        ref.free(h)


if __name__ == "__main__":
    main()
