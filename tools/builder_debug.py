"""GPU box: two GPU builds of the same scene, field by field."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from montecarlopathtracer_b200 import tracer as T
from oracle.pyoracle import SceneArrays
T.initialize(0)
for name in ("scene01", "scene02"):
    s = SceneArrays.load(os.path.join(ROOT, "scenes", name + ".mcpt"))
    trees = []
    for rep in range(3):
        T.create_geometry(s.verts, s.normals, s.tris, s.mats, T.BUILD_GPU)
        trees.append(T.get_tree())
    a = trees[0]
    for rep in (1, 2):
        b = trees[rep]
        for f in ("left", "aabb", "split_axis", "split_value", "leaf_offset", "leaf_count", "leaf_ids"):
            x, y = getattr(a, f), getattr(b, f)
            if x.shape != y.shape:
                print(name, rep, f, "shape", x.shape, y.shape); continue
            d = np.flatnonzero((x.view(np.uint32) != y.view(np.uint32)).reshape(len(x), -1).any(1))
            print(name, rep, f, "differing rows", len(d), d[:8], (x[d[:3]], y[d[:3]]) if len(d) else "")
