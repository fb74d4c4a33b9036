"""GPU box: mcpt_create_geometry wall time for the small scenes with the host SAH builder and with the device builder."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from montecarlopathtracer_b200 import tracer as T
from oracle.pyoracle import SceneArrays
T.initialize(0)
for name in ("scene01", "scene02", "scene03"):
    s = SceneArrays.load(os.path.join(ROOT, "scenes", name + ".mcpt"))
    for mode, tag in ((T.BUILD_SAH, "host SAH"), (T.BUILD_GPU, "device")):
        ts = []
        for i in range(30):
            t = time.perf_counter(); T.create_geometry(s.verts, s.normals, s.tris, s.mats, mode); ts.append(time.perf_counter() - t)
        i = T.tree_info()
        print(f"{name} {tag:9s}: create_geometry min {1e3 * min(ts):.3f} ms median {1e3 * sorted(ts)[15]:.3f} ms  (build {1e3 * i.build_seconds:.3f} ms, {i.num_nodes} nodes, {i.num_leaf_refs} refs)", flush=True)
