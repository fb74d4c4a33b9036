"""Traversal work per ray and render speed for both tree builders (scratch, GPU box)."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from montecarlopathtracer_b200 import tracer as T
T.initialize(0)
g = np.random.Generator(np.random.Philox(7))
for scene in ("scene01_lambert", "scene01", "scene02", "scene03"):
    T.load_scene_file(os.path.join(ROOT, "scenes", scene + ".mcpt"))
    v, _, _, _ = T.get_scene()
    lo, hi = v.min(0), v.max(0)
    n = 1 << 18
    o = g.uniform(lo, hi, (n, 3)); d = g.normal(size=(n, 3)); d /= np.linalg.norm(d, axis=1, keepdims=True)
    rays = np.concatenate([o, d], 1).astype(np.float32)
    for mode in (0, 1):
        T.load_scene_file(os.path.join(ROOT, "scenes", scene + ".mcpt"), mode)
        i = T.tree_info()
        kd = T.traversal_stats(rays, T.ISECT_KD); rf = T.traversal_stats(rays, T.ISECT_REFERENCE)
        line = f"{scene:16s} mode {mode} nodes {i.num_nodes:6d} refs {i.num_leaf_refs:6d} depth {i.max_depth:2d} | kd: nodes/ray {kd[0]/n:6.2f} leaves {kd[1]/n:5.2f} tris {kd[2]/n:5.2f} | ref-order: popped {rf[0]/n:6.2f} entered {rf[1]/n:6.2f} tris {rf[2]/n:5.2f}"
        eye = 23 if scene == "scene02" else (4.75 if scene == "scene03" else 17)
        for flags, tag in ((0, "sm"), (16, "simple")):
            cfg = T.render_cfg(1024, 1024, 32, 8, flags=flags, eye=(0, 5, eye), at=(0, 5, eye - 1))
            T.render_scene(cfg); T.render_scene(cfg)
            c = T.counters()
            line += f" | {tag} {c.paths / c.kernel_ms / 1e3:6.1f} Mp/s {c.rays / c.kernel_ms / 1e3:7.1f} Mr/s"
        print(line, flush=True)
