"""K4 scene (1.0 M-triangle height field) probe: one scene build, a few short renders (GPU box)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from montecarlopathtracer_b200 import tracer as T, scenes
n = int(sys.argv[1]) if len(sys.argv) > 1 else 708
spp = int(sys.argv[2]) if len(sys.argv) > 2 else 16
T.initialize(0)
T.load_scene_file(os.path.join(ROOT, "scenes", "scene01_lambert.mcpt"), 1)
T.create_geometry(*scenes.cornell_heightfield(T.get_scene(), n), 1)
i = T.tree_info()
print(f"nodes {i.num_nodes} leaves {i.num_leaves} refs {i.num_leaf_refs} depth {i.max_depth} build {i.build_seconds:.2f}s device MB {i.device_bytes/1e6:.1f}", flush=True)
cfg = T.render_cfg(1024, 1024, spp, 8)
for k in range(3):
    T.render_scene(cfg)
    c = T.counters()
    print(f"{c.paths / c.kernel_ms / 1e3:7.1f} Mp/s {c.rays / c.kernel_ms / 1e3:7.1f} Mr/s rays/path {c.rays / c.paths:.2f}", flush=True)
import numpy as np
t = T.get_tree()
lc = t.leaf_count[t.split_axis == 0]
print("leaf size: mean %.2f max %d, hist" % (lc.mean(), lc.max()), np.bincount(np.minimum(lc, 16)).tolist(), flush=True)
