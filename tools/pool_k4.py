"""GPU box: K4 (1 M triangles, device-built tree) with the pooled-lane kernel against the shipped one."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from montecarlopathtracer_b200 import tracer as T, scenes as S
T.initialize(0)
T.load_scene_file(os.path.join(ROOT, "scenes", "scene01_lambert.mcpt"), 1)
T.create_geometry(*S.cornell_heightfield(T.get_scene(), 708), T.BUILD_GPU)
ref = None
for pool, mb in ((0, 0), (1, 0)):
    T.set_tuning("defaults", 0); T.set_tuning("pool", pool); T.set_tuning("min_blocks", mb)
    cfg = T.render_cfg(1024, 1024, 256, 8, seed=1)
    img = T.render_scene(cfg)
    best = 0.0
    for rep in range(2):
        T.render_scene(cfg); c = T.counters()
        best = max(best, c.paths / c.kernel_ms / 1e3)
    if ref is None: ref = img
    print(f"pool {pool} min_blocks {mb}: {best:7.1f} Mpaths/s rays {c.rays} image {'same' if np.array_equal(img, ref) else 'differs in %d px, max %.3g' % ((img != ref).any(-1).sum(), np.abs(img - ref).max())}", flush=True)
