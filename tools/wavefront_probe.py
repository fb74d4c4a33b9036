"""What a wavefront's traversal stage would run at: one whole ray per thread (intersect_kernel, stackless walk)
over (a) coherent camera rays and (b) incoherent rays with origins inside the scene -- the population a path
tracer's bounce queue holds.  Run under ncu --metrics smsp__thread_inst_executed_per_inst_executed.ratio."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from montecarlopathtracer_b200 import tracer as T

T.initialize(0)
T.load_scene_file(os.path.join(ROOT, "scenes", "scene01_lambert.mcpt"), 1)
cfg = T.render_cfg(1024, 1024, 1, 8, eye=(0, 5, 17.0), at=(0, 5, 16.0))
cam = T.primary_rays(cfg, 0).reshape(-1, 6)
g = np.random.Generator(np.random.Philox(7))
verts = T.get_scene()[0]
lo, hi = verts.min(0), verts.max(0)
o = g.uniform(lo, hi, (1 << 20, 3))
d = g.normal(size=(1 << 20, 3)); d /= np.linalg.norm(d, axis=1, keepdims=True)
inc = np.concatenate([o, d], 1).astype(np.float32)
for name, rays in (("camera", cam), ("incoherent", inc)):
    for variant in (T.ISECT_ROPES, T.ISECT_KD):
        tri, t, b, gm = T.intersect_batch(rays, variant)
        print(name, "variant", variant, "hits", int((tri >= 0).sum()), "of", len(rays), flush=True)
