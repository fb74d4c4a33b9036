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
# (c) the same incoherent rays SORTED the way a wavefront tracer would sort its bounce queue: direction octant, then a 30-bit
# Morton code of the origin -- rays that start in the same region and head the same way become neighbours in a warp
def morton30(p):
    q = np.clip(((p - lo) / (hi - lo) * 1023.0), 0, 1023).astype(np.uint64)
    def spread(v):
        v = (v | (v << 16)) & 0x030000FF; v = (v | (v << 8)) & 0x0300F00F; v = (v | (v << 4)) & 0x030C30C3; v = (v | (v << 2)) & 0x09249249
        return v
    return spread(q[:, 0]) | (spread(q[:, 1]) << 1) | (spread(q[:, 2]) << 2)
octant = ((inc[:, 3] < 0).astype(np.uint64) | ((inc[:, 4] < 0).astype(np.uint64) << 1) | ((inc[:, 5] < 0).astype(np.uint64) << 2))
srt = inc[np.argsort((octant << 30) | morton30(inc[:, :3].astype(np.float64)), kind="stable")]
for name, rays in (("camera", cam), ("incoherent", inc), ("incoherent sorted by octant + Morton(origin)", srt)):
    for variant in (T.ISECT_ROPES, T.ISECT_KD):
        tri, t, b, gm = T.intersect_batch(rays, variant)
        print(name, "variant", variant, "hits", int((tri >= 0).sum()), "of", len(rays), flush=True)
