"""GPU box: K4 (1 002 550-triangle height field, 1024^2 x 256 spp, depth 8) kernel-only Mpaths/s against residency
(min_blocks 4 / 5 / 6 / 8 resident 256-thread blocks per SM = 64 / 48 / 40 / 32 registers) (the persisting-L2
window over rope_nodes of an earlier version of this probe made no difference and was removed); one geometry build, images must not change (the knobs only move data and registers)."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from montecarlopathtracer_b200 import tracer as T, scenes as S
T.initialize(0)
T.tuning_from_env()
T.load_scene_file(os.path.join(ROOT, "scenes", "scene01_lambert.mcpt"), T.BUILD_SAH)
T.create_geometry(*S.cornell_heightfield(T.get_scene(), 708), T.BUILD_SAH)
i = T.tree_info()
print(f"nodes {i.num_nodes} leaves {i.num_leaves} refs {i.num_leaf_refs} depth {i.max_depth} build {i.build_seconds:.3f} s device MB {i.device_bytes / 1e6:.1f}", flush=True)
spp = int(sys.argv[1]) if len(sys.argv) > 1 else 256
variants = [int(v) for v in sys.argv[2:]] or [503]      # more than one variant needs the tuning build (MCPT_LIB=.../libmcpt_tune.so)
ref = None
for var in variants:
  for mb in ((4, 5, 6, 8) if len(variants) == 1 else (4, 8)):
    for l2 in (0,):
        T.set_tuning("variant", var); T.set_tuning("min_blocks", mb)
        cfg = T.render_cfg(1024, 1024, spp, 8, seed=1)
        img = T.render_scene(cfg)
        best = 0.0
        for rep in range(2):
            T.render_scene(cfg); c = T.counters()
            best = max(best, c.paths / c.kernel_ms / 1e3)
        if ref is None:
            ref = img
        print(f"variant {var} min_blocks {mb} l2_persist {l2}: {best:7.1f} Mpaths/s  {c.rays / c.kernel_ms / 1e3:7.1f} Mrays/s  image {'same' if np.array_equal(img, ref) else 'DIFFERENT'}", flush=True)
