"""How much do the tails cost?  K2 scene, 2048x2048 (27.7 waves of blocks at one slice), kernel-only Mpaths/s against
the samples a lane traces in one go (tuning knob chunks = 1: spp per lane = spp)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from montecarlopathtracer_b200 import tracer as T
T.initialize(0)
T.load_scene_file(os.path.join(ROOT, "scenes", "scene01_lambert.mcpt"), 1)
T.set_tuning("chunks", 1)
for w, spps in ((2048, (16, 64, 256, 1024, 4096)), (1024, (64, 256, 1024, 4096))):
    for spp in spps:
        cfg = T.render_cfg(w, w, spp, 8)
        best = 0
        for rep in range(2):
            T.render_scene(cfg); c = T.counters()
            best = max(best, c.paths / c.kernel_ms / 1e3)
        print(f"{w}x{w} spp/lane {spp:5d}: {best:7.1f} Mpaths/s", flush=True)
