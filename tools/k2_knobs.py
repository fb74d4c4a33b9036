"""GPU box: K2 (Cornell, 1024^2, depth 8) against the run-time knobs of render_kernel_sm (kernel-only Mpaths/s, 512 spp)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from montecarlopathtracer_b200 import tracer as T
T.initialize(0)
T.load_scene_file(os.path.join(ROOT, "scenes", "scene01_lambert.mcpt"), T.BUILD_SAH)
cfg = T.render_cfg(1024, 1024, 512, 8, seed=1)
def run(tag):
    best = 0.0
    for rep in range(3):
        T.render_scene(cfg); c = T.counters()
        best = max(best, c.paths / c.kernel_ms / 1e3)
    print(f"{tag:40s} {best:7.1f} Mpaths/s", flush=True)
run("defaults")
for knob, values in (("shade_threshold", (14, 16, 17, 19, 20, 22)), ("tri_cap", (3, 5, 6)), ("cam_threshold", (8, 12, 20, 24)), ("chunks", (2, 8, 16))):
    for v in values:
        T.set_tuning("defaults", 0); T.set_tuning(knob, v)
        run(f"{knob} = {v}")
T.set_tuning("defaults", 0)
run("defaults again")
