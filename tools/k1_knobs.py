"""GPU box: K1 (Cornell, 512^2 x 16 spp, depth 5: 4.2 M paths, ~2.7 ms) against the slice count and the vote thresholds."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from montecarlopathtracer_b200 import tracer as T
T.initialize(0)
T.load_scene_file(os.path.join(ROOT, "scenes", "scene01.mcpt"), T.BUILD_SAH)
cfg = T.render_cfg(512, 512, 16, 5, seed=1)
def run(tag):
    best = 0.0
    for rep in range(20):
        T.render_scene(cfg); c = T.counters()
        best = max(best, c.paths / c.kernel_ms / 1e3)
    print(f"{tag:40s} {best:7.1f} Mpaths/s  ({c.kernel_ms:.3f} ms last)", flush=True)
run("defaults")
for knob, values in (("chunks", (1, 2, 3, 4, 6, 8, 16)), ("shade_threshold", (12, 16, 22)), ("cam_threshold", (8, 24))):
    for v in values:
        T.set_tuning("defaults", 0); T.set_tuning(knob, v)
        run(f"{knob} = {v}")
T.set_tuning("defaults", 0)
