"""GPU box: K4 (1 M triangles, device-built tree, 8 blocks/SM) against the run-time knobs of render_kernel_sm."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from montecarlopathtracer_b200 import tracer as T, scenes as S
T.initialize(0)
T.load_scene_file(os.path.join(ROOT, "scenes", "scene01_lambert.mcpt"), T.BUILD_SAH)
T.create_geometry(*S.cornell_heightfield(T.get_scene(), 708), T.BUILD_GPU)
cfg = T.render_cfg(1024, 1024, 256, 8, seed=1)
def run(tag):
    best = 0.0
    for rep in range(2):
        T.render_scene(cfg); c = T.counters()
        best = max(best, c.paths / c.kernel_ms / 1e3)
    print(f"{tag:40s} {best:7.1f} Mpaths/s", flush=True)
run("defaults")
for knob, values in (("tri_cap", (2, 3, 6, 8)), ("shade_threshold", (8, 12, 16, 22, 26)), ("cam_threshold", (8, 24)), ("chunks", (1, 2, 8, 16))):
    for v in values:
        T.set_tuning("defaults", 0); T.set_tuning(knob, v)
        run(f"{knob} = {v}")
T.set_tuning("defaults", 0)
