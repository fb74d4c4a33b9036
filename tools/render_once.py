"""One 1024x1024x64spp render of a scene file through the C ABI (for ncu captures of kernel variants)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from montecarlopathtracer_b200 import tracer as T
T.initialize(0)
T.tuning_from_env()   # MCPT_SM_VARIANT etc. (tuning build) for ncu captures of variants
T.load_scene_file(os.path.join(ROOT, "scenes", sys.argv[1] if len(sys.argv) > 1 else "scene01_lambert.mcpt"), 1)
Z = {"scene02.mcpt": 23.0, "scene03.mcpt": 4.75}.get(sys.argv[1] if len(sys.argv) > 1 else "", 17.0)
cfg = T.render_cfg(1024, 1024, int(sys.argv[2]) if len(sys.argv) > 2 else 64, 8, eye=(0, 5, Z), at=(0, 5, Z - 1))
T.render_scene(cfg)
c = T.counters()
print("%7.1f Mp/s" % (c.paths / c.kernel_ms / 1e3))
