"""Variant / residency / threshold sweep of render_kernel_sm on the GPU box (needs a MCPT_TUNE build)."""
import os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
scene = sys.argv[1] if len(sys.argv) > 1 else "scene01_lambert.mcpt"
code = r'''
import sys; sys.path.insert(0, %r)
from montecarlopathtracer_b200 import tracer as T
T.initialize(0)
T.load_scene_file(%r, 1)
cfg = T.render_cfg(1024, 1024, 64, 8)
T.render_scene(cfg); T.render_scene(cfg)
c = T.counters()
print("%%7.1f Mp/s" %% (c.paths / c.kernel_ms / 1e3), flush=True)
''' % (ROOT, os.path.join(ROOT, "scenes", scene))
combos = [(v, 16, 3) for v in range(8)] + [(7, 16, 2), (7, 16, 4), (7, 12, 3), (7, 20, 3), (7, 24, 3), (5, 16, 4), (3, 16, 4), (1, 16, 4)]
if len(sys.argv) > 2:
    combos = [tuple(int(x) for x in c.split(",")) for c in sys.argv[2:]]
for var, sh, mb in combos:
    env = dict(os.environ, MCPT_SM_VARIANT=str(var), MCPT_SM_SHADE=str(sh), MCPT_SM_MINBLOCKS=str(mb))
    r = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True)
    print(f"variant {var} shade>={sh} minblocks={mb}: {r.stdout.strip()} {r.stderr.strip()[-200:]}", flush=True)
