"""Threshold / variant sweep on the GPU box (scratch; prints a table)."""
import os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
code = r'''
import sys; sys.path.insert(0, %r)
from montecarlopathtracer_b200 import tracer as T
T.initialize(0)
out = []
for mode in (0, 1):
    T.load_scene_file(%r, mode)
    for flags, tag in ((0, "sm"), (16, "simple"), (1, "sm+nee")):
        cfg = T.render_cfg(1024, 1024, 64, 8, flags=flags)
        T.render_scene(cfg); T.render_scene(cfg)
        c = T.counters()
        out.append("%%s/m%%d %%6.1f Mp/s" %% (tag, mode, c.paths / c.kernel_ms / 1e3))
print(" | ".join(out), flush=True)
''' % (ROOT, os.path.join(ROOT, "scenes", sys.argv[1] if len(sys.argv) > 1 else "scene01_lambert.mcpt"))
for sh, lf in ((16, 3), (12, 4), (16, 4), (20, 4), (24, 4)):
    env = dict(os.environ, MCPT_SM_SHADE=str(sh), MCPT_SM_MINBLOCKS=str(lf))
    print(f"--- shade>={sh} minblocks={lf}", flush=True)
    subprocess.run([sys.executable, "-c", code], env=env)
