"""Threshold / variant sweep on the GPU box (scratch; prints a table)."""
import os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
code = r'''
import sys; sys.path.insert(0, %r)
from montecarlopathtracer_b200 import tracer as T
T.initialize(0)
T.load_scene_file(%r)
for flags, tag in ((0, "sm"), (16, "simple"), (1, "sm+nee")):
    cfg = T.render_cfg(1024, 1024, 64, 8, flags=flags)
    T.render_scene(cfg); T.render_scene(cfg)
    c = T.counters()
    print("%%-8s kernel_ms %%8.2f Mpaths/s %%7.1f Mrays/s %%8.1f" %% (tag, c.kernel_ms, c.paths / c.kernel_ms / 1e3, c.rays / c.kernel_ms / 1e3), flush=True)
''' % (ROOT, os.path.join(ROOT, "scenes", sys.argv[1] if len(sys.argv) > 1 else "scene01_lambert.mcpt"))
for sh, lf in ((12, 10), (1, 1), (4, 4), (8, 8), (16, 8), (16, 16), (24, 16), (24, 8), (20, 12), (32, 32), (8, 16), (28, 20)):
    env = dict(os.environ, MCPT_SM_SHADE=str(sh), MCPT_SM_LEAF=str(lf))
    print(f"--- shade>={sh} leaf>={lf}", flush=True)
    subprocess.run([sys.executable, "-c", code], env=env)
