"""Variant / residency / threshold sweep of render_kernel_sm on the GPU box (needs a MCPT_TUNE build)."""
import os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
scene = sys.argv[1] if len(sys.argv) > 1 else "scene01_lambert.mcpt"
code = r'''
import sys; sys.path.insert(0, %r)
from montecarlopathtracer_b200 import tracer as T
T.initialize(0)
T.tuning_from_env()   # the sweep's knobs arrive as MCPT_* variables of this child process
name = %r
if "heightfield:" in name:
    from montecarlopathtracer_b200 import scenes
    T.load_scene_file(name.split("heightfield:")[0] + "scene01_lambert.mcpt", 1)
    T.create_geometry(*scenes.cornell_heightfield(T.get_scene(), int(name.split(":")[1])), 1)
else:
    T.load_scene_file(name, 1)
cfg = T.render_cfg(1024, 1024, 64, 8, eye=(0, 5, %r), at=(0, 5, %r - 1))
T.render_scene(cfg); T.render_scene(cfg)
c = T.counters()
i = T.tree_info()
print("%%7.1f Mp/s  %%5.2f rays/path | nodes %%d refs %%d depth %%d" %% (c.paths / c.kernel_ms / 1e3, c.rays / c.paths, i.num_nodes, i.num_leaf_refs, i.max_depth), flush=True)
''' % (ROOT, os.path.join(ROOT, "scenes", scene), {"scene02.mcpt": 23.0, "scene03.mcpt": 4.75}.get(scene, 17.0), {"scene02.mcpt": 23.0, "scene03.mcpt": 4.75}.get(scene, 17.0))
combos = [(v, 20, 4) for v in range(8)] + [(3, 20, 5), (3, 16, 5), (3, 24, 4), (3, 16, 4), (3, 20, 3)]
if len(sys.argv) > 2:
    combos = [tuple(float(x) if "." in x else int(x) for x in c.split(",")) for c in sys.argv[2:]]
for c in combos:
    var, sh, mb = c[:3]
    env = dict(os.environ, MCPT_SM_VARIANT=str(var), MCPT_SM_SHADE=str(sh), MCPT_SM_MINBLOCKS=str(mb))
    if len(c) > 3:
        env["MCPT_SAH_CI"] = str(c[3])
    if len(c) > 4:
        env["MCPT_SAH_EB"] = str(c[4])
    r = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True)
    print(f"variant {var} shade>={sh} minblocks={mb} sah {c[3:]}: {r.stdout.strip()} {r.stderr.strip()[-200:]}", flush=True)
