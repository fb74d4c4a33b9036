"""Kernel variants of render_kernel_sm against each other on the GPU box (MCPT_TUNE build): images must be
bit-identical to the first variant's, ray counts equal; then kernel-only Mpaths/s per scene.

    MCPT_TUNE=1 python -m montecarlopathtracer_b200.build --force
    python tools/var_probe.py 119 247 375 503 [--spp 256] [--k2spp 1024]     (a variant may carry knobs: 119:shade_threshold=18)
"""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from montecarlopathtracer_b200 import tracer as T

args = sys.argv[1:]
spp, k2spp = 256, 1024
if "--spp" in args: i = args.index("--spp"); spp = int(args[i + 1]); del args[i:i + 2]
if "--k2spp" in args: i = args.index("--k2spp"); k2spp = int(args[i + 1]); del args[i:i + 2]
k4 = "--k4" in args
if k4: args.remove("--k4")
variants = [a for a in args]
EYE = {"scene02": 23.0, "scene03": 4.75}


def setenv(v):
    """variant[:knob=value...] -> mcpt_set_tuning (the library does not read the environment)."""
    parts = v.split(":")
    T.set_tuning("defaults", 0)
    T.set_tuning("variant", int(parts[0]))
    for kv in parts[1:]:
        k, x = kv.split("=")
        T.set_tuning(k, float(x))


def render(scene, w, h, n, depth=8):
    z = EYE.get(scene, 17.0)
    out = T.render_scene(T.render_cfg(w, h, n, depth, seed=3, eye=(0, 5, z), at=(0, 5, z - 1)))
    return out, T.counters()


T.initialize(0)
for scene in (("heightfield",) if k4 else ("scene01_lambert", "scene01", "scene02", "scene03")):
    if k4:   # K4: 1 002 550 triangles, device-built tree, residency automatic (8 blocks per SM) unless a variant carries min_blocks
        from montecarlopathtracer_b200 import scenes as S
        T.load_scene_file(os.path.join(ROOT, "scenes", "scene01_lambert.mcpt"), 1)
        T.create_geometry(*S.cornell_heightfield(T.get_scene(), 708), T.BUILD_GPU)
    else:
        T.load_scene_file(os.path.join(ROOT, "scenes", scene + ".mcpt"), 1)
    ref = None
    line = []
    for v in variants:
        # reference image: the first variant's kernel under this variant's knobs (slices change the summation order)
        setenv(":".join([variants[0].split(":")[0]] + v.split(":")[1:]))
        rimg, rc = render(scene, 203, 121, 24, 5 if scene != "scene01_lambert" else 8)
        setenv(v)
        img, c = render(scene, 203, 121, 24, 5 if scene != "scene01_lambert" else 8)
        same = np.array_equal(img, rimg) and c.rays == rc.rays
        n = k2spp if scene == "scene01_lambert" else spp
        render(scene, 1024, 1024, 16)
        best = 0.0
        for rep in range(2):
            _, c = render(scene, 1024, 1024, n)
            best = max(best, c.paths / c.kernel_ms / 1e3)
        line.append(f"{v} {'same' if same else 'DIFFERENT'} {best:7.1f}")
    print(f"{scene:16s} " + " | ".join(line), flush=True)
