"""GPU box: render_kernel_pool (mcpt_set_tuning("pool", sets)) against render_kernel_sm: images must be bit-identical and
ray counts equal; then kernel-only Mpaths/s per scene.   python tools/pool_probe.py [spp] [k2spp] [min_blocks...]"""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from montecarlopathtracer_b200 import tracer as T
spp = int(sys.argv[1]) if len(sys.argv) > 1 else 256
k2spp = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
mbs = [int(x) for x in sys.argv[3:]] or [4]
EYE = {"scene02": 23.0, "scene03": 4.75}
T.initialize(0)


def render(scene, w, h, n, depth=8, **kw):
    z = EYE.get(scene, 17.0)
    out = T.render_scene(T.render_cfg(w, h, n, depth, seed=3, eye=(0, 5, z), at=(0, 5, z - 1), **kw))
    return out, T.counters()


for scene in ("scene01_lambert", "scene01", "scene02", "scene03"):
    T.load_scene_file(os.path.join(ROOT, "scenes", scene + ".mcpt"), 1)
    line = []
    for pool, mb in [(0, 4), (1, 4)]:
        T.set_tuning("defaults", 0)
        ref, rc = render(scene, 203, 121, 24, 5 if scene != "scene01_lambert" else 8)
        ref2, rc2 = render(scene, 64, 48, 3, 8, rng_mode=T.RNG_REFERENCE)
        T.set_tuning("pool", pool); T.set_tuning("min_blocks", mb)
        img, c = render(scene, 203, 121, 24, 5 if scene != "scene01_lambert" else 8)
        img2, c2 = render(scene, 64, 48, 3, 8, rng_mode=T.RNG_REFERENCE)
        same = np.array_equal(img, ref) and c.rays == rc.rays and np.array_equal(img2, ref2) and c2.rays == rc2.rays
        n = k2spp if scene == "scene01_lambert" else spp
        render(scene, 1024, 1024, 16)
        best = 0.0
        for rep in range(2):
            _, c = render(scene, 1024, 1024, n)
            best = max(best, c.paths / c.kernel_ms / 1e3)
        line.append(f"pool {pool} mb {mb} {'same' if same else 'DIFFERENT'} {best:7.1f}")
        if not same:
            d = np.abs(img - ref)
            line.append(f"(max diff {d.max():.3g}, {int((d > 0).any(-1).sum())} px, rays {c.rays} vs {rc.rays})")
    print(f"{scene:16s} " + " | ".join(line), flush=True)
T.set_tuning("defaults", 0)
