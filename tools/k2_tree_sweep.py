"""GPU box: the L1-resident scenes against the SAH constants of the host builder with the CURRENT kernel (the constants
were last swept in round 1, before the stackless walk): kernel-only Mpaths/s at 1024^2."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from montecarlopathtracer_b200 import tracer as T
T.initialize(0)
EYE = {"scene02": 23.0, "scene03": 4.75}
cis = [float(x) for x in sys.argv[1:] if not x.startswith("eb=")] or [0.7, 1.0, 1.3, 1.6, 2.0, 3.0]
ebs = [float(x[3:]) for x in sys.argv[1:] if x.startswith("eb=")] or [1.0, 0.85]
for scene, spp in (("scene01_lambert", 512), ("scene01", 256), ("scene02", 256), ("scene03", 64)):
    z = EYE.get(scene, 17.0)
    for eb in ebs:
        row = []
        for ci in cis:
            T.set_tuning("defaults", 0); T.set_tuning("sah_ci", ci); T.set_tuning("sah_eb", eb)
            T.load_scene_file(os.path.join(ROOT, "scenes", scene + ".mcpt"), T.BUILD_SAH)
            i = T.tree_info()
            cfg = T.render_cfg(1024, 1024, spp, 8, seed=1, eye=(0, 5, z), at=(0, 5, z - 1))
            best = 0.0
            for rep in range(2):
                T.render_scene(cfg); c = T.counters()
                best = max(best, c.paths / c.kernel_ms / 1e3)
            row.append(f"ci {ci:4.2f}: {i.num_leaves:5d} leaves {i.num_leaf_refs:5d} refs {best:7.1f}")
        print(f"{scene:16s} eb {eb:4.2f} | " + " | ".join(row), flush=True)
T.set_tuning("defaults", 0)
