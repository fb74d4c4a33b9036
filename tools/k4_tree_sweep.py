"""GPU box: K4 (1 002 550-triangle height field, 1024^2, depth 8) against the SAH constants of both builders: a scene that
outgrows L2 is bound by the dependent loads of every hop (header -> neighbour link -> point location -> next header), so
fewer, fatter leaves trade hops for triangle tests whose loads are independent.  Prints tree size and kernel-only Mpaths/s."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from montecarlopathtracer_b200 import tracer as T, scenes as S
T.initialize(0)
T.load_scene_file(os.path.join(ROOT, "scenes", "scene01_lambert.mcpt"), T.BUILD_SAH)
hf = S.cornell_heightfield(T.get_scene(), 708)
spp = int(sys.argv[1]) if len(sys.argv) > 1 else 64
cis = [float(x) for x in sys.argv[2:] if not x.startswith("eb=")] or [0.125, 0.25, 0.5, 1.0, 2.0]
ebs = [float(x[3:]) for x in sys.argv[2:] if x.startswith("eb=")] or [1.0]
for mode, name in ((T.BUILD_GPU, "gpu"), (T.BUILD_SAH, "host")):
    for ci in cis:
        for eb in ebs:
            T.set_tuning("defaults", 0); T.set_tuning("sah_ci", ci); T.set_tuning("sah_eb", eb)
            T.create_geometry(*hf, mode)
            i = T.tree_info()
            cfg = T.render_cfg(1024, 1024, spp, 8, seed=1)
            best = 0.0
            for rep in range(2):
                T.render_scene(cfg); c = T.counters()
                best = max(best, c.paths / c.kernel_ms / 1e3)
            print(f"{name:4s} ci {ci:5.3f} eb {eb:4.2f}: nodes {i.num_nodes:8d} leaves {i.num_leaves:8d} refs {i.num_leaf_refs:8d} depth {i.max_depth:2d} "
                  f"build {i.build_seconds:6.3f} s device {i.device_bytes / 1e6:6.1f} MB | {best:7.1f} Mpaths/s {c.rays / c.kernel_ms / 1e3:7.1f} Mrays/s", flush=True)
T.set_tuning("defaults", 0)
