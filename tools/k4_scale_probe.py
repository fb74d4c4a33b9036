"""GPU box: the height-field workload against its own size -- 2 n^2 triangles for n = 125 ... 1000 (31 k ... 2 M), device-built
tree, 1024^2 x 64 spp, depth 8, at 4 and 8 resident blocks per SM.  Separates what the scene's FOOTPRINT costs (L1 / L2 /
DRAM residency) from what its STRUCTURE costs (hops per ray through small cells)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from montecarlopathtracer_b200 import tracer as T, scenes as S
T.initialize(0)
T.load_scene_file(os.path.join(ROOT, "scenes", "scene01_lambert.mcpt"), T.BUILD_SAH)
base = T.get_scene()
for n in [int(a) for a in sys.argv[1:]] or [125, 177, 250, 354, 500, 708, 1000]:
    T.create_geometry(*S.cornell_heightfield(base, n), T.BUILD_GPU)
    i = T.tree_info()
    out = []
    for mb in (4, 8):
        T.set_tuning("min_blocks", mb)
        cfg = T.render_cfg(1024, 1024, 64, 8, seed=1)
        best = 0.0
        for rep in range(2):
            T.render_scene(cfg); c = T.counters()
            best = max(best, c.paths / c.kernel_ms / 1e3)
        out.append(f"{mb} blocks {best:7.1f} Mpaths/s {c.rays / c.kernel_ms / 1e3:7.1f} Mrays/s")
    print(f"n {n:5d} tris {2 * n * n + 14:8d} leaves {i.num_leaves:8d} refs {i.num_leaf_refs:8d} device {i.device_bytes / 1e6:6.1f} MB rays/path {c.rays / c.paths:.2f} | " + " | ".join(out), flush=True)
T.set_tuning("defaults", 0)
