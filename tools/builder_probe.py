"""GPU box: the three KD builders on the benchmark scenes -- build seconds, tree size, kernel-only Mpaths/s of the tree they give."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from montecarlopathtracer_b200 import tracer as T, scenes as S
T.initialize(0)
T.set_tuning("build_timing", 1)
T.load_scene_file(os.path.join(ROOT, "scenes", "scene01_lambert.mcpt"), T.BUILD_SAH)
base = T.get_scene()
cases = [("scene01_lambert", None, 17.0, 256), ("scene02", None, 23.0, 256), ("scene03", None, 4.75, 64), ("heightfield:224", 224, 17.0, 128), ("heightfield:708", 708, 17.0, 128)]
for name, hf, ez, spp in cases:
    if hf:
        tables = S.cornell_heightfield(base, hf)
    else:
        T.load_scene_file(os.path.join(ROOT, "scenes", name + ".mcpt"), T.BUILD_SAH)
        tables = T.get_scene()
    for mode, tag in ((T.BUILD_SAH, "host SAH"), (T.BUILD_GPU, "GPU SAH"), (T.BUILD_GPU, "GPU SAH (2nd)")):
        t0 = time.perf_counter()
        T.create_geometry(*tables, mode)
        dt = time.perf_counter() - t0
        i = T.tree_info()
        cfg = T.render_cfg(1024, 1024, spp, 8, seed=1, eye=(0, 5, ez), at=(0, 5, ez - 1))
        T.render_scene(cfg); T.render_scene(cfg)
        c = T.counters()
        print(f"{name:16s} {tag:14s} create_geometry {dt * 1e3:8.1f} ms (tree {i.build_seconds * 1e3:7.1f} ms) nodes {i.num_nodes:8d} leaves {i.num_leaves:8d} refs {i.num_leaf_refs:8d} depth {i.max_depth:2d} | {c.paths / c.kernel_ms / 1e3:7.1f} Mpaths/s", flush=True)
