"""GPU box, tuning build (MCPT_BUILD_NAME=tune MCPT_TUNE=1 -> lib/libmcpt_tune.so, run with MCPT_LIB pointing at it):
per-warp round statistics of render_kernel_sm (VAR 511 = shipped 503 + STATS) -- lanes per TRAVERSE round, (lane, triangle)
pairs per round against the round's longest triangle loop (what ballot compaction of the tests could pack), point
locations behind exit faces and their walk lengths.  The kernel prints to stderr through mcpt's sm_stats knob."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from montecarlopathtracer_b200 import tracer as T, scenes as S
T.initialize(0)
T.set_tuning("variant", 511); T.set_tuning("sm_stats", 1)
for name, eye, spp in (("scene01_lambert", 17.0, 64), ("scene01", 17.0, 64), ("scene02", 23.0, 64), ("scene03", 4.75, 32), ("heightfield:708", 17.0, 16), ("heightfield_gpu:708", 17.0, 16)):
    if name.startswith("heightfield"):
        T.load_scene_file(os.path.join(ROOT, "scenes", "scene01_lambert.mcpt"), 1)
        T.create_geometry(*S.cornell_heightfield(T.get_scene(), 708), 2 if "gpu" in name else 1)
        i = T.tree_info()
        print(f"nodes {i.num_nodes} leaves {i.num_leaves} refs {i.num_leaf_refs} depth {i.max_depth} device MB {i.device_bytes / 1e6:.1f}", flush=True)
    else:
        T.load_scene_file(os.path.join(ROOT, "scenes", name + ".mcpt"), 1)
    print("==", name, flush=True)
    sys.stderr.write(f"== {name}\n"); sys.stderr.flush()
    T.render_scene(T.render_cfg(1024, 1024, spp, 8, eye=(0, 5, eye), at=(0, 5, eye - 1)))
    c = T.counters()
    print(f"{c.paths / c.kernel_ms / 1e3:.1f} Mpaths/s (with statistics on), {c.rays / c.paths:.3f} rays/path", flush=True)
