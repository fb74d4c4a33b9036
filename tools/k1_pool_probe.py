import os, sys
sys.path.insert(0, os.getcwd())
from montecarlopathtracer_b200 import tracer as T
T.initialize(0)
T.load_scene_file("scenes/scene01.mcpt", T.BUILD_SAH)
cfg = T.render_cfg(512, 512, 16, 5, seed=1)
for pool, chunks in ((0, 0), (1, 2), (1, 4), (1, 8), (0, 4)):
    T.set_tuning("defaults", 0); T.set_tuning("pool", pool); T.set_tuning("chunks", chunks)
    best = 0
    for rep in range(30):
        T.render_scene(cfg); c = T.counters(); best = max(best, c.paths / c.kernel_ms / 1e3)
    print(f"pool {pool} chunks {chunks}: {best:7.1f} Mpaths/s", flush=True)
