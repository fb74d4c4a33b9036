import os, sys
import numpy as np
sys.path.insert(0, os.getcwd())
from montecarlopathtracer_b200 import tracer as T
T.initialize(0)
T.load_scene_file("scenes/scene03.mcpt", 1)
for kw in (dict(), dict(rng_mode=T.RNG_REFERENCE)):
    for (w, h, n, d) in ((203, 121, 24, 5), (64, 48, 3, 8), (64, 48, 16, 5)):
        out = []
        for pool in (0, 2):
            T.set_tuning("defaults", 0); T.set_tuning("pool", pool)
            img = T.render_scene(T.render_cfg(w, h, n, d, seed=3, eye=(0, 5, 4.75), at=(0, 5, 3.75), **kw))
            c = T.counters(); out.append((c.rays, c.paths, float(img.sum())))
        print(kw, w, h, n, d, out, flush=True)
