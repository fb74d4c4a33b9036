"""The two workloads the reference's project report publishes times for (brute-force CUDA
tracer on a GTX 1070: scene 1 800x600x1000 spp in 200 s = 2.40 Mpaths/s, scene 2 800x600x100 spp in 34 s =
1.41 Mpaths/s), rendered in MCPT_TRACER_CUDA_TWIN mode (the twin's estimator: 7 events + gather, no roulette,
+-1 px jitter, hfov 60).  Context only: different algorithm (KD ropes against O(triangles) per ray) and hardware."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from montecarlopathtracer_b200 import tracer as T

T.initialize(0)
for name, scene_id, spp, pub_s in (("scene01.mcpt", 1, 1000, 200.0), ("scene02.mcpt", 2, 100, 34.0), ("scene02.mcpt", 2, 10000, 58 * 60.0)):
    T.load_scene_file(os.path.join(ROOT, "scenes", name), 1)
    cfg = T.cuda_twin_render_cfg(800, 600, spp, scene_id)
    T.render_scene(T.cuda_twin_render_cfg(800, 600, 16, scene_id))  # warm-up
    t0 = time.time()
    img = T.render_scene(cfg)
    dt = time.time() - t0
    c = T.counters()
    print(f"{name} scene_id {scene_id} 800x600x{spp}: {dt:.3f} s wall incl. D2H ({800 * 600 * spp / dt / 1e6:.0f} Mpaths/s), "
          f"kernel {c.kernel_ms:.1f} ms, {c.rays / c.paths:.2f} rays/path, mean radiance {float(img.mean()):.4f}; "
          f"published: {pub_s:.0f} s ({800 * 600 * spp / pub_s / 1e6:.2f} Mpaths/s, GTX 1070, brute force)", flush=True)
