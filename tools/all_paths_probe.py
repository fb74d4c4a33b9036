"""GPU box (compute-sanitizer is closed on this pool: run plain): small invocations of every device code path (render kernels in all modes, the pooled
kernel, the parity hooks, the device KD builder, the no-alloc entry point).
    python tools/sanitize_probe.py"""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from montecarlopathtracer_b200 import tracer as T, scenes as S
import torch
T.initialize(0)
for scene, z in (("scene01", 17.0), ("scene03", 4.75)):
    for mode in (T.BUILD_REFERENCE, T.BUILD_SAH, T.BUILD_GPU):
        T.load_scene_file(os.path.join(ROOT, "scenes", scene + ".mcpt"), mode)
        for flags in (0, T.FLAG_NEE, T.FLAG_SIMPLE_KERNEL, T.FLAG_SIMPLE_KERNEL | T.FLAG_REF_TRAVERSAL):
            for rng in (T.RNG_PHILOX, T.RNG_REFERENCE):
                T.render_scene(T.render_cfg(67, 45, 3, 5, seed=2, flags=flags, rng_mode=rng, eye=(0, 5, z), at=(0, 5, z - 1)), moments=True)
        T.render_scene(T.cuda_twin_render_cfg(80, 60, 2, 1))
        T.set_tuning("pool", 1)
        for chunks in (1, 2, 3):
            T.set_tuning("chunks", chunks)
            T.render_scene(T.render_cfg(67, 45, 5, 5, seed=2, eye=(0, 5, z), at=(0, 5, z - 1)))
        T.set_tuning("defaults", 0)
        for mb in (5, 6, 8):
            T.set_tuning("min_blocks", mb)
            T.render_scene(T.render_cfg(67, 45, 3, 5, seed=2, eye=(0, 5, z), at=(0, 5, z - 1)))
        T.set_tuning("defaults", 0)
        g = np.random.Generator(np.random.Philox(1))
        rays = np.concatenate([g.uniform(-6, 6, (4096, 3)), g.normal(size=(4096, 3))], 1).astype(np.float32)
        for v in (T.ISECT_REFERENCE, T.ISECT_KD, T.ISECT_ROPES):
            T.intersect_batch(rays, v)
        T.primary_rays(T.render_cfg(64, 48, 1, 5), 0)
        acc = torch.zeros(45, 67, 4, device="cuda"); torch.cuda.synchronize()
        T.render_accumulate_device_noalloc(T.render_cfg(67, 45, 3, 5, seed=2), acc.data_ptr())
        print(scene, mode, "ok", flush=True)
T.load_scene_file(os.path.join(ROOT, "scenes", "scene01_lambert.mcpt"), T.BUILD_SAH)
T.create_geometry(*S.cornell_heightfield(T.get_scene(), 60), T.BUILD_GPU)
T.render_scene(T.render_cfg(96, 64, 2, 8, seed=1))
u = np.random.default_rng(0).random((1000, 2), np.float32); n = np.tile(np.array([[0, 1, 0]], np.float32), (1000, 1))
T.sample_batch(T.SAMPLE_HEMI, u, n, None, None, None)
print("all ok")
