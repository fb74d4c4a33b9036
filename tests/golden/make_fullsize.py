"""Freezes matched-spp ORACLE renders of BASELINE.json's configurations at their own size (run HERE, once:
python tests/golden/make_fullsize.py [K2 K3a K3b S3]; ~1 h of 8 cores in total).

The reference render of north_star's image gate ("per-pixel radiance within tolerance of the reference render at
matched spp: mean relative error < 1 %, no pixel beyond 4 sigma of its sample variance") is the oracle in the
reference's own random stream (rtx.hlsl:61-82 TEA + Park-Miller, frame seeds mt19937(1234)) at the FULL resolution
and the FULL sample count of the configuration.  A 1024^2 x 3 float image pair does not belong in git, so what is kept is
  lattice   every `stride`-th pixel in x and y, PER PIXEL: mean (float32) and per-sample standard deviation (float16)
            -- the "no pixel beyond 4 sigma" gate runs on these pixels one by one;
  tiles     4x4 tile means of the WHOLE image (float32) and of the per-sample variance (float16)
            -- the "mean relative error" gate and a tile-level z test cover every pixel;
  totals    image mean, rays per path, oracle counters.
tests/test_fullsize.py compares the CUDA renderer against them on the GPU box (no /root/reference, no oracle run).
"""
import json
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from oracle.pyoracle import Oracle, SceneArrays  # noqa: E402

# name: (scene, width, height, spp, depth, eye z, lattice stride)   -- BASELINE.json configs[1], [2] (3a, 3b) + scene03
CONFIGS = {
    "K2": ("scene01_lambert", 1024, 1024, 1024, 8, 17.0, 2),
    "K3a": ("scene01", 1024, 1024, 2048, 8, 17.0, 4),
    "K3b": ("scene02", 1024, 1024, 2048, 8, 23.0, 4),
    "S3": ("scene03", 512, 512, 512, 8, 4.75, 2),
}


def make(name, threads):
    scene, W, H, spp, depth, ez, stride = CONFIGS[name]
    o = Oracle()
    s = SceneArrays.load(os.path.join(ROOT, "scenes", scene + ".mcpt"))
    sc = o.scene(s)
    t0 = time.time()
    ssum, ssq, _, cnt = sc.render(o.cfg(W, H, spp, depth, eye=(0, 5, ez), at=(0, 5, ez - 1), threads=threads))
    mean = ssum / spp
    var = np.maximum(ssq / spp - mean * mean, 0.0)          # per-sample variance
    tile = lambda a: a.reshape(H // 4, 4, W // 4, 4, 3).mean((1, 3))
    out = dict(lattice_mean=mean[::stride, ::stride].astype(np.float32), lattice_std=np.sqrt(var[::stride, ::stride]).astype(np.float16),
               tile_mean=tile(mean).astype(np.float32), tile_var=tile(var).astype(np.float16))
    np.savez_compressed(os.path.join(HERE, f"fullsize_{name}.npz"), **out)
    meta = dict(scene=scene, width=W, height=H, spp=spp, depth=depth, eye_z=ez, stride=stride, rng="reference (TEA + Park-Miller, mt19937(1234) frame seeds)",
                image_mean=float(mean.mean()), channel_means=[float(x) for x in mean.mean((0, 1))], paths=int(cnt.paths), rays=int(cnt.rays),
                rays_per_path=cnt.rays / cnt.paths, nodes_per_ray=cnt.nodes_popped / cnt.rays, tris_per_ray=cnt.tris_tested / cnt.rays,
                oracle_seconds=time.time() - t0, oracle_threads=threads)
    with open(os.path.join(HERE, f"fullsize_{name}.json"), "w") as f:
        json.dump(meta, f, indent=1, sort_keys=True)
    print(name, json.dumps(meta), flush=True)


if __name__ == "__main__":
    names = [a for a in sys.argv[1:] if a in CONFIGS] or list(CONFIGS)
    threads = int(os.environ.get("OMP_NUM_THREADS", "0") or 0)
    for n in names:
        make(n, threads)
