"""Scratch diagnostics run on the GPU box (writes to gpurun_out/)."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from montecarlopathtracer_b200 import tracer as T
from oracle.pyoracle import Oracle, SceneArrays

T.initialize(0)
o = Oracle()
# 1. scene03 kd mismatches
s = SceneArrays.load(os.path.join(ROOT, 'scenes/scene03.mcpt'))
T.create_geometry(s.verts, s.normals, s.tris, s.mats)
osc = o.scene(s)
cam = o.cfg(512, 512, 1, 5, jitter=0, eye=(0, 5, 4.75), at=(0, 5, 3.75))
g = np.random.Generator(np.random.Philox(2))
lo, hi = s.verts.min(0), s.verts.max(0)
oo = g.uniform(lo, hi, (1 << 18, 3)); dd = g.normal(size=(1 << 18, 3)); dd /= np.linalg.norm(dd, axis=1, keepdims=True)
for name, rays in (('camera', o.primary_rays(cam)), ('random', np.concatenate([oo, dd], 1).astype(np.float32))):
    tri, t, b, gmm, _ = osc.intersect(rays)
    ktri, kt, kb, kg = T.intersect_batch(rays, T.ISECT_KD)
    bad = np.flatnonzero(tri != ktri)
    print(name, 'mismatches', len(bad), 'of', len(rays))
    for i in bad[:8]:
        print('  ray', rays[i], 'oracle', tri[i], t[i], 'kd', ktri[i], kt[i])
# 2. NEE bias structure: GPU vs GPU
s = SceneArrays.load(os.path.join(ROOT, 'scenes/scene01_lambert.mcpt'))
T.create_geometry(s.verts, s.normals, s.tris, s.mats)
W = H = 128
a, a2 = T.render_scene(T.render_cfg(W, H, 4096, 8, seed=1), moments=True)
b, b2 = T.render_scene(T.render_cfg(W, H, 4096, 8, seed=2, flags=T.FLAG_NEE), moments=True)
print('mean plain', a.mean(), 'mean nee', b.mean(), 'ratio', b.mean() / a.mean())
sig = np.sqrt((np.maximum(a2 - a * a, 0) + np.maximum(b2 - b * b, 0)) / 4096) + 1e-6
z = (b - a) / sig
print('z mean', z.mean(), 'frac |z|>4', (np.abs(z) > 4).mean())
for y in range(0, H, 16):
    print(' '.join('%+5.1f' % z[y:y + 16, x:x + 16, 0].mean() for x in range(0, W, 16)))
print('ratio rows (nee/plain) by 16x16 block')
for y in range(0, H, 16):
    print(' '.join('%5.2f' % (b[y:y + 16, x:x + 16].mean() / max(a[y:y + 16, x:x + 16].mean(), 1e-9)) for x in range(0, W, 16)))
np.savez_compressed(os.path.join(ROOT, 'gpurun_out/nee_debug.npz'), a=a, b=b)
