"""Synthetic benchmark scenes (BASELINE.json configs[3], SURVEY.md 8d "K4"): a deterministic tessellated
surface inside the Cornell shell of the reference's scene01 -- KD-tree depth / traversal stress.

The shell (walls ids 0-9, light slab ids 430-441 of QuinEngine/Res/scene01.obj as tinyobj reads it)
is taken from the loaded base scene's own tables; the surface is the height field
    y = 2 + 0.8 sin(3x) cos(3z) + 0.05 sin(40x + 17z),  x in [-5, 5], z in [-4.5, 4.5]
sampled on an n x n quad grid (n = 708 -> 1 002 528 triangles), two triangles per quad, smooth vertex normals
from the analytic gradient, grey Lambert (Kd 0.5).
"""
from __future__ import annotations

import numpy as np

SHELL_TRIANGLES = tuple(range(0, 10)) + tuple(range(430, 442))


def cornell_heightfield(base, n: int = 708):
    """base = (vertices, normals, triangles, materials) of scene01 / scene01_lambert; returns the same four tables."""
    bv, bn, bt, bm = base
    shell = bt[list(SHELL_TRIANGLES)].copy()
    # compact the shell's vertices / normals
    vid, vinv = np.unique(shell[:, :3], return_inverse=True)
    nid, ninv = np.unique(shell[:, 3:6], return_inverse=True)
    shell[:, :3] = vinv.reshape(-1, 3)
    shell[:, 3:6] = ninv.reshape(-1, 3)
    sv, sn = bv[vid], bn[nid]

    xs = np.linspace(-5.0, 5.0, n + 1)
    zs = np.linspace(-4.5, 4.5, n + 1)
    X, Z = np.meshgrid(xs, zs, indexing="xy")
    Y = 2.0 + 0.8 * np.sin(3 * X) * np.cos(3 * Z) + 0.05 * np.sin(40 * X + 17 * Z)
    dYdx = 2.4 * np.cos(3 * X) * np.cos(3 * Z) + 2.0 * np.cos(40 * X + 17 * Z)
    dYdz = -2.4 * np.sin(3 * X) * np.sin(3 * Z) + 0.85 * np.cos(40 * X + 17 * Z)
    nrm = np.stack([-dYdx, np.ones_like(Y), -dYdz], -1)
    nrm /= np.linalg.norm(nrm, axis=-1, keepdims=True)
    hv = np.stack([X, Y, Z], -1).reshape(-1, 3).astype(np.float32)
    hn = nrm.reshape(-1, 3).astype(np.float32)

    idx = np.arange((n + 1) * (n + 1), dtype=np.int64).reshape(n + 1, n + 1)
    a, b, c, d = idx[:-1, :-1].ravel(), idx[:-1, 1:].ravel(), idx[1:, 1:].ravel(), idx[1:, :-1].ravel()
    # counter-clockwise seen from above (+y)
    tv = np.concatenate([np.stack([a, c, b], -1), np.stack([a, d, c], -1)], 0)
    grey = 1  # initialShadingGroup-like grey Lambert of the base scene (Kd 0.5)
    ht = np.empty((len(tv), 7), np.uint32)
    ht[:, :3] = tv + len(sv)
    ht[:, 3:6] = tv + len(sn)
    ht[:, 6] = grey
    return (np.concatenate([sv, hv]).astype(np.float32), np.concatenate([sn, hn]).astype(np.float32),
            np.concatenate([shell, ht]).astype(np.uint32), bm.copy())
