import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

SCENES = os.path.join(ROOT, "scenes")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def oracle():
    from oracle.pyoracle import Oracle
    return Oracle()


@pytest.fixture(scope="session")
def reference():
    """The reference's own host code compiled in oracle/_ref (absent -> skip)."""
    from oracle.pyoracle import Reference, build
    build()
    if not Reference.available():
        pytest.skip("oracle/_ref/libmcpt_ref.so not built (needs /root/reference)")
    return Reference()


@pytest.fixture(scope="session")
def scenes():
    from oracle.pyoracle import SceneArrays
    return {n: SceneArrays.load(os.path.join(SCENES, n + ".mcpt"))
            for n in ("scene01", "scene01_lambert", "scene02", "scene03")}


@pytest.fixture(scope="session")
def tracer():
    """libmcpt.so through its C ABI (host side only; no device is bound here)."""
    from montecarlopathtracer_b200 import build, tracer as T
    build.build()
    return T


@pytest.fixture(scope="session")
def gpu(tracer):
    """libmcpt.so bound to cuda:0.  Fails (not skips) when the library cannot reach the GPU."""
    tracer.initialize(0)
    return tracer


def heightfield(n, seed=0):
    """Deterministic tessellated surface in the Cornell shell, n x n quads."""
    xs = np.linspace(-5.0, 5.0, n + 1, dtype=np.float64)
    zs = np.linspace(-4.5, 4.5, n + 1, dtype=np.float64)
    X, Z = np.meshgrid(xs, zs, indexing="xy")
    Y = 2.0 + 0.8 * np.sin(3 * X) * np.cos(3 * Z) + 0.05 * np.sin(40 * X + 17 * Z)
    verts = np.stack([X, Y, Z], -1).reshape(-1, 3).astype(np.float32)
    idx = np.arange((n + 1) * (n + 1)).reshape(n + 1, n + 1)
    a, b, c, d = idx[:-1, :-1].ravel(), idx[:-1, 1:].ravel(), idx[1:, 1:].ravel(), idx[1:, :-1].ravel()
    tri_v = np.concatenate([np.stack([a, b, c], -1), np.stack([a, c, d], -1)], 0).astype(np.uint32)
    return verts, tri_v
