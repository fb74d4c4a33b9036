"""BSDF samplers against the reference's OWN code.

tests/golden/samplers_ref.npz holds outputs of CVMCTracer/CVMCTracer/CUDA/Utils.hpp:46-137 -- the same math as
rtx.hlsl:213-302 -- compiled unmodified on the host (oracle/ref_bridge_cv.cpp), together with the uniforms its own
cuRAND stream handed each call (tests/golden/make_golden.py --samplers).  CPU: the oracle's restatement reproduces
them bit for bit (this pins the oracle's shading half on any machine, without /root/reference).  GPU: the device
samplers of csrc/shading.cuh, through mcpt_sample_batch, stay within stated bounds of them in both arithmetic forms.
"""
import os

import numpy as np
import pytest

from conftest import ROOT

Z = np.load(os.path.join(ROOT, "tests", "golden", "samplers_ref.npz"))
EPS = 2.0 ** -23   # one ulp of 1.0: directions are unit vectors, bounds are absolute errors per component


def _bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def test_oracle_reproduces_reference_sampler_outputs(oracle):
    n, d = Z["normals"], Z["dirs"]
    assert np.array_equal(_bits(oracle.sample_batch(oracle.HEMI, Z["u_hemi"], n)), _bits(Z["hemi"]))
    assert np.array_equal(_bits(oracle.sample_batch(oracle.PHONG, Z["u_phong"], n, d, Z["Ns"].astype(np.float32))), _bits(Z["phong"]))
    assert np.array_equal(_bits(oracle.sample_batch(oracle.FRESNEL_TWIN, Z["u_fresnel"], n, d, Z["Tr"], Z["Ni"])), _bits(Z["fresnel_twin"]))
    # the shader's form (normalize() on every branch, rtx.hlsl:250) differs from the twin's only by that normalisation
    hl = oracle.sample_batch(oracle.FRESNEL, Z["u_fresnel"], n, d, Z["Tr"], Z["Ni"])
    assert np.abs(hl - Z["fresnel_twin"]).max() < 4 * EPS


def test_fixture_covers_every_branch():
    n, d, u = Z["normals"], Z["dirs"], Z["u_fresnel"]
    assert (np.abs(n[:, 1] + 1) < 1.2e-7).sum() >= 2 and (np.abs(n[:, 1] - 1) < 1.2e-7).sum() >= 2   # both special cases of the rotation
    c = (d * n).sum(1)
    trp = Z["Tr"] * (1 - (1 - np.abs(c)) ** 5)
    refr = u < trp
    tir = refr & (c > 0) & (1 - (1 - c * c) * Z["Ni"] ** 2 < 0)
    assert (refr & (c <= 0)).sum() > 500 and (refr & (c > 0) & ~tir).sum() > 500 and tir.sum() > 100 and (~refr).sum() > 500
    assert set(np.unique(Z["Ns"])) == {2, 5, 10, 20, 50, 1000}


def _device(T, kind, fast):
    n, d = Z["normals"], Z["dirs"]
    if kind == T.SAMPLE_HEMI:
        return T.sample_batch(kind, Z["u_hemi"], n, render_arithmetic=fast), Z["hemi"]
    if kind == T.SAMPLE_PHONG:
        return T.sample_batch(kind, Z["u_phong"], n, d, Z["Ns"].astype(np.float32), render_arithmetic=fast), Z["phong"]
    return T.sample_batch(kind, Z["u_fresnel"], n, d, Z["Tr"], Z["Ni"], render_arithmetic=fast), Z["fresnel_twin"]


# Bounds (absolute, per component of a unit vector, in ulps of 1.0 = EPS), set from tools/sampler_probe.py on a B200
# (gpurun_out/sampler_probe.log, round 2) with ~2x margin.  Measured, IEEE form / render form:
#   hemi 1.5 / 3.5 (x 1/(1 - n.y^2) for the render form); phong Ns <= 50: 9 / 45; phong Ns = 1000: 15 / 365; fresnel 0 (bit-equal) / 36.
#   IEEE form (parity.cu, -fmad=false): the reference's expressions with round-to-nearest mul/add/div/sqrt; what differs is
#     CUDA's sincosf / powf against the host libm the reference build used (each <= 2 ulp of its own result).
#   render form (launch.cu, MCPT_FAST_SHADE, -fmad=true): MUFU.RSQ/SQRT/RCP (<= 2 ulp), sincospif, exp2(log2 x / (Ns+1)), FMA
#     contraction.  Three ill-conditioned spots are stated as such instead of hidden in one loose number:
#       * the rotation's 1 - n.y^2 (rtx.hlsl:269/294) cancels towards the poles: any change of rounding (one FMA -- the more
#         accurate form) moves the result by eps / (1 - n.y^2): the render-form bound carries that factor;
#       * samplePhong at Ns = 1000: sinT = sqrt(1 - cosT^2) with cosT = x^(1/1001) within 1e-3 of 1 amplifies an ulp of cosT
#         ~1/sinT^2-fold whichever implementation computes it (the reference's included);
#       * sampleFresnel next to total internal reflection: sqrt(test) with test -> 0.
BOUNDS = {  # kind -> (IEEE form, render form)
    "hemi": (4, 8), "phong": (20, 96), "phong_ns1000": (32, 768), "fresnel": (2, 64),
}


@pytest.mark.gpu
@pytest.mark.parametrize("fast", [False, True], ids=["ieee", "render_arithmetic"])
def test_device_samplers_match_reference_code(gpu, fast):
    T = gpu
    ny = Z["normals"][:, 1].astype(np.float64)
    special = np.abs(np.abs(ny) - 1) < 1.2e-7                    # the rotation's two exact special cases (rtx.hlsl:262-267)
    cond = np.where(special, 1.0, 1.0 / np.maximum(1 - ny * ny, 1e-9)) if fast else np.ones_like(ny)
    sharp = Z["Ns"] == 1000

    got, want = _device(T, T.SAMPLE_HEMI, fast)
    err = np.abs(got - want).max(1)
    assert (err <= BOUNDS["hemi"][fast] * EPS * cond).all(), (err / cond).max() / EPS
    assert err[special].max() <= BOUNDS["hemi"][fast] * EPS       # the special cases take the reference's branches

    got, want = _device(T, T.SAMPLE_PHONG, fast)
    err = np.abs(got - want).max(1)
    assert (err[~sharp] <= BOUNDS["phong"][fast] * EPS * cond[~sharp]).all(), (err[~sharp] / cond[~sharp]).max() / EPS
    assert (err[sharp] <= BOUNDS["phong_ns1000"][fast] * EPS * cond[sharp]).all(), (err[sharp] / cond[sharp]).max() / EPS

    got, want = _device(T, T.SAMPLE_FRESNEL_TWIN, fast)
    assert np.abs(got - want).max() <= BOUNDS["fresnel"][fast] * EPS, np.abs(got - want).max() / EPS
    # rtx.hlsl's own form (normalize() on every branch) against the oracle's restatement of it
    from oracle.pyoracle import Oracle
    o = Oracle()
    hl = o.sample_batch(o.FRESNEL, Z["u_fresnel"], Z["normals"], Z["dirs"], Z["Tr"], Z["Ni"])
    got = T.sample_batch(T.SAMPLE_FRESNEL, Z["u_fresnel"], Z["normals"], Z["dirs"], Z["Tr"], Z["Ni"], render_arithmetic=fast)
    assert np.abs(got - hl).max() <= BOUNDS["fresnel"][fast] * EPS, np.abs(got - hl).max() / EPS
    assert np.abs(np.linalg.norm(got.astype(np.float64), axis=1) - 1).max() < 4 * EPS


@pytest.mark.gpu
def test_render_path_camera_rays_within_ulps_of_the_exact_form(gpu, scenes):
    """The render kernels generate camera rays with one host-computed factor per axis and the hardware rsqrt
    (camera_ray<false>); mcpt_primary_rays' default form (camera_ray<true>) is the one that is bit-equal to the
    oracle.  Bound the distance between the two, for both cameras."""
    T = gpu
    for cfg in (T.render_cfg(1024, 1024, 1, 8, seed=5), T.cuda_twin_render_cfg(800, 600, 1, 1, seed=5)):
        exact = T.primary_rays(cfg, 3)
        cfg.flags |= T.FLAG_RENDER_ARITHMETIC
        fast = T.primary_rays(cfg, 3)
        assert np.abs(exact[:, :3] - fast[:, :3]).max() <= 4 * 2.0 ** -20          # origins: |coordinate| < 32
        assert np.abs(exact[:, 3:] - fast[:, 3:]).max() <= 4 * EPS                 # unit directions
        assert np.abs(np.linalg.norm(fast[:, 3:].astype(np.float64), axis=1) - 1).max() < 4 * EPS


def test_sample_batch_argument_checks(tracer):
    T = tracer
    u = np.zeros((4, 2), np.float32)
    n = np.zeros((4, 3), np.float32)
    with pytest.raises(T.McptError) as e:
        T.sample_batch(9, u, n)
    assert e.value.code == 1
    with pytest.raises(T.McptError) as e:
        T.sample_batch(T.SAMPLE_PHONG, u, n)       # no incoming directions
    assert e.value.code == 1
