"""GPU box: max / quantile |delta| between the device samplers (both arithmetic forms) and the reference's own Utils.hpp
outputs frozen in tests/golden/samplers_ref.npz -- the numbers the bounds in tests/test_samplers.py are set from."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from montecarlopathtracer_b200 import tracer as T
from oracle.pyoracle import Oracle

z = np.load(os.path.join(ROOT, "tests", "golden", "samplers_ref.npz"))
T.initialize(0)
o = Oracle()
hl = o.sample_batch(o.FRESNEL, z["u_fresnel"], z["normals"], z["dirs"], z["Tr"], z["Ni"])
for fast in (False, True):
    for name, kind, u, par0, par1, want in (("hemi", T.SAMPLE_HEMI, z["u_hemi"], None, None, z["hemi"]),
                                           ("phong", T.SAMPLE_PHONG, z["u_phong"], z["Ns"].astype(np.float32), None, z["phong"]),
                                           ("fresnel", T.SAMPLE_FRESNEL, z["u_fresnel"], z["Tr"], z["Ni"], hl),
                                           ("fresnel_twin", T.SAMPLE_FRESNEL_TWIN, z["u_fresnel"], z["Tr"], z["Ni"], z["fresnel_twin"])):
        got = T.sample_batch(kind, u, z["normals"], z["dirs"], par0, par1, render_arithmetic=fast)
        d = np.abs(got.astype(np.float64) - want).max(1)
        eq = (got.view(np.uint32) == want.view(np.uint32)).all(1).mean()
        line = f"fast={int(fast)} {name:13s} bit-equal {eq:.4f} max {d.max():.3e} p99.9 {np.quantile(d, 0.999):.3e} median {np.median(d):.3e}"
        if name == "phong":
            for ns in (2, 50, 1000):
                m = z["Ns"] == ns
                line += f" | Ns={ns}: max {d[m].max():.2e}"
        print(line)

# raw outputs for offline calibration of the test bounds
out = {}
for fast in (False, True):
    out[f"hemi_{int(fast)}"] = T.sample_batch(T.SAMPLE_HEMI, z["u_hemi"], z["normals"], render_arithmetic=fast)
    out[f"phong_{int(fast)}"] = T.sample_batch(T.SAMPLE_PHONG, z["u_phong"], z["normals"], z["dirs"], z["Ns"].astype(np.float32), render_arithmetic=fast)
    out[f"fresnel_{int(fast)}"] = T.sample_batch(T.SAMPLE_FRESNEL, z["u_fresnel"], z["normals"], z["dirs"], z["Tr"], z["Ni"], render_arithmetic=fast)
    out[f"fresnel_twin_{int(fast)}"] = T.sample_batch(T.SAMPLE_FRESNEL_TWIN, z["u_fresnel"], z["normals"], z["dirs"], z["Tr"], z["Ni"], render_arithmetic=fast)
np.savez_compressed(os.path.join(ROOT, "gpurun_out", "sampler_outputs.npz"), **out)
