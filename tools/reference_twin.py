"""GPU box: runs THE REFERENCE'S OWN CUDA tracer (CUDA/CUTracer.cu compiled unmodified for sm_100a into
oracle/_ref/cutracer_ref, `make -C oracle refcu`) on an OBJ written from the repo's scene tables, and libmcpt's
MCPT_TRACER_CUDA_TWIN render of the same file (through the compiled drop-in shim oracle/_ref/dropin_main); prints the
comparison and freezes the reference's image as 4x4 tile means in tests/golden/twin_reference_scene<id>.npz (copied to
gpurun_out/ for the way back)."""
import json, os, subprocess, sys, tempfile, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests", "dropin"))
from objwriter import write_obj
from oracle.pyoracle import SceneArrays


def load_f32(path):
    with open(path, "rb") as f:
        w, h = np.fromfile(f, "<u4", 2)
        return np.fromfile(f, "<f4").reshape(h, w, 3)


out_dir = os.path.join(ROOT, "gpurun_out")
os.makedirs(out_dir, exist_ok=True)
report = {}
for scene_id, name in ((1, "scene01"), (2, "scene02")):
    s = SceneArrays.load(os.path.join(ROOT, "scenes", name + ".mcpt"))
    d = tempfile.mkdtemp()
    obj = os.path.join(d, name + ".obj")
    write_obj(obj, s.verts, s.normals, s.tris, s.mats)
    imgs = {}
    for tag, exe in (("reference", "cutracer_ref"), ("dropin", "dropin_main")):
        t0 = time.time()
        r = subprocess.run([os.path.join(ROOT, "oracle", "_ref", exe), obj, str(scene_id), os.path.join(d, tag + ".f32")], capture_output=True, text=True, timeout=1500)
        dt = time.time() - t0
        print(tag, name, "rc", r.returncode, r.stdout.strip(), r.stderr.strip()[-300:], f"{dt:.1f} s", flush=True)
        if r.returncode == 0:
            imgs[tag] = load_f32(os.path.join(d, tag + ".f32"))
            report[f"{name} {tag} seconds"] = dt
    if "reference" in imgs:
        ref = imgs["reference"].astype(np.float64)
        H, W, _ = ref.shape
        print(name, "reference image: finite", bool(np.isfinite(ref).all()), "mean", ref.mean(), "max", ref.max(), flush=True)
        tile = ref.reshape(H // 4, 4, W // 4, 4, 3).mean((1, 3)).astype(np.float32)
        for dst in (os.path.join(ROOT, "tests", "golden"), out_dir):
            np.savez_compressed(os.path.join(dst, f"twin_reference_scene{scene_id}.npz"), tile_mean=tile, width=W, height=H, spp=10000)
        if "dropin" in imgs:
            got = imgs["dropin"].astype(np.float64)
            rel = abs(got.mean() - ref.mean()) / ref.mean()
            gt = got.reshape(H // 4, 4, W // 4, 4, 3).mean((1, 3))
            bright = tile > 0.05 * tile.mean()
            trel = (np.abs(gt - tile)[bright] / tile[bright])
            report[name] = {"image_mean_reference": ref.mean(), "image_mean_mcpt": got.mean(), "mean_rel_err": rel,
                            "tile_rel_err_mean": float(trel.mean()), "tile_rel_err_p99": float(np.quantile(trel, 0.99)), "tile_rel_err_max": float(trel.max())}
            print(name, json.dumps(report[name]), flush=True)
json.dump(report, open(os.path.join(out_dir, "reference_twin.json"), "w"), indent=1)
