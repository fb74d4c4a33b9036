"""Freezes the oracle's outputs as committed fixtures (run HERE: python tests/golden/make_golden.py).

The reference ships no golden vectors, so the fixtures are the oracle's own results on fixed
inputs -- the oracle whose host half is pinned against the reference's code in oracle/_ref.  They let the GPU
box (no /root/reference) detect any drift of oracle, tree builder or kernels:
  golden.json   per scene: tree digest (reference-rule tree), digests of the hit records {tri, t, beta, gamma}
                for the 512x512 un-jittered camera batch and 2^18 Philox(1) random rays, work counters, and
                a 64x64 x 4-frame render in the reference's own RNG stream (sum digest + mean);
  samplers_ref.npz   outputs of THE REFERENCE'S OWN samplers (CUDA/Utils.hpp:46-137 compiled on the host,
                oracle/_ref/libmcpt_refcv.so) with the uniforms its cuRAND stream handed them: the oracle must
                reproduce them bit for bit, the CUDA samplers to a stated bound (tests/test_samplers.py);
  hits_scene01.npz   explicit records for 128x128 camera rays + 4096 random rays (compared bit for bit with
                the CUDA kernels in tests/test_gpu_parity.py).
"""
import hashlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from oracle.pyoracle import Oracle, SceneArrays  # noqa: E402

CAMERA_Z = {"scene01": 17.0, "scene01_lambert": 17.0, "scene02": 23.0, "scene03": 4.75}


def digest(*arrays):
    h = hashlib.sha256()
    for a in arrays:
        h.update(np.ascontiguousarray(a).tobytes())
    return h.hexdigest()


def random_rays(s, n, seed=1):
    g = np.random.Generator(np.random.Philox(seed))
    lo, hi = s.verts.min(0), s.verts.max(0)
    o = g.uniform(lo, hi, (n, 3))
    d = g.normal(size=(n, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    return np.concatenate([o, d], 1).astype(np.float32)


def scene_record(o, name):
    s = SceneArrays.load(os.path.join(ROOT, "scenes", name + ".mcpt"))
    sc = o.scene(s)
    t = sc.tree()
    z = CAMERA_Z[name]
    cam = o.cfg(512, 512, 1, 5, jitter=0, eye=(0, 5, z), at=(0, 5, z - 1))
    rec = {"triangles": int(len(s.tris)), "scene_digest": digest(s.verts, s.normals, s.tris, s.mats),
           "tree": {"nodes": int(len(t.left)), "leaves": int(t.num_leaves), "leaf_refs": int(len(t.leaf_ids)), "max_depth": int(t.max_depth),
                    "digest": digest(t.left, t.right, t.aabb, t.split_axis, t.split_value, t.leaf_offset, t.leaf_count, t.leaf_ids)}}
    for tag, rays in (("camera_512", o.primary_rays(cam)), ("random_2p18", random_rays(s, 1 << 18))):
        tri, tt, b, g, cnt = sc.intersect(rays)
        hit = tri >= 0
        rec[tag] = {"rays_digest": digest(rays), "hits": int(hit.sum()), "tri_digest": digest(tri),
                    "t_digest": digest(tt[hit]), "beta_gamma_digest": digest(b[hit], g[hit]),
                    "nodes_popped": int(cnt.nodes_popped), "tris_tested": int(cnt.tris_tested)}
    ssum, ssq, _, cnt = sc.render(o.cfg(64, 64, 4, 5, eye=(0, 5, z), at=(0, 5, z - 1)))
    rec["render_refstream_64x64x4"] = {"sum_digest": digest(ssum), "mean": float(ssum.mean() / 4), "rays": int(cnt.rays)}
    return rec, s, sc


def sampler_inputs(n, seed):
    """Unit normals / incoming directions incl. the rotation's special cases (|n.y +- 1| < FLT_EPSILON,
    rtx.hlsl:262-267 / :287-292), near-special normals, normal and grazing incidence."""
    g = np.random.default_rng(seed)
    nrm = g.normal(size=(n, 3))
    nrm /= np.linalg.norm(nrm, axis=1, keepdims=True)
    nrm = nrm.astype(np.float32)
    nrm[0], nrm[1] = [0, 1, 0], [0, -1, 0]
    nrm[2] = [0, np.float32(1) - np.float32(2.0 ** -24), 0]      # inside the epsilon
    nrm[3] = [3e-4, np.sqrt(1 - 9e-8), 0]                         # just outside it
    nrm[4] = [0, -1 + 2.0 ** -24, 1e-5]
    nrm[5], nrm[6], nrm[7] = [1, 0, 0], [0, 0, -1], [0.6, 0.0, 0.8]
    d = g.normal(size=(n, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    d = d.astype(np.float32)
    d[8] = -nrm[8]                                                # normal incidence
    t = np.cross(nrm[9], [0.3, 0.5, 0.8])
    t /= np.linalg.norm(t)
    d[9] = (t + 1e-3 * nrm[9]).astype(np.float32)                # grazing
    return nrm, d


def sampler_fixture(n=8192):
    """Runs the reference's own Utils.hpp samplers (needs oracle/_ref, i.e. /root/reference at build time)."""
    from oracle.pyoracle import ReferenceCV
    cv = ReferenceCV()
    g = np.random.default_rng(21)
    nrm, d = sampler_inputs(n, 20)
    Ns = g.choice([2, 5, 10, 20, 50, 1000], n).astype(np.uint32)
    Tr = g.choice([0.9, 0.5, 1.0, 0.1], n).astype(np.float32)
    Ni = g.choice([1.5, 1.33, 2.4, 1.0], n).astype(np.float32)
    hemi, u_hemi = cv.sample_hemi(1234, nrm)
    phong, u_phong = cv.sample_phong(99, nrm, d, Ns)
    fres, u_fres = cv.sample_fresnel(7, nrm, d, Tr, Ni)
    return dict(normals=nrm, dirs=d, Ns=Ns, Tr=Tr, Ni=Ni, hemi=hemi, u_hemi=u_hemi, phong=phong, u_phong=u_phong,
                fresnel_twin=fres, u_fresnel=u_fres)


def main():
    o = Oracle()
    if "--samplers" in sys.argv or not os.path.exists(os.path.join(HERE, "samplers_ref.npz")):
        np.savez_compressed(os.path.join(HERE, "samplers_ref.npz"), **sampler_fixture())
        if "--samplers" in sys.argv:
            return
    out = {}
    for name in ("scene01", "scene01_lambert", "scene02", "scene03"):
        out[name], s, sc = scene_record(o, name)
        if name == "scene01":
            rays = np.concatenate([o.primary_rays(o.cfg(128, 128, 1, 5, jitter=0)), random_rays(s, 4096, seed=9)])
            tri, t, b, g, _ = sc.intersect(rays)
            np.savez_compressed(os.path.join(HERE, "hits_scene01.npz"), rays=rays, tri=tri, t=t, beta=b, gamma=g)
        print(name, json.dumps(out[name]["tree"]))
    with open(os.path.join(HERE, "golden.json"), "w") as f:
        json.dump(out, f, indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
