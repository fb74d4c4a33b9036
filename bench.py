#!/usr/bin/env python
"""bench.py -- headline benchmark of the path-sampling hot path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload K2|K1|K3a|K3b|K4|K5]

A "step" is one full render of the workload (one pass of the hot path over every pixel x sample).
N = 1 workload: K2 = Cornell box (scene01 geometry, all-Lambert + the area light), 1024x1024, 1024 spp, depth 8.
N > 1 (torchrun, one rank per GPU): spp-split -- the job is 1024*N spp per pixel (weak scaling), rank r renders
the r-th contiguous 1024-sample share with disjoint Philox streams and ONE ncclReduce over NVLink, issued by
libmcpt.so itself (mcpt_render_*_distributed), combines the accumulation buffers on rank 0.

Prints one JSON line (rank 0).  `value` = Mpaths/s with the scene resident in HBM; `e2e` = the same through the
reference-facing calls (create_geometry with HOST tables -> render_scene[_distributed] into a HOST buffer);
`roofline` and `cpu_baseline` as DESIGN.md section "Measurement" defines them; `other_workloads` carries the other
BASELINE configurations (K1, K3a, K3b, K4) at a step or two each, `k5_strong` (N > 1) the 4096^2 x 4096 spp job split
over the ranks, `multi_rank_equal` (N > 1) the on-GPU check that the R-rank image equals the 1-rank image.
"""
from __future__ import annotations

import argparse
import hashlib
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (scene file, width, height, spp, depth, eye z, description)
    "K1": ("scene01.mcpt", 512, 512, 16, 5, 17.0, "Cornell Box (repo scene01) 512x512, 16 spp, depth 5"),
    "K2": ("scene01_lambert.mcpt", 1024, 1024, 1024, 8, 17.0, "Cornell Box 1024x1024, 1024 spp, depth 8, Lambert + area light"),
    "K3a": ("scene01.mcpt", 1024, 1024, 2048, 8, 17.0, "Cornell Box, mirror (Ns 1000) + glass (Tr 0.9) spheres, 1024x1024, 2048 spp, depth 8"),
    "K3b": ("scene02.mcpt", 1024, 1024, 2048, 8, 23.0, "scene02: 4 sphere lights + glossy Phong slabs, 1024x1024, 2048 spp, depth 8"),
    "K4": ("heightfield:708", 1024, 1024, 256, 8, 17.0, "1 002 550-triangle height field in the Cornell shell, 1024x1024, 256 spp, depth 8"),
    "K5": ("scene01_lambert.mcpt", 4096, 4096, 4096, 8, 17.0, "Cornell Box 4096x4096, 4096 spp (use --scaling strong to split the 4096 spp over the ranks), depth 8"),
}

# Algorithmic work per unit, counted from the reference's math:
F_NODE, B_NODE = 22, 24          # slab test per popped node: flops, bytes (two float3)
F_TRI, B_TRI = 71, 40            # Cramer test: 68 flops + 3 divides; three float3 vertices + id
# nodes popped / triangles tested per ray by the REFERENCE's traversal of the REFERENCE's tree, from the oracle's counters on
# the full-size matched-spp renders (tests/golden/fullsize_*.json); only used when the cpu_baseline leg (which measures them
# live) is switched off.
ORACLE_WORK = {"K2": "fullsize_K2.json", "K3a": "fullsize_K3a.json", "K3b": "fullsize_K3b.json"}
KERNEL_SOURCES = ("render_sm.cuh", "render_pool.cuh", "render.cuh", "intersect.cuh", "shading.cuh", "launch.cu", "device_scene.h")


def kernel_source_hash() -> str:
    """Identity of the render kernel a number belongs to: the sources that define it plus the build script's flags.
    profiles/ncu_constants.json records the hash its ncu capture was taken with; constants from another kernel are not used."""
    h = hashlib.sha256()
    for f in KERNEL_SOURCES:
        h.update(open(os.path.join(ROOT, "montecarlopathtracer_b200", "csrc", f), "rb").read())
    h.update(open(os.path.join(ROOT, "montecarlopathtracer_b200", "build.py"), "rb").read())
    return h.hexdigest()[:16]


def ncu_constants(workload: str):
    """Counters that only ncu can give (warp instructions per ray, threads per instruction, issue slots busy, DRAM bytes per
    launch), written by tools/ncu_constants.py from a committed capture.  None unless that capture is of THIS kernel."""
    try:
        rec = json.load(open(os.path.join(ROOT, "profiles", "ncu_constants.json"))).get(workload)
    except (OSError, ValueError):
        return None
    if not rec or rec.get("kernel_source_hash") != kernel_source_hash():
        return None
    return rec


def load_scene(name):
    from oracle.pyoracle import SceneArrays  # only used by the cpu_baseline / reference legs below
    if name.startswith("heightfield:"):
        from montecarlopathtracer_b200 import scenes
        b = SceneArrays.load(os.path.join(ROOT, "scenes", "scene01_lambert.mcpt"))
        return SceneArrays(*scenes.cornell_heightfield((b.verts, b.normals, b.tris, b.mats), int(name.split(":")[1])))
    return SceneArrays.load(os.path.join(ROOT, "scenes", name))


def load_scene_gpu(T, name, build_mode):
    """Product-side scene set-up for the GPU arm (no oracle code)."""
    if name.startswith("heightfield:"):
        from montecarlopathtracer_b200 import scenes
        T.load_scene_file(os.path.join(ROOT, "scenes", "scene01_lambert.mcpt"), T.BUILD_SAH)
        T.create_geometry(*scenes.cornell_heightfield(T.get_scene(), int(name.split(":")[1])), build_mode)
    else:
        T.load_scene_file(os.path.join(ROOT, "scenes", name), build_mode)


class ClockSampler:
    """nvidia-smi clocks/throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int, period_ms: int = 200):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                          "-lms", str(period_ms)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), [x.strip() for x in line.split(",")]))

    def mark(self):
        return time.time()

    def window(self, t0: float, t1: float) -> dict:
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        rows = [r for t, r in self.rows if t0 <= t <= t1 and len(r) >= 7]
        if not rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        sm = [float(r[0]) for r in rows if r[0].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[3 + i].lower().startswith("active") for r in rows)]
        pw = [float(r[2]) for r in rows if r[2].replace(".", "").isdigit()]
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": float(rows[0][1]) if rows[0][1].isdigit() else None,
                "power_w_max": max(pw) if pw else None, "samples": len(rows), "reasons": reasons}

    def close(self):
        if self.proc:
            self.proc.terminate()


def host_cpu() -> dict:
    """CPU model, physical and logical core counts of this host (BASELINE.md section 3 asks for them beside the CPU rate)."""
    model, cores = None, set()
    try:
        phys = core = None
        for line in open("/proc/cpuinfo"):
            k, _, v = line.partition(":")
            k, v = k.strip(), v.strip()
            if k == "model name" and model is None:
                model = v
            elif k == "physical id":
                phys = v
            elif k == "core id":
                core = v
            elif not k and phys is not None:
                cores.add((phys, core)); phys = core = None
    except OSError:
        pass
    physical = len(cores) or None
    try:
        import psutil
        physical = psutil.cpu_count(logical=False) or physical
    except Exception:
        pass
    usable = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    return {"cpu_model": model, "physical_cores": physical, "logical_cpus": os.cpu_count(), "usable_cpus": usable}


def cpu_leg(workload, seconds_target=8.0, threads=0):
    """The reference's CPU path on the host cores: the oracle's restatement of rtx.hlsl driven by the
    reference's tree (built by the reference's own KDTree.hpp when oracle/_ref is present), on a bounded
    sample of the workload: full resolution, reduced spp (Mpaths/s is spp-independent)."""
    from oracle.pyoracle import Oracle, Reference
    scene_file, W, H, spp, depth, eyez, _ = WORKLOADS[workload]
    s = load_scene(scene_file)
    o = Oracle()
    kind, tree = "port", None
    if Reference.available() and len(s.tris) < 200000:   # the reference's own builder needs 25 s for K4's 1 M triangles
        ref = Reference()
        h = ref.from_arrays(s.verts, s.tris[:, :3])
        tree, _ = ref.build_tree(h)       # KDTree.hpp:58-287, the reference's own code
        ref.free(h)
        kind = "port"                     # the shader half is still a port: rtx.hlsl cannot run on this box
    sc = o.scene(s, tree)
    hc = host_cpu()
    ncores = threads or hc["usable_cpus"]
    probe = o.cfg(W, H, 1, depth, eye=(0, 5, eyez), at=(0, 5, eyez - 1), threads=ncores, y0=0, y1=max(1, H // 8), x0=0, x1=W)
    _, _, _, c = sc.render(probe)
    rate = c.paths / max(c.seconds, 1e-6)
    n = int(max(1, min(64, round(seconds_target * rate / (W * H)))))
    return sc, o, dict(W=W, H=H, depth=depth, eyez=eyez, spp=n, cores=ncores, kind=kind, host=hc,
                       tree="reference KDTree.hpp (oracle/_ref)" if tree is not None else "restated builder")


def run_cpu_sample(sc, o, p, threads=None, rows=None):
    kw = dict(eye=(0, 5, p["eyez"]), at=(0, 5, p["eyez"] - 1), threads=threads or p["cores"])
    if rows:
        kw.update(x0=0, x1=p["W"], y0=rows[0], y1=rows[1])
    cfg = o.cfg(p["W"], p["H"], p["spp"], p["depth"], **kw)
    _, _, _, c = sc.render(cfg)
    return c


def cpu_single_thread(sc, o, p, all_core_rate, seconds_target=3.0):
    """One host thread (the reference's own host code is single-threaded) on a band of rows in the middle of the image."""
    guess = all_core_rate / max(p["cores"], 1) * 1.3
    rows = int(max(1, min(p["H"], seconds_target * guess / (p["W"] * p["spp"]))))
    y0 = (p["H"] - rows) // 2
    c = run_cpu_sample(sc, o, p, threads=1, rows=(y0, y0 + rows))
    return {"value": c.paths / c.seconds / 1e6, "unit": "Mpaths/s", "sample": f"rows {y0}..{y0 + rows} of {p['W']}x{p['H']} at {p['spp']} spp"}


def reference_arm(args):
    """--impl reference: the reference's CPU implementation of the path (see cpu_leg), all host threads,
    each step a bounded sample of the same workload."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    _, W, H, spp, depth, _, desc = WORKLOADS[args.workload]
    sc, o, p = cpu_leg(args.workload, seconds_target=10.0)
    for _ in range(args.warmup):
        run_cpu_sample(sc, o, dict(p, spp=1))
    secs, paths, rays = 0.0, 0, 0
    for _ in range(args.steps):
        c = run_cpu_sample(sc, o, p)
        secs += c.seconds; paths += c.paths; rays += c.rays
    value = paths / secs / 1e6
    sample = f"{W}x{H} at {p['spp']} spp per step ({p['spp'] * W * H / 1e6:.1f} M paths), depth {depth}, {p['tree']}"
    line = {"impl": "reference", "metric": "Mpaths/s", "value": value, "unit": "Mpaths/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * secs / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "mrays_per_s": rays / secs / 1e6,
            "config": {"workload": desc, "sample": sample},
            "cpu_baseline": {"value": value, "unit": "Mpaths/s", "cores": p["cores"], "kind": p["kind"], "sample": sample, **p["host"],
                             "single_thread": cpu_single_thread(sc, o, p, paths / secs)},
            "e2e": {"value": value, "unit": "Mpaths/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)
    return 0


class Bench:
    """One workload on this rank's GPU through libmcpt.so; timing as the contract prescribes (CUDA events on the launching
    stream, barrier + synchronize on both sides, max over ranks, L2 flushed between timed steps)."""

    def __init__(self, T, torch, dist, rank, world, sampler, flush):
        self.T, self.torch, self.dist, self.rank, self.world, self.sampler, self.flush = T, torch, dist, rank, world, sampler, flush
        self.stream = torch.cuda.current_stream()

    def barrier(self):
        if self.dist is not None:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def allmax(self, *vals):
        t = self.torch.tensor(list(vals), device="cuda", dtype=self.torch.float64)
        if self.dist is not None:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return t.tolist()

    def allsum(self, *vals):
        t = self.torch.tensor(list(vals), device="cuda", dtype=self.torch.float64)
        if self.dist is not None:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM)
        return t.tolist()

    def run(self, workload, steps, warmup, e2e_steps, build_mode, flags=0, scaling="weak", spp_override=0, inner=1, e2e_warm=True):
        """-> dict of measurements.  `inner` renders per timed step (K1 is 3 ms: the step is repeated so the clock sampler sees it)."""
        T, torch = self.T, self.torch
        scene_file, W, H, spp, depth, eyez, desc = WORKLOADS[workload]
        if spp_override:
            spp = spp_override
        t_geo0 = time.perf_counter()
        load_scene_gpu(T, scene_file, build_mode)
        t_geo = time.perf_counter() - t_geo0
        info = T.tree_info()
        v, n, tr, m = T.get_scene()   # host tables for the e2e leg
        total_spp = spp * self.world if scaling == "weak" else spp
        job = T.render_cfg(W, H, total_spp, depth, seed=0, flags=flags, eye=(0, 5, eyez), at=(0, 5, eyez - 1))   # the WHOLE job
        my_off, my_spp = T.rank_share(total_spp, 0, self.rank, self.world)
        acc = torch.zeros(H, W, 4, device="cuda", dtype=torch.float32)

        def step():
            acc.zero_()
            # render this rank's share + the path's one exchange step (ncclReduce inside libmcpt.so, on this stream)
            T.render_accumulate_distributed_device(job, acc.data_ptr(), self.stream.cuda_stream)

        for _ in range(max(warmup, 0)):
            step()
        self.barrier()
        t0 = self.sampler.mark() if self.sampler else 0.0
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        kernel_ms, rays, launches = 0.0, 0, 0
        for a, b in ev:
            self.flush.zero_()           # flush L2 between timed iterations (outside the timed events)
            self.barrier()
            a.record(self.stream)
            for _ in range(inner):
                step()
                c = T.counters()        # the call is blocking: counters belong to this step
                kernel_ms += c.kernel_ms; rays += c.rays; launches += c.launches
            b.record(self.stream)
        self.barrier()
        t1 = self.sampler.mark() if self.sampler else 0.0
        ms = sum(a.elapsed_time(b) for a, b in ev)
        ms, kms = self.allmax(ms, kernel_ms)
        rays_all, = self.allsum(float(rays))
        renders = steps * inner
        paths_all = float(W) * H * total_spp * renders
        out = {"workload": desc + ("" if not spp_override else f" [spp overridden to {spp}]"), "value": paths_all / (ms * 1e-3) / 1e6,
               "unit": "Mpaths/s", "ms_per_step": ms / steps, "renders_per_step": inner, "steps": steps, "mrays_per_s": rays_all / (ms * 1e-3) / 1e6,
               "rays_per_path": rays_all / paths_all, "kernel_ms_per_render": kms / renders, "gpu_launches": int(launches),
               "spp_per_gpu": my_spp, "spp_total": total_spp, "tree_nodes": info.num_nodes, "scene_bytes_hbm": int(info.device_bytes),
               "clocks": self.sampler.window(t0, t1) if self.sampler else None,
               "_rays_per_launch": rays_all / self.world / renders, "_t_launch": kms / renders * 1e-3, "_W": W, "_H": H}

        # ---- e2e: the reference-facing calls with HOST buffers, host<->device copies inside the timed region ----
        host = np.empty((H, W, 3), np.float32) if self.rank == 0 else None
        split = {"create_geometry_ms": 0.0, "kd_build_ms": 0.0, "render_call_ms": 0.0, "render_kernel_ms": 0.0}

        def e2e_step(record):
            ta = time.perf_counter()
            T.create_geometry(v, n, tr, m, build_mode)     # CreateGeometry: host tables -> KD build -> pack -> H2D
            tb = time.perf_counter()
            T.render_scene_distributed(job, host)          # RenderScene: share -> ncclReduce -> mean on the device -> ONE float3 D2H on rank 0
            tc = time.perf_counter()
            if record:
                split["create_geometry_ms"] += 1e3 * (tb - ta); split["kd_build_ms"] += 1e3 * T.tree_info().build_seconds
                split["render_call_ms"] += 1e3 * (tc - tb); split["render_kernel_ms"] += T.counters().kernel_ms

        if e2e_warm:
            e2e_step(False)
        self.barrier()
        ta = time.perf_counter()
        for _ in range(e2e_steps * inner):
            e2e_step(True)
        torch.cuda.synchronize()
        e2e_s, = self.allmax(time.perf_counter() - ta)
        k = e2e_steps * inner
        # host -> device per step: with the device builder the caller's tables go up and everything else is made there; with a host
        # builder the packed render-path tables go up
        device_made = build_mode == T.BUILD_GPU and T.get_tuning("device_pack") != 0
        h2d = int(v.nbytes + n.nbytes + tr.nbytes + m.nbytes) if device_made else int(info.device_bytes)
        out["e2e"] = {"value": float(W) * H * total_spp * k / e2e_s / 1e6, "unit": "Mpaths/s", "h2d_bytes_per_step": h2d,
                      "h2d_what": "the caller's scene tables (tree and render-path tables are built on the device)" if device_made else "packed render-path tables (built on the host)",
                      "d2h_bytes_per_step": W * H * 3 * 4, "steps": e2e_steps, "timer": "host wall clock (perf_counter) around blocking C-ABI calls, max over ranks",
                      "includes": "mcpt_create_geometry (KD build + pack + H2D) and mcpt_render_scene_distributed (render, ncclReduce, D2H of the float3 image) every step",
                      "split_ms_per_render": {a: b / k for a, b in split.items()}}
        out["e2e"]["split_ms_per_render"]["pack_and_upload_ms"] = out["e2e"]["split_ms_per_render"]["create_geometry_ms"] - out["e2e"]["split_ms_per_render"]["kd_build_ms"]
        out["first_load_s"] = t_geo
        return out


def roofline(T, m, nn, nt, clocks, peaks, workload):
    """Roofline of render_kernel_sm.  The scene's node and triangle records are served by L1 (K1-K3, K5) or L2 (K4); HBM sees only
    the accumulation slices.  What binds is FP32 / instruction issue at the kernel's SIMT efficiency, so the headline fraction is
    ALGORITHMIC flops (the reference tree's work per ray x the reference's flop counts) over the measured FP32 FMA
    peak; the issue-slot view (ncu) and the byte view sit beside it under their own names."""
    rays, t = m["_rays_per_launch"], m["_t_launch"]
    aflops = rays * (nn * F_NODE + nt * F_TRI)
    abytes = rays * (nn * B_NODE + nt * B_TRI) + m["_W"] * m["_H"] * 16.0
    fp32_peak = T.measure_fp32_peak()    # GFLOP/s, fp32_peak_kernel on this device (MEASURED_PEAKS.json has no FP32 figure)
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    nc = ncu_constants(workload)
    sm_mhz = (clocks or {}).get("sm_mhz") or 1965.0
    r = {"bound": "fp32_issue", "achieved": aflops / t / 1e12, "peak": fp32_peak / 1e3, "unit": "TFLOP/s", "frac": aflops / t / 1e9 / fp32_peak,
         "peak_source": "mcpt_measure_fp32_peak (fp32_peak_kernel: 8 independent FMA chains per thread, measured in this run)",
         "kernel": "render_kernel_sm", "launch_ms": t * 1e3, "rays_per_launch": rays,
         "algorithmic_flops_per_ray": nn * F_NODE + nt * F_TRI, "algorithmic_bytes_per_ray": nn * B_NODE + nt * B_TRI,
         "reference_tree_nodes_per_ray": nn, "reference_tree_tris_per_ray": nt,
         "traffic": nc["dram_bytes_per_launch"] if nc else None,
         "hbm": {"mandatory_bytes_per_launch": m["_W"] * m["_H"] * 16.0, "measured_dram_bytes_per_launch": nc["dram_bytes_per_launch"] if nc else None,
                 "achieved_gbs": (nc["dram_bytes_per_launch"] / t / 1e9) if nc else None, "peak_gbs": hbm_peak,
                 "peak_source": "MEASURED_PEAKS.json" if peaks else "fallback"},
         "l1_delivered": {"algorithmic_gbs": abytes / t / 1e9, "note": "reference bytes per ray x rays / s: served by L1/L2, NOT an HBM figure"},
         "issue": ({"warp_inst_per_ray": nc["warp_inst_per_ray"], "threads_per_inst": nc["threads_per_inst"], "issue_slots_busy": nc["issue_slots_busy"],
                    "achieved_ginst_s": nc["warp_inst_per_ray"] * rays / t / 1e9, "peak_ginst_s": 148 * 4 * sm_mhz / 1e3,
                    "frac": nc["warp_inst_per_ray"] * rays / t / 1e9 / (148 * 4 * sm_mhz / 1e3),
                    "lane_issue_frac": nc["warp_inst_per_ray"] * rays / t / 1e9 / (148 * 4 * sm_mhz / 1e3) * nc["threads_per_inst"] / 32.0,
                    "source": nc["source"], "kernel_source_hash": nc["kernel_source_hash"]} if nc else None),
         "kernel_source_hash": kernel_source_hash(),
         "note": "bound = FP32/instruction issue x SIMT efficiency; `issue` is null when profiles/ncu_constants.json was not captured from this exact kernel"}
    return r


def ncu_view(workload, m, peaks):
    """What only ncu can tell about a workload's render kernel (profiles/ncu_constants.json, keyed by the kernel-source hash; None
    for any other kernel) next to this run's own launch time: achieved DRAM GB/s against the measured HBM peak, cache hit rates,
    SIMT efficiency, issue-slot use.  For K4 -- the one workload whose scene outgrows the caches -- this is the memory-side figure."""
    nc = ncu_constants(workload)
    if not nc:
        return None
    t = m["_t_launch"]
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    return {"source": nc["source"], "kernel_source_hash": nc["kernel_source_hash"], "dram_bytes_per_launch": nc["dram_bytes_per_launch"],
            "dram_gbs": nc["dram_bytes_per_launch"] / t / 1e9, "hbm_peak_gbs": hbm_peak, "hbm_frac": nc["dram_bytes_per_launch"] / t / 1e9 / hbm_peak,
            "l1_hit_rate": nc["l1_hit_rate"], "l2_hit_rate": nc["l2_hit_rate"], "warp_inst_per_ray": nc["warp_inst_per_ray"],
            "threads_per_inst": nc["threads_per_inst"], "issue_slots_busy": nc["issue_slots_busy"], "registers_per_thread": nc["registers_per_thread"]}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="mcpt", choices=["mcpt", "reference"])
    ap.add_argument("--workload", default="K2", choices=sorted(WORKLOADS))
    ap.add_argument("--spp", type=int, default=0, help="override the workload's spp (a reduced-spp run is NOT the headline number)")
    ap.add_argument("--flags", default="", help="comma list: nee,ref_traversal")
    ap.add_argument("--tree", default="auto", choices=["auto", "sah", "gpu", "reference"],
                    help="KD builder for the render path (same hits whichever): auto = host SAH, device SAH builder for K4's 1 M triangles")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip other_workloads / k5_strong / multi_rank_equal")
    ap.add_argument("--tune", action="append", default=[], help="knob=value through mcpt_set_tuning (recorded in config)")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="weak: every rank renders the workload's spp (default, the driver's contract); strong: the spp are split")
    args = ap.parse_args()
    if args.impl == "reference":
        return reference_arm(args)

    import torch
    from montecarlopathtracer_b200 import multigpu as M
    from montecarlopathtracer_b200 import tracer as T

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the tracer has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))   # barriers and the max-over-ranks of the timings

    flags = 0
    for f in filter(None, args.flags.split(",")):
        flags |= {"nee": T.FLAG_NEE, "ref_traversal": T.FLAG_REF_TRAVERSAL}[f]
    T.initialize(local)
    tuned = {}
    for kv in args.tune:
        k, _, x = kv.partition("=")
        T.set_tuning(k, float(x)); tuned[k] = T.get_tuning(k)
    M.comm_init_from_torch(T, dist, rank, world)   # libmcpt's own NCCL communicator; torch only carries the 128-byte id
    def mode_for(workload):
        if args.tree == "auto":
            return T.BUILD_GPU if workload == "K4" else T.BUILD_SAH
        return {"sah": T.BUILD_SAH, "gpu": T.BUILD_GPU, "reference": T.BUILD_REFERENCE}[args.tree]
    build_mode = mode_for(args.workload)
    tree_name = {T.BUILD_SAH: "own SAH builder on the host (MCPT_BUILD_SAH)", T.BUILD_GPU: "own SAH builder on the device (MCPT_BUILD_GPU)",
                 T.BUILD_REFERENCE: "reference-rule tree (MCPT_BUILD_REFERENCE)"}
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")   # > 126 MB L2
    sampler = ClockSampler(local) if rank == 0 else None
    B = Bench(T, torch, dist, rank, world, sampler, flush)

    main_inner = 100 if args.workload == "K1" else 1
    m = B.run(args.workload, args.steps, args.warmup, max(1, min(args.steps, 3)), build_mode, flags, args.scaling, args.spp, inner=main_inner)

    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except OSError:
        pass
    extras = {}
    if not args.no_extras and args.workload == "K2" and not args.spp:
        if world == 1:
            others = []
            for w, inner in (("K1", 100), ("K3a", 1), ("K3b", 1), ("K4", 1)):
                o = B.run(w, 2, 1, 2 if w == "K4" else 1, mode_for(w), flags, "weak", inner=inner)
                o["tree"] = tree_name[mode_for(w)]
                o["ncu"] = ncu_view(w, o, peaks)
                others.append({k: v for k, v in o.items() if not k.startswith("_")})
            extras["other_workloads"] = others
        else:
            # BASELINE configs[4]: 4096^2 x 4096 spp, the 4096 samples SPLIT over the ranks (strong), 256 MiB reduce
            k5 = B.run("K5", 1, 0, 1, build_mode, flags, "strong", e2e_warm=False)
            extras["k5_strong"] = {k: v for k, v in k5.items() if not k.startswith("_")}
            # the R-rank reduced image equals the 1-rank image up to fp32 reassociation -- on the GPUs, through the ABI
            load_scene_gpu(T, "scene01_lambert.mcpt", build_mode)
            cfg = T.render_cfg(256, 256, 64, 8, seed=7)
            split_img = T.render_scene_distributed(cfg)
            if rank == 0:
                whole = T.render_scene(T.render_cfg(256, 256, 64, 8, seed=7))
                extras["multi_rank_equal"] = bool(np.allclose(split_img, whole, rtol=2e-5, atol=1e-7))
                extras["multi_rank_max_abs_diff"] = float(np.abs(split_img - whole).max())
            B.barrier()

    if rank != 0:
        T.comm_destroy()
        if dist is not None:
            dist.destroy_process_group()
        return 0

    _, W, H, spp, depth, eyez, desc = WORKLOADS[args.workload]
    line = {"metric": "Mpaths/s", "value": m["value"], "unit": "Mpaths/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": m["ms_per_step"], "higher_is_better": True, "scaling": args.scaling,
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "mrays_per_s": m["mrays_per_s"], "rays_per_path": m["rays_per_path"],
            "config": {"workload": m["workload"], "spp_per_gpu": m["spp_per_gpu"], "spp_total": m["spp_total"],
                       "parallelism": f"spp-split x{world}, one ncclReduce inside libmcpt.so (NCCL {T.comm_info()[2]})" if world > 1 else "1 GPU",
                       "estimator": "NEE" if flags & T.FLAG_NEE else "emission-on-hit (reference)", "rng": "philox4x32-10",
                       "l2": "flushed between timed steps (256 MiB memset); scene is L1/L2-resident by design",
                       "tree": tree_name[build_mode],
                       "tree_nodes": m["tree_nodes"], "scene_bytes_hbm": m["scene_bytes_hbm"], "renders_per_step": m["renders_per_step"],
                       "tuning": tuned or "shipped defaults"},
            "e2e": m["e2e"], "gpu_launches": m["gpu_launches"], "clocks": m["clocks"]}

    cpu, nn, nt = None, None, None
    if not args.no_cpu_baseline:
        sc, o, p = cpu_leg(args.workload, seconds_target=8.0)
        c = run_cpu_sample(sc, o, p)
        sample = f"{p['W']}x{p['H']} at {p['spp']} spp ({c.paths / 1e6:.1f} M paths), depth {p['depth']}, {p['tree']}"
        cpu = {"value": c.paths / c.seconds / 1e6, "unit": "Mpaths/s", "cores": p["cores"], "kind": p["kind"], "sample": sample, **p["host"],
               "mrays_per_s": c.rays / c.seconds / 1e6, "nodes_per_ray": c.nodes_popped / c.rays, "tris_per_ray": c.tris_tested / c.rays,
               "single_thread": cpu_single_thread(sc, o, p, c.paths / c.seconds)}
        nn, nt = cpu["nodes_per_ray"], cpu["tris_per_ray"]
    elif args.workload in ORACLE_WORK:
        try:
            meta = json.load(open(os.path.join(ROOT, "tests", "golden", ORACLE_WORK[args.workload])))
            nn, nt = meta["nodes_per_ray"], meta["tris_per_ray"]
        except OSError:
            pass
    line["cpu_baseline"] = cpu
    line["ncu"] = ncu_view(args.workload, m, peaks)
    line["roofline"] = roofline(T, m, nn, nt, m["clocks"], peaks, args.workload) if nn is not None else None
    line.update(extras)
    if sampler:
        sampler.close()
    print(json.dumps(line), flush=True)
    T.comm_destroy()
    if dist is not None:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
