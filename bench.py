#!/usr/bin/env python
"""bench.py -- headline benchmark of the path-sampling hot path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload K2|K1|K3a|K3b]

A "step" is one full render of the workload (one pass of the hot path over every pixel x sample).
N = 1 workload: K2 = Cornell box (scene01 geometry, all-Lambert + the area light), 1024x1024,
1024 spp, depth 8.  N > 1 (torchrun, one rank per GPU): spp-split -- every rank renders its own
1024-sample range of every pixel with disjoint Philox streams (weak scaling: the job's total is
1024*N spp) and ONE NCCL reduce over NVLink combines the accumulation buffers on rank 0.

Prints one JSON line (rank 0).  `value` = Mpaths/s with the scene resident in HBM; `e2e` = the same
through the reference-facing calls (create_geometry with HOST tables -> render_scene into a HOST
buffer); `roofline` and `cpu_baseline` as DESIGN.md section "Measurement" defines them.
"""
from __future__ import annotations

import argparse
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

    def __init__(self, index: int):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                          "-lms", "200"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), [x.strip() for x in line.split(",")]))

    def stop(self, t0: float, t1: float) -> dict:
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        rows = [r for t, r in self.rows if t0 <= t <= t1 and len(r) >= 7] or [r for _, r in self.rows if len(r) >= 7]
        if not rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        sm = [float(r[0]) for r in rows if r[0].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[3 + i].lower().startswith("active") for r in rows)]
        pw = [float(r[2]) for r in rows if r[2].replace(".", "").isdigit()]
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": float(rows[0][1]) if rows[0][1].isdigit() else None,
                "power_w_max": max(pw) if pw else None, "samples": len(rows), "reasons": reasons}


def cpu_leg(workload, seconds_target=8.0, threads=0):
    """The reference's CPU path on the host cores: the oracle's restatement of rtx.hlsl driven by the
    reference's tree (built by the reference's own KDTree.hpp when oracle/_ref is present), on a bounded
    sample of the workload: full resolution, reduced spp (Mpaths/s is spp-independent)."""
    from oracle.pyoracle import Oracle, Reference
    scene_file, W, H, spp, depth, eyez, _ = WORKLOADS[workload]
    s = load_scene(scene_file)
    o = Oracle()
    kind, tree = "port", None
    if Reference.available():
        ref = Reference()
        h = ref.from_arrays(s.verts, s.tris[:, :3])
        tree, _ = ref.build_tree(h)       # KDTree.hpp:58-287, the reference's own code
        ref.free(h)
        kind = "port"                     # the shader half is still a port: rtx.hlsl cannot run on this box
    sc = o.scene(s, tree)
    ncores = threads or (os.cpu_count() or 1)
    probe = o.cfg(W, H, 1, depth, eye=(0, 5, eyez), at=(0, 5, eyez - 1), threads=ncores, y0=0, y1=max(1, H // 8), x0=0, x1=W)
    _, _, _, c = sc.render(probe)
    rate = c.paths / max(c.seconds, 1e-6)
    n = int(max(1, min(64, round(seconds_target * rate / (W * H)))))
    return sc, o, dict(W=W, H=H, depth=depth, eyez=eyez, spp=n, cores=ncores, kind=kind,
                       tree="reference KDTree.hpp (oracle/_ref)" if tree is not None else "restated builder")


def run_cpu_sample(sc, o, p):
    cfg = o.cfg(p["W"], p["H"], p["spp"], p["depth"], eye=(0, 5, p["eyez"]), at=(0, 5, p["eyez"] - 1), threads=p["cores"])
    _, _, _, c = sc.render(cfg)
    return c


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
            "cpu_baseline": {"value": value, "unit": "Mpaths/s", "cores": p["cores"], "kind": p["kind"], "sample": sample},
            "e2e": {"value": value, "unit": "Mpaths/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="mcpt", choices=["mcpt", "reference"])
    ap.add_argument("--workload", default="K2", choices=sorted(WORKLOADS))
    ap.add_argument("--spp", type=int, default=0, help="override the workload's spp (a reduced-spp run is NOT the headline number)")
    ap.add_argument("--flags", default="", help="comma list: nee,ref_traversal")
    ap.add_argument("--tree", default="sah", choices=["sah", "reference"], help="KD builder for the render path (same hits either way)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="weak: every rank renders the workload's spp (default, the driver's contract); strong: the spp are split")
    args = ap.parse_args()
    if args.impl == "reference":
        return reference_arm(args)

    import torch
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
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    scene_file, W, H, spp, depth, eyez, desc = WORKLOADS[args.workload]
    if args.spp:
        spp = args.spp
    flags = 0
    for f in filter(None, args.flags.split(",")):
        flags |= {"nee": T.FLAG_NEE, "ref_traversal": T.FLAG_REF_TRAVERSAL}[f]

    T.initialize(local)
    build_mode = T.BUILD_REFERENCE if args.tree == "reference" else T.BUILD_SAH
    load_scene_gpu(T, scene_file, build_mode)
    info = T.tree_info()
    v, n, tr, m = T.get_scene()   # host tables for the e2e leg

    stream = torch.cuda.current_stream()
    acc = torch.zeros(H, W, 4, device="cuda", dtype=torch.float32)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")   # > 126 MB L2
    from montecarlopathtracer_b200 import multigpu as M
    my_off, my_spp = M.rank_samples(rank, world, spp, args.scaling)
    total_spp = spp * world if args.scaling == "weak" else spp
    cfg = T.render_cfg(W, H, my_spp, depth, seed=0, sample_offset=my_off, flags=flags,
                       eye=(0, 5, eyez), at=(0, 5, eyez - 1))

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def step():
        acc.zero_()
        T.render_accumulate_device(cfg, acc.data_ptr(), 0, stream.cuda_stream)
        M.reduce_accumulators(acc, dist)   # the one exchange step of the path: NCCL sum over NVLink

    for _ in range(max(args.warmup, 0)):
        step()
    barrier()

    sampler = ClockSampler(local) if rank == 0 else None
    t_wall0 = time.time()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    kernel_ms, rays, launches = [], 0, 0
    for a, b in ev:
        flush.zero_()           # flush L2 between timed iterations (outside the timed events)
        barrier()
        a.record(stream)
        step()
        b.record(stream)
        c = T.counters()        # render_accumulate_device is blocking: counters belong to this step
        kernel_ms.append(c.kernel_ms); rays += c.rays; launches += c.launches
    barrier()
    t_wall1 = time.time()
    ms = sum(a.elapsed_time(b) for a, b in ev)
    tot = torch.tensor([ms, float(rays), sum(kernel_ms)], device="cuda", dtype=torch.float64)
    if dist is not None:
        mx = tot.clone(); dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        sm = tot.clone(); dist.all_reduce(sm, op=dist.ReduceOp.SUM)
        ms, rays_all, kms = mx[0].item(), sm[1].item(), mx[2].item()
    else:
        rays_all, kms = float(rays), sum(kernel_ms)
    clocks = sampler.stop(t_wall0, t_wall1) if sampler else None
    paths_all = float(W) * H * total_spp * args.steps
    value = paths_all / (ms * 1e-3) / 1e6
    mrays = rays_all / (ms * 1e-3) / 1e6

    # ---- e2e: the reference-facing calls with HOST buffers, copies inside the timed region ----
    host = np.empty((H, W, 3), np.float32)
    e2e_cfg = T.render_cfg(W, H, my_spp, depth, seed=0, sample_offset=my_off, flags=flags, eye=(0, 5, eyez), at=(0, 5, eyez - 1))
    e2e_steps = max(1, min(args.steps, 2))
    pinned = torch.empty(H, W, 4, dtype=torch.float32).pin_memory() if world > 1 else None

    def e2e_step():
        T.create_geometry(v, n, tr, m, build_mode)     # CreateGeometry: host tables -> tree build -> H2D
        if world == 1:
            T.render_scene(e2e_cfg, host)              # RenderScene: kernel -> mean -> D2H of the image
        else:                                          # N ranks: render share -> one NCCL reduce -> D2H on rank 0
            acc.zero_()
            T.render_accumulate_device(e2e_cfg, acc.data_ptr(), 0, stream.cuda_stream)
            M.reduce_accumulators(acc, dist)
            if rank == 0:
                pinned.copy_(acc, non_blocking=False)
                host[...] = M.finalize(pinned)

    e2e_step()      # warm
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_step()
    torch.cuda.synchronize()
    e2e_s = torch.tensor([time.perf_counter() - t0], device="cuda", dtype=torch.float64)
    if dist is not None:
        dist.all_reduce(e2e_s, op=dist.ReduceOp.MAX)
    e2e_value = float(W) * H * total_spp * e2e_steps / e2e_s.item() / 1e6

    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return 0

    # ---- roofline of the dominant kernel (render_kernel), per launch ----
    line = {"metric": "Mpaths/s", "value": value, "unit": "Mpaths/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": args.scaling,
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "mrays_per_s": mrays, "rays_per_path": rays_all / paths_all,
            "config": {"workload": desc + ("" if not args.spp else f" [spp overridden to {spp}]"),
                       "spp_per_gpu": my_spp, "spp_total": total_spp, "parallelism": f"spp-split x{world}, one NCCL reduce" if world > 1 else "1 GPU",
                       "estimator": "NEE" if flags & T.FLAG_NEE else "emission-on-hit (reference)", "rng": "philox4x32-10",
                       "l2": "flushed between timed steps (256 MiB memset); scene is L1/L2-resident by design",
                       "tree": "own SAH builder (MCPT_BUILD_SAH)" if build_mode == T.BUILD_SAH else "reference-rule tree (MCPT_BUILD_REFERENCE)",
                       "tree_nodes": info.num_nodes, "scene_bytes_hbm": int(info.device_bytes)},
            "e2e": {"value": e2e_value, "unit": "Mpaths/s", "h2d_bytes_per_step": int(info.device_bytes),
                    "d2h_bytes_per_step": W * H * (3 if world == 1 else 4) * 4, "steps": e2e_steps,
                    "includes": "mcpt_create_geometry (host KD build + H2D) and the image D2H every step"},
            "gpu_launches": int(launches), "clocks": clocks}

    cpu = None
    if not args.no_cpu_baseline:
        sc, o, p = cpu_leg(args.workload, seconds_target=8.0)
        c = run_cpu_sample(sc, o, p)
        sample = f"{p['W']}x{p['H']} at {p['spp']} spp ({c.paths / 1e6:.1f} M paths), depth {p['depth']}, {p['tree']}"
        cpu = {"value": c.paths / c.seconds / 1e6, "unit": "Mpaths/s", "cores": p["cores"], "kind": p["kind"], "sample": sample,
               "mrays_per_s": c.rays / c.seconds / 1e6, "nodes_per_ray": c.nodes_popped / c.rays, "tris_per_ray": c.tris_tested / c.rays}
        nn, nt = cpu["nodes_per_ray"], cpu["tris_per_ray"]
    else:
        nn, nt = 32.1, 5.07   # oracle counters on scene01 (recorded in DESIGN.md)
    line["cpu_baseline"] = cpu
    rays_per_launch = rays_all / world / args.steps
    t_launch = kms / args.steps * 1e-3
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except OSError:
        pass
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    abytes = rays_per_launch * (nn * B_NODE + nt * B_TRI) + W * H * 16.0
    aflops = rays_per_launch * (nn * F_NODE + nt * F_TRI)
    fp32_peak = T.measure_fp32_peak()
    line["roofline"] = {"bound": "hbm", "achieved": abytes / t_launch / 1e9, "peak": hbm_peak, "unit": "GB/s",
                        "frac": abytes / t_launch / 1e9 / hbm_peak,
                        # dram__bytes_read.sum + dram__bytes_write.sum of ONE render_kernel_sm launch at the K2 size
                        # (profiles/r1i_ncu_full_render_kernel_sm_K2_raw.csv; the four 16 MB accumulation slices mostly stay in L2); other workloads were not captured
                        "traffic": 23419904 if (args.workload == "K2" and not args.spp) else None,
                        "peak_source": "MEASURED_PEAKS.json" if peaks else "fallback",
                        "kernel": "render_kernel_sm", "launch_ms": t_launch * 1e3,
                        "algorithmic_bytes_per_ray": nn * B_NODE + nt * B_TRI, "algorithmic_flops_per_ray": nn * F_NODE + nt * F_TRI,
                        "fp32": {"achieved_gflops": aflops / t_launch / 1e9, "peak_gflops": fp32_peak, "frac": aflops / t_launch / 1e9 / fp32_peak},
                        # The resource that actually binds: warp-instruction issue slots (148 SMs x 4 schedulers x 1 per
                        # clock).  warp instructions per ray and threads per instruction are constants of the committed
                        # ncu capture (profiles/r1i_*: smsp__inst_executed.sum / rays, thread_inst / inst), K2 only.
                        "issue": ({"warp_inst_per_ray": 82.9, "threads_per_inst": 11.37,
                                   "achieved_ginst_s": 82.9 * rays_per_launch / t_launch / 1e9,
                                   "peak_ginst_s": 148 * 4 * ((clocks or {}).get("sm_mhz") or 1965.0) / 1e3,
                                   "frac": 82.9 * rays_per_launch / t_launch / 1e9 / (148 * 4 * ((clocks or {}).get("sm_mhz") or 1965.0) / 1e3),
                                   "source": "profiles/r1i_ncu_full_render_kernel_sm_K2_raw.csv"}
                                  if (args.workload == "K2" and not args.spp) else None),
                        "note": "node/triangle fetches are served by L1/L2 (scene < 1 MB); HBM traffic is only the accumulation buffer. "
                                "The binding limit is FP32/issue rate x SIMT efficiency, see profiles/."}
    print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
