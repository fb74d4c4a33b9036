#!/usr/bin/env python
"""HERE: turns one `ncu --set full` capture of render_kernel_sm (exported with `ncu -i X.ncu-rep --page raw --csv`) plus the
bench line of the SAME build into the record bench.py reads from profiles/ncu_constants.json -- the counters only ncu can
give, keyed by workload and by the hash of the kernel sources they were taken from (bench.py emits null for any other kernel).

    python tools/ncu_constants.py K2 profiles/r2a_ncu_full_K2_raw.csv profiles/r2a_bench_K2.json
"""
import csv, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench

workload, raw, line = sys.argv[1], sys.argv[2], sys.argv[3]
rows = list(csv.reader(open(raw)))
hdr, units, val = rows[0], rows[1], rows[2]
get = lambda name: float(val[hdr.index(name)])
scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
byt = lambda name: get(name) * scale[units[hdr.index(name)]]
b = json.loads(open(line).read().strip().splitlines()[-1])
rays = b["roofline"]["rays_per_launch"] if b.get("roofline") else b["mrays_per_s"] * 1e6 * b["ms_per_step"] * 1e-3
rec = {"kernel_source_hash": bench.kernel_source_hash(), "source": os.path.relpath(raw, ROOT),
       "warp_inst": get("smsp__inst_executed.sum"), "rays_per_launch": rays,
       "warp_inst_per_ray": get("smsp__inst_executed.sum") / rays,
       "threads_per_inst": get("smsp__thread_inst_executed_per_inst_executed.ratio"),
       "issue_slots_busy": get("smsp__issue_active.avg.pct_of_peak_sustained_active") / 100.0,
       "dram_bytes_per_launch": byt("dram__bytes_read.sum") + byt("dram__bytes_write.sum"),
       "l1_hit_rate": get("l1tex__t_sector_hit_rate.pct") / 100.0, "l2_hit_rate": get("lts__t_sector_hit_rate.pct") / 100.0,
       "registers_per_thread": get("launch__registers_per_thread"), "ncu_duration_ms": get("gpu__time_duration.sum") * {"ms": 1.0, "us": 1e-3, "s": 1e3, "ns": 1e-6}[units[hdr.index("gpu__time_duration.sum")]]}
path = os.path.join(ROOT, "profiles", "ncu_constants.json")
allrec = json.load(open(path)) if os.path.exists(path) else {}
allrec[workload] = rec
json.dump(allrec, open(path, "w"), indent=1, sort_keys=True)
print(json.dumps(rec, indent=1))
