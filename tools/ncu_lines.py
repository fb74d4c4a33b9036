#!/usr/bin/env python
"""Aggregate `ncu --page source --print-source cuda,sass --csv` per CUDA source line.

usage: ncu -i X.ncu-rep --page source --print-source cuda,sass --csv > src.csv; python tools/ncu_lines.py src.csv [top]
Prints, per file:line, warp instructions executed, share of the kernel, threads active per instruction, stall samples.
"""
import csv, sys
from collections import defaultdict

path = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
rows = list(csv.reader(open(path)))
agg = defaultdict(lambda: [0, 0, 0, 0])  # inst, thread inst, samples, n sass
src = {}
cur_file = None
hdr = None
for r in rows:
    if len(r) == 2 and r[0] == "File Path":
        cur_file = r[1].split("/")[-1]
        continue
    if r and r[0] == "Line No":
        hdr = r
        iI = hdr.index("Instructions Executed"); iT = hdr.index("Thread Instructions Executed"); iS = hdr.index("# Samples")
        continue
    if hdr is None or len(r) < len(hdr) or not r[0].isdigit():
        continue
    key = (cur_file, int(r[0]))
    if r[2] == "":  # CUDA line row: remember text
        src[key] = r[1].strip()
        continue
    try:
        a = agg[key]
        a[0] += int(r[iI] or 0); a[1] += int(r[iT] or 0); a[2] += int(r[iS] or 0); a[3] += 1
    except ValueError:
        pass
tot = sum(a[0] for a in agg.values()) or 1
tots = sum(a[2] for a in agg.values()) or 1
print(f"total warp instructions {tot:.4g}, thread instructions {sum(a[1] for a in agg.values()):.4g}, "
      f"avg threads/inst {sum(a[1] for a in agg.values()) / tot:.2f}")
byfile = defaultdict(lambda: [0, 0, 0])
for (f, l), a in agg.items():
    byfile[f][0] += a[0]; byfile[f][1] += a[1]; byfile[f][2] += a[2]
for f, a in sorted(byfile.items(), key=lambda x: -x[1][0]):
    print(f"  {f:20s} {100 * a[0] / tot:6.2f}% of warp inst, {a[1] / max(a[0], 1):5.2f} thr/inst, {100 * a[2] / tots:6.2f}% of samples")
print(f"{'file:line':28s} {'%inst':>6s} {'thr/inst':>8s} {'%smpl':>6s} {'sass':>4s}  source")
for (f, l), a in sorted(agg.items(), key=lambda x: -x[1][0])[:top]:
    print(f"{f + ':' + str(l):28s} {100 * a[0] / tot:6.2f} {a[1] / max(a[0], 1):8.2f} {100 * a[2] / tots:6.2f} {a[3]:4d}  {src.get((f, l), '')[:90]}")
