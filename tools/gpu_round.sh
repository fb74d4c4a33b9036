#!/bin/bash
# After a kernel change: GPU parity tests, then one bench line per workload (short runs, no CPU baseline).
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/pytest_gpu.log
for w in K2 K1 K3a K3b K4; do
  timeout 600 python bench.py --workload $w --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/quick_$w.json 2> gpurun_out/quick_$w.err; echo "bench $w exit $?"
  python - "$w" <<'PY'
import json, sys
try:
    d = json.loads(open(f"gpurun_out/quick_{sys.argv[1]}.json").read().strip().splitlines()[-1])
    print(sys.argv[1], "value", d["value"], "e2e", d["e2e"]["value"], "mrays", d.get("mrays_per_s"), "ms", d["ms_per_step"])
except Exception as e:
    print("no line", e)
PY
done
