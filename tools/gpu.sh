#!/bin/bash
# One entry point for the GPU box (replaces the per-experiment wrappers of round 1).
#   tools/gpu.sh tests                      GPU parity tests
#   tools/gpu.sh quick [workloads...]       one short bench line per workload (default K2 K1 K3a K3b K4)
#   tools/gpu.sh ab NAME [workloads...]     same, shipped library against lib/libmcpt_NAME.so (MCPT_BUILD_NAME builds)
#   tools/gpu.sh probe SCRIPT [args]        python tools/SCRIPT args, output kept in gpurun_out/
#   tools/gpu.sh ncu [workload] [tag]       launch list + one --set full capture of render_kernel_sm, exported as csv
mkdir -p gpurun_out
line() {  # $1 = json file
python - "$1" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(d["config"]["workload"][:28], "| value %.1f e2e %.1f Mrays/s %.0f ms/step %.2f" % (d["value"], d["e2e"]["value"], d.get("mrays_per_s", 0), d["ms_per_step"]))
except Exception as e:
    print("no line in", sys.argv[1], e)
PY
}
cmd=$1; shift
case $cmd in
tests)
  timeout 1500 python -m pytest tests -m gpu -x -q "$@" > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -15 gpurun_out/pytest_gpu.log;;
quick)
  for w in ${@:-K2 K1 K3a K3b K4}; do
    timeout 600 python bench.py --workload $w --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/quick_$w.json 2> gpurun_out/quick_$w.err; echo "bench $w exit $?"; line gpurun_out/quick_$w.json
  done;;
ab)
  name=$1; shift
  for w in ${@:-K2 K1 K3a K3b K4}; do
    for lib in "" $name; do
      MCPT_LIB=${lib:+$PWD/montecarlopathtracer_b200/lib/libmcpt_$lib.so} timeout 600 python bench.py --workload $w --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ab_${w}_${lib:-ship}.json 2> gpurun_out/ab_${w}_${lib:-ship}.err
      echo -n "${lib:-ship} "; line gpurun_out/ab_${w}_${lib:-ship}.json
    done
  done;;
probe)
  s=$1; shift
  timeout 1200 python tools/$s "$@" > gpurun_out/${s%.py}.log 2>&1; echo "$s exit $?"; tail -40 gpurun_out/${s%.py}.log;;
ncu)
  w=${1:-K2}; tag=${2:-r2}
  args="--workload $w --steps 1 --warmup 1 --no-cpu-baseline --no-extras"
  python bench.py $args > gpurun_out/${tag}_bench_${w}_short.json 2> gpurun_out/plain.err &&
  ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_launches_bench_$w.csv \
      python bench.py $args > gpurun_out/ncu_list.log 2>&1
  ncu --set full --clock-control none --import-source on -k regex:render_kernel_sm -s 1 -c 1 -f -o gpurun_out/${tag}_ncu_full_$w \
      python bench.py $args > gpurun_out/ncu_full.log 2>&1
  ncu -i gpurun_out/${tag}_ncu_full_$w.ncu-rep --page raw --csv > gpurun_out/${tag}_ncu_full_${w}_raw.csv 2>/dev/null
  ncu -i gpurun_out/${tag}_ncu_full_$w.ncu-rep --page details > gpurun_out/${tag}_ncu_full_${w}_details.txt 2>/dev/null
  ncu -i gpurun_out/${tag}_ncu_full_$w.ncu-rep --page source --print-source cuda,sass --csv > gpurun_out/${tag}_ncu_full_${w}_source.csv 2>/dev/null
  python tools/ncu_lines.py gpurun_out/${tag}_ncu_full_${w}_source.csv 70 > gpurun_out/${tag}_ncu_source_lines_$w.txt 2>&1
  rm -f gpurun_out/${tag}_ncu_full_${w}_source.csv
  ls -la gpurun_out | tail -8;;
metrics)
  # tools/gpu.sh metrics TAG KERNEL_REGEX python-script args...: a handful of counters per launch (SIMT efficiency, issue slots, instructions)
  tag=$1; k=$2; shift 2
  python tools/"$@" > gpurun_out/${tag}_plain.log 2>&1 &&
  ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,smsp__thread_inst_executed_per_inst_executed.ratio,smsp__issue_active.avg.pct_of_peak_sustained_active \
      --clock-control none -k regex:$k --csv --log-file gpurun_out/${tag}_metrics.csv python tools/"$@" > gpurun_out/${tag}_ncu.log 2>&1
  python - gpurun_out/${tag}_metrics.csv <<'PY'
import csv, sys
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10 and r[0].isdigit()]
by = {}
for r in rows:
    by.setdefault(r[0], {})[r[-3]] = r[-1]
for k, v in by.items():
    print(k, {a.split("__")[-1][:40]: b for a, b in v.items()})
PY
  ;;
*) echo "unknown command $cmd"; exit 2;;
esac
