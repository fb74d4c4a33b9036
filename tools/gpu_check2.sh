#!/bin/bash
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -5 gpurun_out/pytest_gpu.log
python bench.py > gpurun_out/bench_K2.json 2> gpurun_out/bench_K2.err; echo "bench exit $?"; cat gpurun_out/bench_K2.json; tail -3 gpurun_out/bench_K2.err
for w in K3a K3b K4; do python bench.py --workload $w --steps 2 --warmup 1 > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err; echo "bench $w exit $?"; cat gpurun_out/bench_$w.json; tail -3 gpurun_out/bench_$w.err; done
python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:render_kernel_sm -s 1 -c 1 -f -o gpurun_out/prof_render_K2 \
    python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1
ls -la gpurun_out
