#!/bin/bash
# Round-end evidence, short form (GPU budget): tests, smoke, K2 bench line with the CPU baseline, the other workloads without it,
# ncu launch list + full capture of render_kernel_sm.
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/pytest_gpu.log
python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"; tail -1 gpurun_out/smoke.log
python bench.py > gpurun_out/bench_K2.json 2> gpurun_out/bench_K2.err; echo "bench exit $?"; cut -c1-300 gpurun_out/bench_K2.json
for w in K1 K3a K3b K4; do python bench.py --workload $w --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err; echo "bench $w exit $?"; cut -c1-200 gpurun_out/bench_$w.json; done
python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_list.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:render_kernel_sm -s 1 -c 1 -f -o gpurun_out/prof_final_K2 \
    python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1
ls -la gpurun_out/prof_final_K2.ncu-rep
