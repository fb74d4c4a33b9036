#!/bin/bash
# One gpurun call: GPU parity tests, smoke, bench, ncu launch list + one full capture of the render kernel.
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"; tail -2 gpurun_out/smoke.log
python bench.py > gpurun_out/bench_K2.json 2> gpurun_out/bench_K2.err; echo "bench exit $?"; cat gpurun_out/bench_K2.json
python bench.py --impl reference --steps 1 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; cat gpurun_out/bench_ref.json
python tools/stats_gpu.py > gpurun_out/stats.log 2>&1; cat gpurun_out/stats.log
python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_list.log 2>&1
python bench.py --steps 1 --warmup 1 --spp 64 --no-cpu-baseline > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:render_kernel -s 1 -c 1 -f -o gpurun_out/prof_render \
    python bench.py --steps 1 --warmup 1 --spp 64 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1
ls -la gpurun_out
