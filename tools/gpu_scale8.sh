#!/bin/bash
# 8-GPU lines with the current kernel (run with gpurun --gpus 8): K2 weak (1024 spp per GPU) and the full K5
# configuration (4096x4096, 4096 spp split 8 x 512).
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus 8 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/scale2_K2_n8.json 2> gpurun_out/scale2_K2_n8.err; echo "K2 n8 exit $?"; grep '^{' gpurun_out/scale2_K2_n8.json | cut -c1-330
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 8 --steps 1 --warmup 1 --workload K5 --scaling strong --no-cpu-baseline > gpurun_out/scale2_K5_n8.json 2> gpurun_out/scale2_K5_n8.err; echo "K5 n8 exit $?"; grep '^{' gpurun_out/scale2_K5_n8.json | cut -c1-330
