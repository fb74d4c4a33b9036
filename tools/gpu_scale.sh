#!/bin/bash
# 1/2/4/8-GPU scaling of the headline workload (run with gpurun --gpus 8)
mkdir -p gpurun_out
for n in 1 2 4 8; do
  if [ $n -eq 1 ]; then python bench.py --gpus 1 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/scale_K2_n$n.json 2> gpurun_out/scale_K2_n$n.err
  else python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29520+n)) bench.py --gpus $n --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/scale_K2_n$n.json 2> gpurun_out/scale_K2_n$n.err; fi
  echo "n=$n exit $?"; python -c "
import json,sys
d=json.loads([l for l in open('gpurun_out/scale_K2_n$n.json') if l.startswith('{')][-1])
print('n', d['n_gpus'], 'value', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'ms/step', round(d['ms_per_step'],1), d['clocks'])"
done
# K5 at reduced spp, strong split over 8 GPUs
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29540 bench.py --gpus 8 --steps 1 --warmup 1 --workload K5 --spp 512 --scaling strong --no-cpu-baseline > gpurun_out/scale_K5_512spp_n8.json 2> gpurun_out/scale_K5_512spp_n8.err; echo "K5 exit $?"; grep '^{' gpurun_out/scale_K5_512spp_n8.json | cut -c1-300
