#!/bin/bash
# Full ncu capture (with source correlation) of render_kernel_sm on a short K2-scene render; the report comes back in gpurun_out/.
mkdir -p gpurun_out
SPP=${1:-64}
python tools/render_once.py scene01_lambert.mcpt $SPP > gpurun_out/once_plain.log 2>&1 && cat gpurun_out/once_plain.log &&
ncu --set full --clock-control none --import-source on -k regex:render_kernel_sm -c 1 -f -o gpurun_out/prof_src python tools/render_once.py scene01_lambert.mcpt $SPP > gpurun_out/ncu_src.log 2>&1
ls -la gpurun_out/prof_src.ncu-rep
