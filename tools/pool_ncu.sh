#!/bin/bash
# GPU box: one --set full capture of render_kernel_pool (K2 scene, 1024^2 x 64 spp) with per-line view
export MCPT_POOL=${1:-1}
ncu --set full --clock-control none --import-source on -k regex:render_kernel -c 1 -f -o gpurun_out/pool_full python tools/render_once.py scene01_lambert.mcpt 256 > gpurun_out/pool_full.log 2>&1
ncu -i gpurun_out/pool_full.ncu-rep --page source --print-source cuda,sass --csv > gpurun_out/pool_full_source.csv 2>/dev/null
python tools/ncu_lines.py gpurun_out/pool_full_source.csv 400 > gpurun_out/pool_source_lines.txt 2>&1
rm -f gpurun_out/pool_full_source.csv
head -12 gpurun_out/pool_source_lines.txt
