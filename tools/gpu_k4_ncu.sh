#!/bin/bash
# K4 (1 M-triangle height field): what bounds render_kernel_sm there?  One metrics pass, 16 spp.
mkdir -p gpurun_out
python tools/k4_probe.py 708 16 > gpurun_out/k4_plain.log 2>&1; cat gpurun_out/k4_plain.log
M="gpu__time_duration.sum,smsp__inst_executed.sum,smsp__thread_inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,l1tex__t_sector_hit_rate.pct,lts__t_sector_hit_rate.pct,dram__bytes_read.sum,dram__bytes_write.sum,dram__throughput.avg.pct_of_peak_sustained_elapsed,lts__t_bytes.sum,l1tex__t_bytes.sum,smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio,smsp__average_warp_latency_per_inst_issued.ratio,smsp__warps_active.avg.per_cycle_active"
timeout 600 ncu --clock-control none -k regex:render_kernel_sm -c 1 --csv --metrics $M python tools/k4_probe.py 708 16 2>&1 | grep -v "^==" | grep "render_kernel" | cut -d, -f13- > gpurun_out/k4_ncu.csv
cat gpurun_out/k4_ncu.csv
