M=gpu__time_duration.sum,smsp__inst_executed.sum,smsp__thread_inst_executed_per_inst_executed.ratio,smsp__issue_active.avg.pct_of_peak_sustained_active,l1tex__t_sector_hit_rate.pct,l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum,smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio,smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio,smsp__average_warps_issue_stalled_wait_per_issue_active.ratio,smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio,smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio,smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio,smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio,smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio,smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio
for pool in 0 1; do
  MCPT_POOL=$pool ncu --metrics $M --clock-control none -k regex:render_kernel --csv --log-file gpurun_out/pool_metrics_$pool.csv python tools/render_once.py scene01_lambert.mcpt 256 > gpurun_out/pool_ncu_$pool.log 2>&1
  python - gpurun_out/pool_metrics_$pool.csv <<'PY'
import csv, sys
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10 and r[0].isdigit()]
for r in rows: print(r[4][:40], r[-3].replace("smsp__average_warps_issue_stalled_","stall_").replace("_per_issue_active.ratio","")[:60], r[-1])
PY
done
