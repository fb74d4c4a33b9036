#!/bin/bash
# Instruction counts of kernel variants (MCPT_TUNE build): one ncu metrics pass per variant.
for v in "$@"; do
  MCPT_SM_VARIANT=$v python tools/render_once.py > gpurun_out/once_$v.log 2>&1 || { echo "variant $v failed"; continue; }
  MCPT_SM_VARIANT=$v ncu --clock-control none -k regex:render_kernel_sm --csv --metrics gpu__time_duration.sum,smsp__inst_executed.sum,smsp__thread_inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__warps_active.avg.pct_of_peak_sustained_active,smsp__inst_executed_op_local_ld.sum,smsp__inst_executed_op_local_st.sum \
     python tools/render_once.py 2>&1 | grep -v "^==" | tail -9 | cut -d, -f13- > gpurun_out/ncu_var_$v.csv
  echo "variant $v: $(cat gpurun_out/once_$v.log | tail -1)"; cat gpurun_out/ncu_var_$v.csv
done
