#!/bin/bash
# HERE (not on the GPU box): gpurun with retries while the pod answers busy / transient (exit 3, or "transient").
#   tools/gpurun_retry.sh LOGFILE TIMEOUT 'command'
log=$1; to=$2; shift 2
for i in $(seq 1 30); do
  /usr/local/graft/bin/gpurun --timeout $to -- "$@" > $log 2>&1
  if grep -q "status=transient\|status=busy\|no box or slot" $log; then sleep 150; continue; fi
  break
done
