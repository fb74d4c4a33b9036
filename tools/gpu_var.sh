#!/bin/bash
# Variant comparison on the GPU box: needs the library built with MCPT_TUNE=1 (done on the CPU side before gpurun).
mkdir -p gpurun_out
timeout 600 python tools/var_probe.py "$@" > gpurun_out/var_probe.log 2>&1; echo "exit $?"; cat gpurun_out/var_probe.log
