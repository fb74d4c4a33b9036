#!/bin/bash
# HISTORICAL (fourth sweep): at the time VAR bit 5 = divides deferred to after the triangle loop, bit 6 = stepped point
# location.  Both lost and were removed (profiles/r1_tuning.md); the bits now mean div3_rn and the triangle cap, so this
# script no longer reproduces that sweep.  Kept for the record of what was run.
for v in 55 87 119; do MCPT_SM_VARIANT=$v timeout 300 python -m pytest tests/test_gpu_parity.py -x -q -k "variants or same_paths or tessellated or ragged or reference_stream or sample_ranges or intersect" 2>&1 | tail -2; done
python tools/tune_gpu.py scene01_lambert.mcpt 23,20,4 55,20,4 87,20,4 119,20,4
for n in 1 3 4 8; do echo "loc_steps $n"; MCPT_SM_LOCSTEPS=$n python tools/tune_gpu.py scene01_lambert.mcpt 87,20,4 119,20,4; done
python tools/tune_gpu.py scene01_lambert.mcpt 119,16,4 119,24,4 119,12,4
for s in scene01.mcpt scene02.mcpt scene03.mcpt heightfield:300; do echo $s; python tools/tune_gpu.py $s 23,20,4 55,20,4 119,20,4; done
