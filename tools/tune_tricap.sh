#!/bin/bash
# Seventh sweep: capped triangle tests per lane and round (VAR bit 6).  Needs a MCPT_TUNE build.
MCPT_SM_VARIANT=119 timeout 300 python -m pytest tests/test_gpu_parity.py -x -q -k "variants or same_paths or tessellated or ragged or reference_stream or sample_ranges" 2>&1 | tail -1
echo "VAR 55: $(MCPT_SM_VARIANT=55 python tools/render_once.py scene01_lambert.mcpt 1024)"
for c in 1 2 3 4 6; do echo "VAR 119 cap $c: $(MCPT_SM_VARIANT=119 MCPT_SM_TRICAP=$c python tools/render_once.py scene01_lambert.mcpt 1024)"; done
for s in scene01.mcpt scene02.mcpt scene03.mcpt; do echo "$s VAR 55: $(MCPT_SM_VARIANT=55 python tools/render_once.py $s 256)"; for c in 2 3; do echo "$s VAR 119 cap $c: $(MCPT_SM_VARIANT=119 MCPT_SM_TRICAP=$c python tools/render_once.py $s 256)"; done; done
for c in 2 3; do for sh in 16 24; do echo "VAR 119 cap $c shade $sh: $(MCPT_SM_VARIANT=119 MCPT_SM_TRICAP=$c MCPT_SM_SHADE=$sh python tools/render_once.py scene01_lambert.mcpt 1024)"; done; done
