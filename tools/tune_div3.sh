#!/bin/bash
# Fifth sweep: shared-reciprocal division (VAR bit 5) and spp chunks per tile (tail effect).  Needs a MCPT_TUNE build.
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -k "division or intersect" 2>&1 | tail -2
for v in 55 53; do MCPT_SM_VARIANT=$v timeout 300 python -m pytest tests/test_gpu_parity.py -x -q -k "variants or same_paths or tessellated or ragged or reference_stream or sample_ranges" 2>&1 | tail -2; done
python tools/tune_gpu.py scene01_lambert.mcpt 23,20,4 55,20,4 5,20,4 53,20,4
for s in scene01.mcpt scene02.mcpt scene03.mcpt heightfield:300; do echo $s; python tools/tune_gpu.py $s 23,20,4 55,20,4; done
for v in 23 55; do for c in 1 2 4 8 16 32; do echo "variant $v chunks $c: $(MCPT_SM_VARIANT=$v MCPT_CHUNKS=$c python tools/render_once.py scene01_lambert.mcpt 1024)"; done; done
