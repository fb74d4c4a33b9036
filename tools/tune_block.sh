#!/bin/bash
# Block size of render_kernel_sm (MCPT_TUNE build): 256 x 4/SM, 128 x 8/SM, 64 x 16/SM, with slices 1..4.
for mb in 4 8 16; do MCPT_SM_MINBLOCKS=$mb timeout 300 python -m pytest tests/test_gpu_parity.py -x -q -k "variants or same_paths or ragged or sample_ranges" 2>&1 | tail -1; done
for mb in 4 8 16; do for c in 1 2 4; do echo "minblocks $mb chunks $c: $(MCPT_SM_MINBLOCKS=$mb MCPT_CHUNKS=$c python tools/render_once.py scene01_lambert.mcpt 1024)"; done; done
for mb in 4 8 16; do echo "scene01 minblocks $mb: $(MCPT_SM_MINBLOCKS=$mb python tools/render_once.py scene01.mcpt 512)";  echo "scene03 minblocks $mb: $(MCPT_SM_MINBLOCKS=$mb python tools/render_once.py scene03.mcpt 256)"; done
