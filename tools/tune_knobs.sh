#!/bin/bash
# Runtime knobs of the shipped kernel at full K2 size: shade threshold, camera-queue refill threshold, slices.
for sh in 14 16 18 20 22 24; do echo "shade $sh: $(MCPT_SM_SHADE=$sh python tools/render_once.py scene01_lambert.mcpt 1024)"; done
for r in 8 12 16 20 24; do echo "camr $r: $(MCPT_SM_CAMR=$r python tools/render_once.py scene01_lambert.mcpt 1024)"; done
for c in 2 4 6 8; do echo "chunks $c: $(MCPT_CHUNKS=$c python tools/render_once.py scene01_lambert.mcpt 1024)"; done
