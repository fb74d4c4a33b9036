#!/bin/bash
# Quick confirmation of a kernel change: GPU parity tests, then kernel-only Mpaths/s of the shipped variant.
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -2
echo "K2 full: $(python tools/render_once.py scene01_lambert.mcpt 1024)"
for s in scene01.mcpt scene02.mcpt scene03.mcpt; do echo "$s 256 spp: $(python tools/render_once.py $s 256)"; done
