#!/bin/bash
# SAH constants under the current kernel's cost balance (full K2 size, kernel-only).
for ci in 0.5 0.7 0.85 1.0 1.2 1.5; do echo "CI $ci: $(MCPT_SAH_CI=$ci python tools/render_once.py scene01_lambert.mcpt 1024)"; done
for eb in 0.6 0.8 0.9; do echo "EB $eb: $(MCPT_SAH_EB=$eb python tools/render_once.py scene01_lambert.mcpt 1024)"; done
for ci in 0.7 1.0 1.3; do echo "scene01 CI $ci: $(MCPT_SAH_CI=$ci python tools/render_once.py scene01.mcpt 512)"; echo "scene03 CI $ci: $(MCPT_SAH_CI=$ci python tools/render_once.py scene03.mcpt 256)"; done
