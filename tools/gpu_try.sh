#!/bin/bash
echo "== default"; timeout 300 python tools/profile_step.py 16384 256 6 2>&1 | tail -2
for l in gpurun_ab/*.so; do [ -f "$l" ] || continue; echo "== $l"; CHAQSDF_LIB=$PWD/$l timeout 300 python tools/profile_step.py 16384 256 6 2>&1 | tail -2; done
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -2
python tools/time_configs.py c2 c3
