#!/bin/bash
mkdir -p gpurun_out
echo "== default"; timeout 300 python tools/profile_step.py 16384 256 4 2>&1 | tail -2
timeout 300 python tools/time_bands.py | grep -E "world (1|8)"
for l in gpurun_ab/*.so; do [ -f "$l" ] || continue; echo "== $l"; CHAQSDF_LIB=$PWD/$l timeout 300 python tools/profile_step.py 16384 256 4 2>&1 | tail -2;  CHAQSDF_LIB=$PWD/$l timeout 300 python tools/time_bands.py | grep -E "world (1|8)"; done
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python tools/time_configs.py c2 c3
