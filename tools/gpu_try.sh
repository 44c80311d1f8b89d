#!/bin/bash
mkdir -p gpurun_out
echo "== hull"; timeout 300 python tools/profile_step.py 16384 256 4 2>&1 | tail -2
for l in gpurun_ab/*.so; do [ -f "$l" ] || continue; echo "== $l"; CHAQSDF_LIB=$PWD/$l timeout 300 python tools/profile_step.py 16384 256 4 2>&1 | tail -2; done
echo "== 4096/64"; timeout 300 python tools/profile_step.py 4096 64 5 2>&1 | tail -2
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
python tools/time_configs.py c2 c3
bash tools/gpu_launches.sh 16384 256 | grep "us$" | head -6
