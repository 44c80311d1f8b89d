#!/bin/bash
mkdir -p gpurun_out
echo "== hull"; timeout 300 python tools/profile_step.py 16384 256 4 2>&1 | tail -2
echo "== 4096/64"; timeout 300 python tools/profile_step.py 4096 64 5 2>&1 | tail -2
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
python tools/time_configs.py c2 c3
python tools/time_bands.py
python tools/profile_step.py 16384 256 2 > gpurun_out/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:'hull_build|hull_emit' -s 2 -c 2 -f -o gpurun_out/prof_hull \
    python tools/profile_step.py 16384 256 2 > gpurun_out/ncu_full.log 2>&1
tail -2 gpurun_out/ncu_full.log
