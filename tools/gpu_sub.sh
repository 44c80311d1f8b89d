#!/bin/bash
# A/B of the build kernel's sub-run shape: band timings and the full step, new library against gpurun_ab/base.so
echo "== new, auto"; timeout 300 python tools/time_bands.py 2>&1 | tail -4
echo "== new, sub=1"; CHAQSDF_HULL_SUB=1 timeout 300 python tools/time_bands.py 2>&1 | tail -4
echo "== new, sub=4"; CHAQSDF_HULL_SUB=4 timeout 300 python tools/time_bands.py 2>&1 | tail -4
echo "== base"; CHAQSDF_LIB=$PWD/gpurun_ab/base.so timeout 300 python tools/time_bands.py 2>&1 | tail -4
echo "== step new"; timeout 300 python tools/profile_step.py 16384 256 6 2>&1 | tail -2
echo "== step sub=4"; CHAQSDF_HULL_SUB=4 timeout 300 python tools/profile_step.py 16384 256 6 2>&1 | tail -2
echo "== step base"; CHAQSDF_LIB=$PWD/gpurun_ab/base.so timeout 300 python tools/profile_step.py 16384 256 6 2>&1 | tail -2
CHAQSDF_HULL_SUB=4 timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -2
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -2
python bench.py --steps 10 --warmup 3 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['e2e'])"
