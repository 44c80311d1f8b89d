#!/bin/bash
# Parity suite, then pass-2 timing under different work-split cost models (c0,cp,ce) and the equal-rows split.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
echo "== rows split"; CHAQSDF_SWEEP_SPLIT=rows timeout 300 python tools/profile_step.py 16384 256 4 2>&1 | tail -2
for c in "1,8,16" "1,4,8" "1,6,20" "1,12,24" "2,16,48" "1,20,40" "1,10,10" "1,30,60"; do
  echo "== cost $c"; CHAQSDF_SWEEP_COST=$c timeout 300 python tools/profile_step.py 16384 256 4 2>&1 | tail -2
done
echo "== 4096/64"; timeout 300 python tools/profile_step.py 4096 64 5 2>&1 | tail -2
echo "== noise 8192/256"; timeout 300 python tools/profile_step.py 8192 256 4 noise 2>&1 | tail -2
CHAQSDF_SWEEP_SPLIT=rows timeout 300 python tools/profile_step.py 8192 256 4 noise 2>&1 | tail -2
