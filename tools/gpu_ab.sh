#!/bin/bash
# A/B of the working library against gpurun_ab/base.so: per-kernel pass-2 times (full image and an 8-way band), step, parity
for lib in "" $PWD/gpurun_ab/base.so; do
  echo "== ${lib:-new}"
  CHAQSDF_LIB=$lib timeout 300 python tools/time_bands.py 16384 256 1 8 2>&1 | tail -2
  CHAQSDF_LIB=$lib timeout 300 python tools/profile_step.py 16384 256 6 2>&1 | tail -2
  CHAQSDF_LIB=$lib timeout 300 python tools/time_configs.py c2 c3 2>&1 | tail -2
done
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -2
