#!/bin/bash
# Compare the two ring modes of pass 2 (timing only) and run the parity suite in the forced mode.
for m in S G; do
  echo "== mode $m"
  CHAQSDF_SWEEP_MODE=$m timeout 300 python tools/profile_step.py 16384 256 4 2>&1 | tail -2
  CHAQSDF_SWEEP_MODE=$m timeout 300 python tools/profile_step.py 4096 64 4 2>&1 | tail -2
  CHAQSDF_SWEEP_MODE=$m timeout 300 python tools/profile_step.py 8192 256 4 noise 2>&1 | tail -2
done
CHAQSDF_SWEEP_MODE=G timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
