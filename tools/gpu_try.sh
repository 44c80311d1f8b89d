#!/bin/bash
# Quick pass-2 experiments: resident CTAs per SM, A/B builds under gpurun_ab/, then the parity suite.
for c in 8 7 6; do echo "== ctas/SM $c"; CHAQSDF_SWEEP_CTAS=$c timeout 300 python tools/profile_step.py 16384 256 4 2>&1 | tail -2; done
for l in gpurun_ab/*.so; do [ -f "$l" ] || continue; echo "== $l"; CHAQSDF_LIB=$PWD/$l timeout 300 python tools/profile_step.py 16384 256 4 2>&1 | tail -2; done
echo "== 4096/64"; timeout 300 python tools/profile_step.py 4096 64 5 2>&1 | tail -2
echo "== noise 8192/256"; timeout 300 python tools/profile_step.py 8192 256 4 noise 2>&1 | tail -2
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
