#!/bin/bash
# Quick pass-2 experiments: resident CTAs per SM (L2 footprint of the re-read window), ring modes, A/B builds.
for c in 7 6 5 4 3; do echo "== ctas/SM $c"; CHAQSDF_SWEEP_CTAS=$c timeout 300 python tools/profile_step.py 16384 256 4 2>&1 | tail -2; done
echo "== S mode"; CHAQSDF_SWEEP_MODE=S timeout 300 python tools/profile_step.py 16384 256 4 2>&1 | tail -2
for l in gpurun_ab/*.so; do echo "== $l"; CHAQSDF_LIB=$PWD/$l timeout 300 python tools/profile_step.py 16384 256 4 2>&1 | tail -2; done
