#!/bin/bash
for k in 1 2 3 4 6; do echo "== runs/cta $k"; CHAQSDF_SWEEP_RUNS=$k timeout 300 python tools/profile_step.py 16384 256 4 2>&1 | tail -2; done
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
