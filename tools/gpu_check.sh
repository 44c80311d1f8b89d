#!/bin/bash
# Parity tests + quick timing of the two passes (no ncu).
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
timeout 300 python tools/profile_step.py 16384 256 5 2>&1 | tail -3
timeout 300 python tools/profile_step.py 4096 64 5 2>&1 | tail -3
