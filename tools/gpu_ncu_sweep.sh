#!/bin/bash
# ncu: full-set capture of the two kernels of one step (after the same command ran clean without ncu).
mkdir -p gpurun_out
SIDE=${1:-16384}; SPREAD=${2:-256}
python tools/profile_step.py $SIDE $SPREAD 2 > gpurun_out/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:'sweep|rowscan8' -s 2 -c 2 -f -o gpurun_out/prof2 \
    python tools/profile_step.py $SIDE $SPREAD 2 > gpurun_out/ncu_full.log 2>&1
cat gpurun_out/plain.log; tail -5 gpurun_out/ncu_full.log
