#!/bin/bash
# one paint-kernel iteration on the GPU box: times + CRCs of the two reference workloads, then an ncu capture of paint
export CHAQSDF_HULL_PARTS=1
python tools/ab_pass2.py 16384 256 5
python tools/ab_pass2.py 8192 256 3 noise
python tools/ab_pass2.py 4096 64 3
ncu --set full --clock-control none --import-source on -k regex:"hull_paint" -s 2 -c 1 -f -o gpurun_out/paint_it python tools/ab_pass2.py 16384 256 2 > gpurun_out/ncu_full.log 2>&1
