#!/bin/bash
# ncu launch list (per-kernel durations, cold-cache and serialised) of one workload
SIDE=${1:-16384}; SPREAD=${2:-256}
mkdir -p gpurun_out
python tools/profile_step.py $SIDE $SPREAD 2 > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_$SIDE.csv \
    python tools/profile_step.py $SIDE $SPREAD 2 > gpurun_out/ncu_launches.log 2>&1
cat gpurun_out/plain.log
python - <<PY
import csv
rows=[r for r in csv.reader(open("gpurun_out/launches_$SIDE.csv")) if len(r)>10 and r[0].isdigit()]
for r in rows: print(r[4][:60].ljust(60), r[8], r[7], int(r[-1])/1e3, "us")
PY
