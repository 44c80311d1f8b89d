#!/bin/bash
# Parity suite + bench with different band schedules of the host-buffer pipeline.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
for r in "2048 512" "4096 512" "4096 1024" "3072 256" "8192 512" "4096 4096"; do set -- $r; echo "== pipe rows $1 first $2"; CHAQSDF_PIPE_ROWS=$1 CHAQSDF_PIPE_FIRST=$2 timeout 600 python bench.py --steps 5 --warmup 3 2>&1 | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['value'], d['e2e']['ms_per_step'])"; done
echo "== default"; timeout 600 python bench.py --steps 10 --warmup 3 | tee gpurun_out/bench.json
