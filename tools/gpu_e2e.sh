#!/bin/bash
# bench e2e with different band schedules of the host-buffer pipeline
for r in "1024 1024" "2048 512" "2048 1024" "4096 1024" "1536 768" "3072 1024"; do set -- $r; echo "== pipe rows $1 first $2"; CHAQSDF_PIPE_ROWS=$1 CHAQSDF_PIPE_FIRST=$2 timeout 600 python bench.py --steps 5 --warmup 3 2>&1 | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['value'], d['e2e']['ms_per_step'])"; done
