#!/bin/bash
# N-GPU bench through torchrun (as the driver launches it) with both halo exchanges + a sharded-vs-unsharded byte check.
N=${1:-2}
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29612 tools/check_sharded.py 2>&1 | grep -v "^W\|^\*" | tail -30
for halo in peer p2p; do
  echo "== bench N=$N halo $halo"
  CHAQSDF_HALO=$halo timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus $N --steps 20 --warmup 5 2> gpurun_out/bench_n${N}_$halo.err | tee gpurun_out/bench_n${N}_$halo.json | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print(round(d['value']), round(d['ms_per_step'], 4), d.get('phases'), round(d['e2e']['ms_per_step'], 3))"
  tail -3 gpurun_out/bench_n${N}_$halo.err
done
