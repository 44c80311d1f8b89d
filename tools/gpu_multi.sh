#!/bin/bash
# N-GPU bench through torchrun (as the driver launches it) + a sharded-vs-unsharded byte check.
N=${1:-2}
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus $N --steps 10 --warmup 3 2> gpurun_out/bench_n$N.err | tee gpurun_out/bench_n$N.json
tail -5 gpurun_out/bench_n$N.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29612 tools/check_sharded.py 2>&1 | tail -5
