#!/bin/bash
bash tools/gpu_multi.sh 8 2>&1 | grep -v "identical=True"
echo "== bench N=4 halo peer"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29613 bench.py --gpus 4 --steps 20 --warmup 5 2> gpurun_out/bench_n4_peer.err | tee gpurun_out/bench_n4_peer.json | cut -c1-400
