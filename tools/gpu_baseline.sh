#!/bin/bash
# One GPU-box visit: parity tests, bench lines (ours + reference arm), ncu launch list and one full capture.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "bench ref rc=$?"
python tools/profile_step.py 16384 256 3 > gpurun_out/plain.log 2>&1
# launch list of the bench command itself (it exited 0 without ncu above); numbers printed under ncu are not bench values
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 3 --warmup 3 > gpurun_out/ncu_launches.log 2>&1
python tools/profile_step.py 16384 256 2 > gpurun_out/plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:'hull|rowscan8' -s 5 -c 5 -f -o gpurun_out/prof \
    python tools/profile_step.py 16384 256 2 > gpurun_out/ncu_full.log 2>&1
tail -3 gpurun_out/pytest_gpu.log; cat gpurun_out/bench.json gpurun_out/bench_ref.json; cat gpurun_out/plain.log; tail -3 gpurun_out/bench.err
