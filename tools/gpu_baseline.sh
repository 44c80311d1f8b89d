#!/bin/bash
# One visit to a B200 box: parity suite, bench (ours + the reference arm), ncu launch list of the bench command, one
# --set full capture of every kernel of a step and of the fused small-image kernel. Results land in gpurun_out/
# (summaries are then copied into profiles/ here, with tools/ncu_summary.py, ncu_lines.py, launch_shares.py).
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 2>/dev/null | tail -1 > gpurun_out/bench_reference.json
timeout 600 python bench.py --steps 20 --warmup 3 2>/dev/null | tail -1 > gpurun_out/bench.json
CHAQ_BENCH_SKIP_OTHER=1 python bench.py --steps 2 --warmup 3 > gpurun_out/plain_bench.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
    env CHAQ_BENCH_SKIP_OTHER=1 python bench.py --steps 2 --warmup 3 > gpurun_out/ncu_launches.log 2>&1
# whole-image kernels for the per-kernel capture: one column part (in production two parts share the SMs)
CHAQSDF_HULL_PARTS=1 python tools/ab_pass2.py 16384 256 2 > gpurun_out/plain.log 2>&1 && \
CHAQSDF_HULL_PARTS=1 ncu --set full --clock-control none --import-source on -k regex:"hull_|rowscan8" -s 10 -c 5 -f -o gpurun_out/step \
    python tools/ab_pass2.py 16384 256 2 > gpurun_out/ncu_full.log 2>&1
python tools/time_configs.py c2 c3 c3atlas c5 > gpurun_out/configs.jsonl 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"glyph" -s 2 -c 1 -f -o gpurun_out/glyph \
    python tools/time_configs.py c3 > gpurun_out/ncu_glyph.log 2>&1
python tools/time_bands.py 16384 256 1 2 4 8 > gpurun_out/bands.txt 2>&1
cat gpurun_out/configs.jsonl; tail -4 gpurun_out/bands.txt; tail -1 gpurun_out/plain.log
python -c "
import json
d=json.load(open('gpurun_out/bench.json')); r=json.load(open('gpurun_out/bench_reference.json'))
print('ours', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], d['e2e']['ms_per_step'], 'ref', r['value'], 'ratio', d['value']/r['value'], d['e2e']['value']/r['value'])
print(d['roofline']); print(d['clocks'])"
