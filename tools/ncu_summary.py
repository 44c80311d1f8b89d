"""Summarise an .ncu-rep here (no GPU needed): key metrics per kernel, and optionally the hottest source lines.
usage: python tools/ncu_summary.py report.ncu-rep [kernel-regex-for-source-page] [top-n]"""
import csv, io, subprocess, sys, collections
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
r = list(csv.reader(io.StringIO(raw)))
h = r[0]
want = ['Kernel Name', 'launch__grid_size', 'launch__block_size', 'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'launch__registers_per_thread', 'launch__occupancy_limit_shared_mem', 'launch__occupancy_limit_registers',
        'smsp__inst_executed.sum', 'sm__inst_executed.avg.per_cycle_elapsed', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'smsp__thread_inst_executed_per_inst_executed.ratio', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'l1tex__t_sector_hit_rate.pct', 'lts__t_sector_hit_rate.pct', 'lts__t_bytes.sum',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'smsp__average_warp_latency_issue_stalled_short_scoreboard.pct',
        'smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio']
for row in r[2:]:
    for k in want:
        if k in h:
            print(f"{k} [{r[1][h.index(k)]}] = {row[h.index(k)]}")
    for i, k in enumerate(h):
        if 'issue_stalled' in k and k.endswith('per_issue_active.ratio'):
            try:
                if float(row[i]) > 0.3: print(f"  stall {k.split('issue_stalled_')[1].split('_per')[0]} = {row[i]}")
            except ValueError: pass
    print()
if len(sys.argv) > 2:
    src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass,cuda" if False else "sass", "--kernel-name", "regex:" + sys.argv[2]],
                         capture_output=True, text=True).stdout
    print(src[:300])
