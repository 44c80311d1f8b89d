"""Group an ncu launch list (--metrics gpu__time_duration.sum --csv) by kernel and grid size.
usage: python tools/launch_shares.py profiles/r01_v7_launches.csv"""
import collections, csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr = [r for r in rows if 'Kernel Name' in r][0]
ki, vi, gi = hdr.index('Kernel Name'), hdr.index('Metric Value'), hdr.index('Grid Size')
agg = collections.defaultdict(list)
for r in rows:
    if len(r) == len(hdr) and r[ki] != 'Kernel Name' and r[vi].replace('.', '').isdigit():
        agg[(r[ki].split('(')[0], r[gi])].append(float(r[vi]))
tot = sum(sum(v) for v in agg.values())
print("kernel, grid, launches, mean us, share of all captured time")
for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
    print(f"  {k[0]:26s} grid {k[1]:16s} n={len(v):3d} mean={sum(v) / len(v) / 1e3:8.1f} us  {sum(v) / tot * 100:5.1f}%")
