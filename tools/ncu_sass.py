"""SASS of one kernel from an .ncu-rep with the executed-instruction count and stall samples per instruction.
usage: python tools/ncu_sass.py report.ncu-rep kernel-regex > file"""
import csv, io, subprocess, sys
rep, kre = sys.argv[1], sys.argv[2]
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass", "--kernel-name", "regex:" + kre],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
hdr = rows[1]
ia, isrc, ismp, iex, ithr = (hdr.index(k) for k in ("Address", "Source", "# Samples", "Instructions Executed", "Avg. Threads Executed"))
base = None
tot = 0
for r in rows[2:]:
    if len(r) != len(hdr): continue
    a = int(r[ia], 16)
    base = a if base is None else base
    ex = int(r[iex] or 0); tot += ex
    print(f"{a - base:05x} {ex:>10} {r[ithr]:>5} {int(r[ismp] or 0):>6}  {r[isrc].strip()}")
print("total", tot)
