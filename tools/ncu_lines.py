"""Per-source-line instruction counts and stall samples from an .ncu-rep (needs -lineinfo and --import-source on).
usage: python tools/ncu_lines.py report.ncu-rep kernel-regex [top-n]"""
import csv, io, subprocess, sys
rep, kre = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--kernel-name", "regex:" + kre],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
fname, hdr, out = None, None, []
for r in rows:
    if len(r) == 2 and r[0] == "File Path": fname = r[1].split("/")[-1]; continue
    if len(r) > 5 and r[0] == "Line No": hdr = r; continue
    if hdr and len(r) == len(hdr) and r[0] not in ("", "Line No"):
        d = dict(zip(hdr, r))
        # first "Source" column is the cuda line; dict keeps the last duplicate, so take by index
        out.append((fname, int(r[0]), r[1].strip(), int(r[hdr.index("Instructions Executed")] or 0), int(r[hdr.index("# Samples")] or 0)))
tot_i = sum(o[3] for o in out) or 1
tot_s = sum(o[4] for o in out) or 1
print(f"total warp instructions {tot_i}, samples {tot_s}")
for f, ln, src, ins, smp in sorted(out, key=lambda o: -o[3])[:top]:
    print(f"{100*ins/tot_i:5.1f}% inst {100*smp/tot_s:5.1f}% smp  {f}:{ln}  {src[:100]}")
