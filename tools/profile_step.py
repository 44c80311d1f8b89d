"""Runs the two device passes a few times on one workload, timing each with CUDA events (also the command
profiled under ncu: no CPU baseline, no e2e).
usage: python tools/profile_step.py [side] [spread] [iters] [kind]      kind: discs (default) | noise"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402
import bench  # noqa: E402

side = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
spread = int(sys.argv[2]) if len(sys.argv) > 2 else 256
iters = int(sys.argv[3]) if len(sys.argv) > 3 else 3
kind = sys.argv[4] if len(sys.argv) > 4 else "discs"
pkg = ge.load_package()
if kind == "noise":
    g = torch.Generator().manual_seed(1)
    d = ((torch.rand((side, side), generator=g) < 1 / 64).to(torch.uint8) * 255).cuda()
else:
    d = torch.from_numpy(bench.syn_discs(side)).cuda()
out = torch.empty((side, side), dtype=torch.uint8, device="cuda")
rb = pkg.inter_row_bytes(side, spread)
inter = torch.empty((side, rb), dtype=torch.uint8, device="cuda")
nb = int(pkg.lib().chaqsdf_envelope_workspace_bytes(side, side, spread))
ws = torch.empty(max(nb, 256), dtype=torch.uint8, device="cuda")
ev = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(iters)]
for i in range(iters):
    ev[i][0].record()
    pkg.rowscan_device(d, side, side, spread, inter)
    ev[i][1].record()
    pkg.envelope_device(inter, side, side, 0, side, spread, False, out, workspace=ws)
    ev[i][2].record()
torch.cuda.synchronize()
p1 = [round(e[0].elapsed_time(e[1]), 3) for e in ev]
p2 = [round(e[1].elapsed_time(e[2]), 3) for e in ev]
px = side * side
print(f"{kind} {side}^2 spread {spread}: pass1 ms {p1} pass2 ms {p2} inter row bytes {rb}")
print(f"  pass1 {(px + side * rb) / min(p1) / 1e6:.0f} GB/s  pass2 {(px + side * rb) / min(p2) / 1e6:.0f} GB/s  "
      f"total {px / (min(p1) + min(p2)) / 1e3:.0f} Mpix/s  checksum {int(out.sum(dtype=torch.int64))}")
