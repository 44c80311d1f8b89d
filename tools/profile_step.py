"""Runs the device path a few times on one workload, for ncu (no CPU baseline, no e2e).
usage: python tools/profile_step.py [side] [spread] [iters]"""
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
pkg = ge.load_package()
d = torch.from_numpy(bench.syn_discs(side)).cuda()
out = torch.empty((side, side), dtype=torch.uint8, device="cuda")
ws = torch.empty(pkg.workspace_bytes(side, side, spread), dtype=torch.uint8, device="cuda")
ev = [torch.cuda.Event(enable_timing=True) for _ in range(iters + 1)]
for i in range(iters):
    ev[i].record()
    pkg.sdf_u8_device(d, side, side, spread=spread, out=out, workspace=ws)
ev[iters].record()
torch.cuda.synchronize()
print("ms per step:", [round(ev[i].elapsed_time(ev[i + 1]), 3) for i in range(iters)], "checksum", int(out.sum(dtype=torch.int64)))
