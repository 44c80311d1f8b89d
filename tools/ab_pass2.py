"""Per-kernel times of the banded pass 2 (library CUDA events) on one workload, plus a CRC of the bytes, so two
builds or two settings of an environment switch can be compared:  python tools/ab_pass2.py [side] [spread] [iters] [kind]"""
import os
import sys
import zlib

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402
import bench  # noqa: E402

side = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
spread = int(sys.argv[2]) if len(sys.argv) > 2 else 256
iters = int(sys.argv[3]) if len(sys.argv) > 3 else 5
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
for i in range(2):
    pkg.rowscan_device(d, side, side, spread, inter)
    pkg.envelope_device(inter, side, side, 0, side, spread, False, out, workspace=ws)
torch.cuda.synchronize()
ev = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(iters)]
for i in range(iters):  # the step as it runs in production (column parts of pass 2 overlap on streams)
    ev[i][0].record()
    pkg.rowscan_device(d, side, side, spread, inter)
    ev[i][1].record()
    pkg.envelope_device(inter, side, side, 0, side, spread, False, out, workspace=ws)
    ev[i][2].record()
torch.cuda.synchronize()
pkg.profile_enable(True)  # per-kernel times: one part, kernels back to back
for i in range(iters):
    pkg.rowscan_device(d, side, side, spread, inter)
    pkg.envelope_device(inter, side, side, 0, side, spread, False, out, workspace=ws)
torch.cuda.synchronize()
ms = pkg.profile_mean_ms()
pkg.profile_enable(False)
p1 = min(e[0].elapsed_time(e[1]) for e in ev)
p2 = min(e[1].elapsed_time(e[2]) for e in ev)
crc = zlib.crc32(out.cpu().numpy().tobytes())
print(f"{kind} {side}^2 s{spread} parts[{os.environ.get('CHAQSDF_HULL_PARTS','auto')}]: "
      f"pass1 {p1:.3f} pass2 {p2:.3f} ms  kernels {ms}  crc {crc:08x}")
