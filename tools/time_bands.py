"""Per-band pass-2 time on one GPU: what each rank of an N-way row-band split would spend in the envelope pass.
usage: python tools/time_bands.py [side] [spread] [world ...]"""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge
import bench
pkg = ge.load_package()
side = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
spread = int(sys.argv[2]) if len(sys.argv) > 2 else 256
d = torch.from_numpy(bench.syn_discs(side)).cuda()
rb = pkg.inter_row_bytes(side, spread)
inter = torch.empty((side, rb), dtype=torch.uint8, device="cuda")
pkg.rowscan_device(d, side, side, spread, inter)
out = torch.empty((side, side), dtype=torch.uint8, device="cuda")
ws = torch.empty(max(int(pkg.lib().chaqsdf_envelope_workspace_bytes(side, side, spread)), 256), dtype=torch.uint8, device="cuda")
worlds = [int(a) for a in sys.argv[3:]] or [1, 2, 4, 8]
for world in worlds:
    ts = []
    for rank in range(world):
        r0, r1, ht, hb, _, _ = pkg.halo_plan(side, world, rank, spread)
        buf = inter[r0 - ht:r1 + hb]
        best = 1e9
        for it in range(4):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            pkg.envelope_device(buf, side, r1 + hb - (r0 - ht), ht, ht + (r1 - r0), spread, False, out[r0:r1], workspace=ws)
            b.record()
            torch.cuda.synchronize()
            best = min(best, a.elapsed_time(b))
        ts.append(round(best, 3))
    pkg.profile_enable(True)
    r0, r1, ht, hb, _, _ = pkg.halo_plan(side, world, world // 2, spread)
    pkg.envelope_device(inter[r0 - ht:r1 + hb], side, r1 + hb - (r0 - ht), ht, ht + (r1 - r0), spread, False, out[r0:r1], workspace=ws)
    km = pkg.profile_mean_ms()
    pkg.profile_enable(False)
    print(f"world {world}: envelope ms per rank {ts}  max {max(ts)}  sum {sum(ts):.3f}  kernels(rank {world // 2}) " +
          " ".join(f"{k.split('_')[1]}={v:.3f}" for k, v in km.items() if k != "calls"))
