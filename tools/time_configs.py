"""Device-timed throughput of the other BASELINE.json configs (C2, C3, C5-on-one-GPU crop) -- not bench lines.
usage: python tools/time_configs.py [c2] [c3] [c5]"""
import os, sys, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge
import bench
pkg = ge.load_package()
which = sys.argv[1:] or ["c2", "c3", "c5"]

def timed(fn, iters=5):
    fn(); torch.cuda.synchronize()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(iters)]
    for a, b in ev:
        a.record(); fn(); b.record()
    torch.cuda.synchronize()
    return min(a.elapsed_time(b) for a, b in ev)

if "c2" in which:
    n, s = 4096, 64
    d = torch.from_numpy(bench.syn_discs(n)).cuda()
    ws = torch.empty(pkg.workspace_bytes(n, n, s), dtype=torch.uint8, device="cuda")
    out = torch.empty((n, n), dtype=torch.uint8, device="cuda")
    ms = timed(lambda: pkg.sdf_u8_device(d, n, n, spread=s, out=out, workspace=ws))
    print(f"C2 4096^2 s64: {ms:.3f} ms  {n*n/ms/1e3:.0f} Mpix/s")
if "c3" in which:
    g, n, s = 4096, 256, 16
    rng = np.random.default_rng(3)
    yy, xx = np.mgrid[0:n, 0:n]
    base = []
    for i in range(64):
        m = np.zeros((n, n), bool)
        for _ in range(int(rng.integers(2, 6))):
            cy, cx, r = rng.integers(40, n - 40), rng.integers(40, n - 40), rng.integers(10, 60)
            ring = np.abs(np.hypot(yy - cy, xx - cx) - r) < rng.integers(3, 10)
            m |= ring
        base.append(m)
    imgs = (np.stack([base[i % 64] for i in range(g)]) * 255).astype(np.uint8)
    d = torch.from_numpy(imgs).cuda()
    ws = torch.empty(pkg.workspace_bytes(n, n, s, g), dtype=torch.uint8, device="cuda")
    out = torch.empty((g, n, n), dtype=torch.uint8, device="cuda")
    ms = timed(lambda: pkg.sdf_u8_batched_device(d, g, n, n, spread=s, out=out, workspace=ws))
    pkg.profile_enable(True)
    pkg.sdf_u8_batched_device(d, g, n, n, spread=s, out=out, workspace=ws)
    km = pkg.profile_mean_ms()
    pkg.profile_enable(False)
    k2 = sum(v for k, v in km.items() if k != "calls")
    print(f"C3 4096 x 256^2 s16: {ms:.3f} ms  {g*n*n/ms/1e3:.0f} Mpix/s   pass 2 kernels " +
          " ".join(f"{k.split('_')[1]}={v:.3f}" for k, v in km.items() if k != "calls") + f"  -> pass 1 ~{ms - k2:.3f} ms")
if "c5" in which:
    for n, s in ((16384, 1024), (32768, 1024)):
        try:
            img = bench.syn_discs(n)
            d = torch.from_numpy(img).cuda()
            ws = torch.empty(pkg.workspace_bytes(n, n, s), dtype=torch.uint8, device="cuda")
            out = torch.empty((n, n), dtype=torch.uint8, device="cuda")
            ms = timed(lambda: pkg.sdf_u8_device(d, n, n, spread=s, out=out, workspace=ws), 3)
            print(f"C5 {n}^2 s{s}: {ms:.3f} ms  {n*n/ms/1e3:.0f} Mpix/s  checksum {int(out.sum(dtype=torch.int64))}")
            del d, ws, out
        except Exception as ex:
            print(f"C5 {n}^2 s{s}: failed: {ex}")
