"""Device-timed runs of the other BASELINE configs on one GPU (parity cases, not bench lines):
   python tools/time_configs.py [c2] [c3] [c4] [c5]   c2: 4096^2 s64   c3: 4096 glyphs 256^2 s16   c4: 16384^2 s256 (one
   call for both passes, unlike bench.py's two)   c5: 32768^2 s1024"""
import json
import os
import sys
import zlib

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402
import bench  # noqa: E402

pkg = ge.load_package()
which = sys.argv[1:] or ["c2", "c3"]


def timed(fn, iters=5):
    fn()
    fn()
    torch.cuda.synchronize()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(iters)]
    for a, b in ev:
        a.record()
        fn()
        b.record()
    torch.cuda.synchronize()
    return min(a.elapsed_time(b) for a, b in ev)


for c in which:
    if c in ("c2", "c4"):
        n, spread = (4096, 64) if c == "c2" else (16384, 256)
        d = torch.from_numpy(bench.syn_discs(n)).cuda()
        out = torch.empty((n, n), dtype=torch.uint8, device="cuda")
        ws = torch.empty(pkg.workspace_bytes(n, n, spread), dtype=torch.uint8, device="cuda")
        ms = timed(lambda: pkg.sdf_u8_device(d, n, n, spread=spread, out=out, workspace=ws))
        px = n * n
    elif c == "c3":
        n, s, spread = 4096, 256, 16
        d = torch.from_numpy(bench.syn_glyphs(n, s)).cuda()
        out = torch.empty((n, s, s), dtype=torch.uint8, device="cuda")
        ws = torch.empty(pkg.workspace_bytes(s, s, spread, n), dtype=torch.uint8, device="cuda")
        ms = timed(lambda: pkg.sdf_u8_batched_device(d, n, s, s, spread=spread, out=out, workspace=ws))
        px = n * s * s
    else:
        n, spread = 32768, 1024
        d = torch.from_numpy(bench.syn_discs(n)).cuda()
        out = torch.empty((n, n), dtype=torch.uint8, device="cuda")
        ws = torch.empty(pkg.workspace_bytes(n, n, spread), dtype=torch.uint8, device="cuda")
        ms = timed(lambda: pkg.sdf_u8_device(d, n, n, spread=spread, out=out, workspace=ws), iters=3)
        px = n * n
    print(json.dumps({"config": c, "ms": round(ms, 4), "Gpix/s": round(px / ms / 1e6, 1), "GB/s_at_2B_per_px": round(2 * px / ms / 1e6, 1),
                      "crc": f"{zlib.crc32(out.cpu().numpy().tobytes()):08x}", "env_glyph": os.environ.get("CHAQSDF_GLYPH", "")}))
