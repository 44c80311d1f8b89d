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
    elif c == "c3atlas":  # config 3 through the atlas front-end: boxes, SDF batch, host packing, blit (device pieces timed)
        import time
        import numpy as np
        n, s, spread, aw = 4096, 256, 16, 8192
        d = torch.from_numpy(bench.syn_glyphs(n, s)).cuda()
        out = torch.empty((n, s, s), dtype=torch.uint8, device="cuda")
        ws = torch.empty(pkg.workspace_bytes(s, s, spread, n), dtype=torch.uint8, device="cuda")
        boxes = pkg.glyph_boxes_device(d, n, s, s, margin=spread)
        b = boxes.cpu().numpy()
        t0 = time.perf_counter()
        xy, ah = pkg.atlas_pack(np.stack([b[:, 2] - b[:, 0], b[:, 3] - b[:, 1]], 1), aw, 1)
        pack_ms = (time.perf_counter() - t0) * 1e3
        d_xy = torch.from_numpy(xy).cuda()
        atlas = torch.zeros((ah, aw), dtype=torch.uint8, device="cuda")

        def run():
            pkg.glyph_boxes_device(d, n, s, s, margin=spread)
            pkg.sdf_u8_batched_device(d, n, s, s, spread=spread, out=out, workspace=ws)
            atlas.zero_()
            pkg.atlas_blit_device(out, n, s, s, boxes, d_xy, atlas, aw)
        ms = timed(run)
        px = n * s * s
        used = int(((b[:, 2] - b[:, 0]) * (b[:, 3] - b[:, 1])).sum())
        print(json.dumps({"config": "c3atlas", "device_ms": round(ms, 4), "host_pack_ms": round(pack_ms, 2), "atlas": [int(ah), aw],
                          "fill": round(used / (ah * aw), 3), "crc": f"{zlib.crc32(atlas.cpu().numpy().tobytes()):08x}"}))
        continue
    else:
        n, spread = 32768, 1024
        d = torch.from_numpy(bench.syn_discs(n)).cuda()
        out = torch.empty((n, n), dtype=torch.uint8, device="cuda")
        ws = torch.empty(pkg.workspace_bytes(n, n, spread), dtype=torch.uint8, device="cuda")
        ms = timed(lambda: pkg.sdf_u8_device(d, n, n, spread=spread, out=out, workspace=ws), iters=3)
        px = n * n
    print(json.dumps({"config": c, "ms": round(ms, 4), "Gpix/s": round(px / ms / 1e6, 1), "GB/s_at_2B_per_px": round(2 * px / ms / 1e6, 1),
                      "crc": f"{zlib.crc32(out.cpu().numpy().tobytes()):08x}", "env_glyph": os.environ.get("CHAQSDF_GLYPH", "")}))
