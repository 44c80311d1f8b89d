"""Every kernel of the library once, at small sizes, for compute-sanitizer (memcheck / racecheck):
   compute-sanitizer --tool memcheck python tools/sanitize_small.py
Checks the bytes against the oracle as it goes, so a silent wrong result under the tool shows too."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import __graft_entry__ as ge  # noqa: E402
from oracle import Oracle  # noqa: E402

pkg, o = ge.load_package(), Oracle()
rng = np.random.default_rng(1)
ok = True


def shapes(h, w):
    yy, xx = np.mgrid[0:h, 0:w]
    m = np.zeros((h, w), np.uint8)
    for _ in range(6):
        cy, cx, r = rng.integers(0, h), rng.integers(0, w), rng.integers(2, max(3, min(h, w) // 3))
        m |= ((yy - cy) ** 2 + (xx - cx) ** 2 <= r * r).astype(np.uint8)
    return m * 255


def check(name, got, want):
    global ok
    same = np.array_equal(got, want)
    ok &= same
    print(f"{name}: {'ok' if same else 'MISMATCH'}", flush=True)


for (w, h, spread) in ((300, 200, 12), (130, 700, 256), (640, 96, 300), (96, 1300, 1024), (33, 77, 3000)):
    img = shapes(h, w)
    d = torch.from_numpy(img).cuda()
    check(f"u8_device {w}x{h} s{spread}", pkg.sdf_u8_device(d, w, h, spread=spread).cpu().numpy(), o.sdf(img, w, h, spread=spread))
imgs = np.stack([shapes(64, 96) for _ in range(5)])
got = pkg.sdf_u8_batched_device(torch.from_numpy(imgs).cuda(), 5, 96, 64, spread=16).cpu().numpy()
check("glyph batch", got, np.stack([o.sdf(imgs[k], 96, 64, spread=16) for k in range(5)]))
imgs = np.stack([shapes(50, 41) for _ in range(3)])
got = pkg.sdf_u8_batched_device(torch.from_numpy(imgs).cuda(), 3, 41, 50, spread=16).cpu().numpy()
check("narrow batch", got, np.stack([o.sdf(imgs[k], 41, 50, spread=16) for k in range(3)]))
img = shapes(260, 300)
res = pkg.sdf_fields_device(torch.from_numpy(img).cuda(), 300, 260, spread=20)
out, sdf, din, dout = o.sdf(img, 300, 260, spread=20, fields=True)
check("fields d2_in", res["d2_in"].cpu().numpy(), din)
check("fields out", res["out"].cpu().numpy(), out)
res = pkg.sdf_fields_a2a(torch.from_numpy(img).cuda(), 300, 260, spread=20)
check("fields via scatter", res["d2_out"].cpu().numpy(), dout)
check("host u8", pkg.sdf_u8(img, 300, 260, spread=20), out)
check("multi [0,0,0]", pkg.sdf_u8_multi(img, 300, 260, [0, 0, 0], spread=20), out)
f = np.where(rng.random((90, 90)) < 0.01, np.float32(0), np.float32(np.inf)).astype(np.float32)
check("dist_transform_2d", pkg.dist_transform_2d(f.copy()).view(np.uint32), o.dist_transform_2d_binary(f).view(np.uint32))
g = rng.random((40, 40)).astype(np.float32) * 50
r = pkg.dist_transform_2d(g.copy())
print("dt general finite:", bool(np.isfinite(r).all()), flush=True)
# round-2 additions: transposed bands (transposition + column-range pass 2), atlas front-end, static scratch, tall runs
img = shapes(600, 333)
want = o.sdf(img, 333, 600, spread=40)
d = torch.from_numpy(img).cuda()
for rank in range(3):
    r0, r1, ht, hb, _, _ = pkg.halo_plan(600, 3, rank, 40)
    ext = d[r0 - ht:r1 + hb].contiguous()
    check(f"tband rank {rank}", pkg.sdf_u8_tband_device(ext, 333, ext.shape[0], ht, r1 - r0, 40).cpu().numpy(), want[r0:r1])
gl = np.stack([shapes(96, 96) for _ in range(6)])
atlas, boxes, xy = pkg.atlas_u8(gl, 6, 96, 96, spread=8, atlas_w=256, pad=1)
okb = True
for i in range(6):
    x0, y0, x1, y1 = boxes[i].tolist()
    okb &= np.array_equal(atlas[xy[i][1]:xy[i][1] + y1 - y0, xy[i][0]:xy[i][0] + x1 - x0], o.sdf(gl[i], 96, 96, spread=8)[y0:y1, x0:x1])
check("atlas", np.array([okb]), np.array([True]))
outs = torch.empty((600, 333), dtype=torch.uint8, device="cuda")
check("static scratch", pkg.sdf_u8_device_static(d, 333, 600, outs, spread=40).cpu().numpy(), want)
img = shapes(2100, 96)
check("tall u8_device 96x2100 s100", pkg.sdf_u8_device(torch.from_numpy(img).cuda(), 96, 2100, spread=100).cpu().numpy(), o.sdf(img, 96, 2100, spread=100))
torch.cuda.synchronize()
print("ALL OK" if ok else "FAILURES", flush=True)
sys.exit(0 if ok else 1)
