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
torch.cuda.synchronize()
print("ALL OK" if ok else "FAILURES", flush=True)
sys.exit(0 if ok else 1)
