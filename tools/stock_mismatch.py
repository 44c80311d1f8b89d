"""Reports what SURVEY.md 8(c) Tier P2 asks for: how many bytes of the STOCK reference build (float32
parabola_intersect, inexact beyond side 4098) differ from the exact bytes (oracle/edt_oracle.c, pinned above 4098 by
the double-intersect build oracle/_ref/libchaqref_dbl.so) on the disc masks of configs 4 and 5's kind. CPU only.
   python tools/stock_mismatch.py [out.json]        writes profiles/r02_stock_mismatch.json by default"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
from oracle import Oracle, Ref  # noqa: E402

o, stock, dbl = Oracle(), Ref(), Ref("_dbl")
rows = []
for n, spread in ((4096, 64), (8192, 64), (16384, 256)):
    img = o.syn_discs(n)
    t = time.time()
    exact = o.sdf(img, n, n, spread=spread)
    a = stock.sdf_u8(img, n, n, spread=spread)
    d = a.astype(np.int16) - exact.astype(np.int16)
    row = {"side": n, "spread": spread, "mask": f"SYN-DISCS({n})", "stock_vs_exact_mismatching_bytes": int((d != 0).sum()),
           "stock_vs_exact_fraction": float((d != 0).mean()), "stock_vs_exact_max_abs_delta": int(np.abs(d).max())}
    if n <= 8192:
        b = dbl.sdf_u8(img, n, n, spread=spread)
        row["double_intersect_build_vs_exact_mismatching_bytes"] = int((b != exact).sum())
    row["seconds"] = round(time.time() - t, 1)
    rows.append(row)
    print(row, flush=True)
out = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "profiles", "r02_stock_mismatch.json")
with open(out, "w") as f:
    json.dump({"what": "bytes of the stock float32 reference build that differ from the exact EDT bytes (which the CUDA path "
               "reproduces bit for bit); the stock build is exact up to side 4098 and wrong in the near field beyond",
               "host": f"{os.cpu_count()} vCPU container, oracle/_ref builds of /root/reference", "rows": rows}, f, indent=1)
