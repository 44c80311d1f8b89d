"""End-to-end time of chaqsdf_u8 (pinned host buffers, H2D + kernels + D2H) for the pipeline's band heights:
   python tools/time_e2e.py [side] [spread]     honours CHAQSDF_PIPE_ROWS / CHAQSDF_PIPE_FIRST"""
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402
import bench  # noqa: E402

side = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
spread = int(sys.argv[2]) if len(sys.argv) > 2 else 256
pkg = ge.load_package()
img = torch.from_numpy(bench.syn_discs(side)).pin_memory()
out = torch.empty((side, side), dtype=torch.uint8).pin_memory()
for _ in range(3):
    pkg.sdf_u8(img, side, side, spread=spread, out=out)
ts = []
for _ in range(10):
    t = time.perf_counter()
    pkg.sdf_u8(img, side, side, spread=spread, out=out)
    ts.append(time.perf_counter() - t)
print(f"PIPE_ROWS={os.environ.get('CHAQSDF_PIPE_ROWS','')} PIPE_FIRST={os.environ.get('CHAQSDF_PIPE_FIRST','')} bands={pkg.pipeline_plan(side, side, 1, spread)} "
      f"e2e ms min {min(ts)*1e3:.3f} median {sorted(ts)[5]*1e3:.3f}")
