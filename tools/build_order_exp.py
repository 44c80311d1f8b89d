"""Experiment (library built with -DCQ_BUILD_TRACE): what would heaviest-first ordering of the build grid give?
Runs the build once, sorts the CTAs by their measured duration, feeds that order back and times the build again."""
import ctypes, sys, numpy as np, torch
sys.path.insert(0, "/root/repo")
import __graft_entry__ as ge, bench
pkg = ge.load_package(); L = pkg.lib()
side, spread = 16384, 256
d = torch.from_numpy(bench.syn_discs(side)).cuda()
out = torch.empty((side, side), dtype=torch.uint8, device="cuda")
ws = torch.empty(pkg.workspace_bytes(side, side, spread), dtype=torch.uint8, device="cuda")
n = 8192
def run(label):
    for _ in range(2):
        pkg.sdf_u8_device(d, side, side, spread=spread, out=out, workspace=ws)
    torch.cuda.synchronize()
    pkg.profile_enable(True)
    for _ in range(5):
        pkg.sdf_u8_device(d, side, side, spread=spread, out=out, workspace=ws)
    torch.cuda.synchronize()
    ms = pkg.profile_mean_ms(); pkg.profile_enable(False)
    buf = (ctypes.c_ulonglong * (3 * n))()
    L.chaqsdf_debug_build_trace(buf, n)
    a = np.array(buf, dtype=np.uint64).reshape(n, 3).astype(np.int64)
    dur = (a[:, 1] - a[:, 0]) / 1e3
    print(label, "build ms", round(ms["hull_build_kernel"], 4), "span us", (a[:, 1].max() - a[:, 0].min()) / 1e3, "sum of CTA us", round(dur.sum()), "max", dur.max())
    return dur
import zlib
dur = run("spatial order ")
crc0 = zlib.crc32(out.cpu().numpy().tobytes())
order = torch.from_numpy(np.argsort(-dur).astype(np.uint32).view(np.int32)).cuda()
L.chaqsdf_debug_set_build_order.argtypes = [ctypes.c_void_p, ctypes.c_uint]
L.chaqsdf_debug_set_build_order(ctypes.c_void_p(order.data_ptr()), n)
run("heaviest first")
print("same bytes", zlib.crc32(out.cpu().numpy().tobytes()) == crc0)
# a coarse order: only 8 classes of cost
q = np.digitize(dur, np.quantile(dur, [0.5, 0.75, 0.9, 0.95, 0.98, 0.99]))
order2 = torch.from_numpy(np.argsort(-q, kind="stable").astype(np.uint32).view(np.int32)).cuda()
L.chaqsdf_debug_set_build_order(ctypes.c_void_p(order2.data_ptr()), n)
run("7 cost classes")

# ---- which cheap statistic of pass 1's record predicts the cost? (per item = 64 columns x 512 rows)
rb = pkg.inter_row_bytes(side, spread)
inter = torch.empty((side, rb), dtype=torch.uint8, device="cuda")
pkg.rowscan_device(d, side, side, spread, inter)
g = inter[:, :side]                                   # g8 = min(G, C) - 1
def tiles(x):                                         # [rows, cols] -> per (strip of 64 cols, run of 512 rows), item = run + 32 * strip
    return x.reshape(32, 512, 256, 64).float().sum(dim=(1, 3)).t().reshape(-1).cpu().numpy()
stats = {
    "px with g<255": tiles(g < 255), "px with g<128": tiles(g < 128), "px with g<64": tiles(g < 64), "px with g<16": tiles(g < 16),
    "sum of 1/(g+1)": tiles(1.0 / (g.float() + 1.0)),
}
busy = inter[:, side + side // 8: side + side // 8 + side // 128]      # one byte per 128-column strip
bz = (busy != 0).reshape(32, 512, 128).float().sum(dim=1).t()            # [strip128, run]
stats["busy rows (128-col strips)"] = bz.repeat_interleave(2, dim=0).reshape(-1).cpu().numpy()
from scipy.stats import spearmanr
L.chaqsdf_debug_set_build_order(None, 0)
dur = run("spatial again ")
for k, v in stats.items():
    rho = spearmanr(v, dur).correlation
    cls = np.digitize(v, np.quantile(v, [0.5, 0.75, 0.9, 0.95, 0.98, 0.99]))
    o = torch.from_numpy(np.argsort(-cls, kind="stable").astype(np.uint32).view(np.int32)).cuda()
    L.chaqsdf_debug_set_build_order(ctypes.c_void_p(o.data_ptr()), n)
    print(f"proxy {k}: spearman {rho:.3f}")
    run("   ordered by it")
