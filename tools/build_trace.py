"""Debug: per-CTA start/end times of hull_build_kernel (library built with -DCQ_BUILD_TRACE)."""
import ctypes, sys, numpy as np, torch
sys.path.insert(0, "/root/repo")
import __graft_entry__ as ge, bench
pkg = ge.load_package(); L = pkg.lib()
side, spread = 16384, 256
d = torch.from_numpy(bench.syn_discs(side)).cuda()
for _ in range(3):
    out = pkg.sdf_u8_device(d, side, side, spread=spread)
torch.cuda.synchronize()
n = 8192
buf = (ctypes.c_ulonglong * (3 * n))()
L.chaqsdf_debug_build_trace(buf, n)
a = np.array(buf, dtype=np.uint64).reshape(n, 3)
t0 = a[:, 0].astype(np.int64); t1 = a[:, 1].astype(np.int64); sm = a[:, 2]
base = t0.min()
dur = (t1 - t0) / 1e3
print("kernel span us", (t1.max() - base) / 1e3, "CTAs", n)
print("CTA duration us: mean %.1f median %.1f p90 %.1f p99 %.1f max %.1f" % (dur.mean(), np.median(dur), np.percentile(dur, 90), np.percentile(dur, 99), dur.max()))
print("sum of durations / (148 SMs * span): %.2f CTAs resident on average" % (dur.sum() / ((t1.max() - base) / 1e3) / 148))
order = np.argsort(t1)[::-1][:10]
for i in order:
    print("late CTA", i, "run", i % 32, "strip", i // 32, "start %.1f end %.1f dur %.1f sm %d" % ((t0[i] - base) / 1e3, (t1[i] - base) / 1e3, dur[i], sm[i]))
# how many CTAs are running over time
ts = np.linspace(0, (t1.max() - base) / 1e3, 21)
print("resident CTAs over time:", [int(((t0 - base) / 1e3 <= t).sum() - ((t1 - base) / 1e3 <= t).sum()) for t in ts])
# duration by start time decile
st = (t0 - base) / 1e3
for lo in range(0, 500, 50):
    m = (st >= lo) & (st < lo + 50)
    if m.any(): print("started in [%d,%d) us: n=%d mean dur %.1f max %.1f" % (lo, lo + 50, m.sum(), dur[m].mean(), dur[m].max()))
