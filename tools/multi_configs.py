"""torchrun worker: the multi-GPU cases of BASELINE.json besides config 4 -- config 3 (4096 glyphs of 256^2, spread 16,
sharded by image, no communication) and config 5 (32768^2, spread 1024, row bands + halo exchange) -- device-timed
(CUDA events, barrier on both sides, max over ranks) and checked byte for byte against single-GPU results.
Not bench lines. usage: torchrun --nproc-per-node N tools/multi_configs.py [c3] [c5]"""
import json, os, sys
import numpy as np, torch, torch.distributed as dist
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge
import bench
pkg = ge.load_package()
rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dev = torch.device("cuda", lr)
dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
which = [a for a in sys.argv[1:] if a in ("c3", "c5")] or ["c3", "c5"]


def timed(fn, iters=5, warm=2):
    for _ in range(warm):
        fn()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize(); dist.barrier()
    t = torch.tensor([a.elapsed_time(b) / iters], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def agree(ok):
    t = torch.tensor([1 if ok else 0], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    return bool(t.item())


if "c3" in which:
    g, n, s = 4096, 256, 16
    rng = np.random.default_rng(3)
    yy, xx = np.mgrid[0:n, 0:n]
    base = []
    for i in range(64):
        m = np.zeros((n, n), bool)
        for _ in range(int(rng.integers(2, 6))):
            cy, cx, r = rng.integers(40, n - 40), rng.integers(40, n - 40), rng.integers(10, 60)
            m |= np.abs(np.hypot(yy - cy, xx - cx) - r) < rng.integers(3, 10)
        base.append(m)
    imgs = (np.stack([base[i % 64] for i in range(g)]) * 255).astype(np.uint8)
    g0, g1 = pkg.shard_rows(g, world, rank)  # glyphs [g0, g1) belong to this rank
    mine = torch.from_numpy(imgs[g0:g1]).to(dev)
    ng = g1 - g0
    ws = torch.empty(pkg.workspace_bytes(n, n, s, ng), dtype=torch.uint8, device=dev)
    out = torch.empty((ng, n, n), dtype=torch.uint8, device=dev)
    ms = timed(lambda: pkg.sdf_u8_batched_device(mine, ng, n, n, spread=s, out=out, workspace=ws))
    full = pkg.sdf_u8_batched_device(torch.from_numpy(imgs).to(dev), g, n, n, spread=s)  # the whole atlas on one GPU
    torch.cuda.synchronize()
    ok = agree(bool(torch.equal(out, full[g0:g1])))
    if rank == 0:
        print(json.dumps({"config": "config3: 4096 glyphs of 256x256, spread 16, sharded by image", "n_gpus": world,
                          "ms": ms, "Mpix/s": g * n * n / ms / 1e3, "identical_to_single_gpu": ok, "collective": "none"}), flush=True)
    del full, out, mine, ws

if "c5" in which:
    n, s = 32768, 1024
    r0, r1, ht, hb, _, _ = pkg.halo_plan(n, world, rank, s)
    ctx = bench.syn_discs(n, rows=(r0 - ht, r1 + hb))       # this rank's band with the rows its pass 2 can see
    band = torch.from_numpy(ctx[ht:ht + (r1 - r0)]).to(dev)
    state = {}
    ms = timed(lambda: pkg.sdf_u8_sharded(band, n, n, s, False, rank, world, None, state), iters=3, warm=2)
    out = pkg.sdf_u8_sharded(band, n, n, s, False, rank, world, None, state)
    # single-GPU check: the unsharded path on the band plus `spread` rows of context gives the same middle rows
    # (every pixel's distance is clamped at spread, so rows further away cannot matter)
    ref = pkg.sdf_u8_device(torch.from_numpy(ctx).to(dev), n, ctx.shape[0], spread=s)
    torch.cuda.synchronize()
    ok = agree(bool(torch.equal(out, ref[ht:ht + (r1 - r0)])))
    if rank == 0:
        print(json.dumps({"config": "config5: SYN-DISCS(32768), spread 1024, row bands + halo exchange", "n_gpus": world,
                          "ms": ms, "Mpix/s": n * n / ms / 1e3, "identical_to_single_gpu_on_band_plus_context": ok,
                          "halo": state.get("halo"), "halo_rows_per_side": s}), flush=True)
dist.destroy_process_group()
