"""Multi-rank check + timing of the exact fields over several GPUs (all-to-all between the passes, variant E1), and of
the row-band (halo) variant next to it. Launch:
   python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 tools/check_a2a.py [side] [iters]
Every rank checks its column part of d2_in / d2_out / out against the single-GPU exact fields (byte for byte) in both
modes ("store": pass 1 writes into the peers' buffers over NVLink; "a2a": NCCL all_to_all), rank 0 prints one JSON line."""
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402
import bench  # noqa: E402

side = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 5
spread = 256
rank, world, lr = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(lr)
dev = torch.device("cuda", lr)
if world > 1:
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
pkg = ge.load_package()
img = bench.syn_discs(side)
r0, r1 = pkg.shard_rows(side, world, rank)
c0, c1 = pkg.col_part(side, world, rank)
d_band = torch.from_numpy(img[r0:r1]).to(dev)
whole = pkg.sdf_fields_device(torch.from_numpy(img).to(dev), side, side, spread=spread, want=("d2_in", "d2_out", "out"))
line = {"what": "exact fields (d2_in, d2_out, u8) of SYN-DISCS(%d), spread %d, over %d GPU(s): row bands -> all-to-all -> column parts" % (side, spread, world),
        "n_gpus": world}
for mode in (("store", "a2a") if world > 1 else ("a2a",)):
    state = {}
    ok = True
    for it in range(2):
        res = pkg.sdf_fields_a2a(d_band, side, side, spread=spread, rank=rank, world=world, state=state, mode=mode)
        torch.cuda.synchronize()
        for k in ("d2_in", "d2_out", "out"):
            ok = ok and torch.equal(res[k], whole[k][:, c0:c1])
    evs = [[torch.cuda.Event(enable_timing=True) for _ in range(4)] for _ in range(iters)]
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    for e in evs:
        pkg.sdf_fields_a2a(d_band, side, side, spread=spread, rank=rank, world=world, state=state, mode=mode, events=e)
    torch.cuda.synchronize()
    t = torch.tensor([min(e[0].elapsed_time(e[3]) for e in evs), min(e[0].elapsed_time(e[1]) for e in evs),
                      min(e[1].elapsed_time(e[2]) for e in evs), min(e[2].elapsed_time(e[3]) for e in evs), 1.0 if ok else 0.0], device=dev)
    if world > 1:
        tmax = t.clone()
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        tmin = t.clone()
        dist.all_reduce(tmin, op=dist.ReduceOp.MIN)
        t = torch.cat([tmax[:4], tmin[4:]])
    line[state["a2a"] if mode == "store" else "a2a"] = {
        "requested": mode, "ran": state["a2a"], "ms": round(float(t[0]), 3), "pass1_ms": round(float(t[1]), 3),
        "exchange_ms": round(float(t[2]), 3), "pass2_ms": round(float(t[3]), 3), "parity_vs_single_gpu": bool(t[4] > 0.5),
        "Mpix/s": round(side * side / float(t[0]) / 1e3, 1), "peer_error": state.get("peer_error")}
if rank == 0:
    print(json.dumps(line), flush=True)
if world > 1:
    dist.destroy_process_group()
