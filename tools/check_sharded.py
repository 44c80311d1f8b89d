"""torchrun worker: the row-band sharded path over NCCL must give the same bytes as the single-GPU path."""
import os, sys
import numpy as np, torch, torch.distributed as dist
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge
import bench
pkg = ge.load_package()
rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", lr))
ok = True
for (n, spread) in ((4096, 64), (8192, 256), (3000, 300)):
    img = bench.syn_discs(n)
    variants = [img, np.ascontiguousarray(img[::-1]), np.ascontiguousarray(img.T), img]  # several calls per state:
    r0, r1 = pkg.shard_rows(n, world, rank)                                           # the two slots alternate
    for halo in ("peer", "p2p"):
        state = {}
        for k, v in enumerate(variants):
            band = torch.from_numpy(v[r0:r1]).cuda()
            out = pkg.sdf_u8_sharded(band, n, n, spread, False, rank, world, None, state, halo=halo)
            full = pkg.sdf_u8_device(torch.from_numpy(v).cuda(), n, n, spread=spread)
            torch.cuda.synchronize()
            same = bool(torch.equal(out, full[r0:r1]))
            ok &= same
            print(f"rank {rank}: {n}^2 spread {spread} halo {halo}->{state['halo']} call {k} rows {r0}:{r1} identical={same}"
                  + (f" ({state.get('peer_error')})" if state["halo"] != halo else ""), flush=True)
t = torch.tensor([1 if ok else 0], device="cuda")
dist.all_reduce(t, op=dist.ReduceOp.MIN)
if rank == 0:
    print("SHARDED CHECK", "PASS" if int(t.item()) == 1 else "FAIL", flush=True)
dist.destroy_process_group()
