"""torchrun --nproc-per-node G profiles/dbg_e2e_multi.py — where the N-GPU end-to-end step of bench.py spends its time
(host wall clock per segment, with a device sync after each: serialised, so the sum is an upper bound of the real step)."""
import os, sys, time
import numpy as np, torch, torch.distributed as dist
sys.path.insert(0, '.')
import gappa_b200
from gappa_b200 import workloads
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
n = int(round(1000 * np.sqrt(world) / 128.0)) * 128
spec, tree, masses = workloads.build("cfg2", samples=n, workers=8)
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
N = spec["samples"]
h_off = torch.from_numpy(masses["off"]); h = [torch.from_numpy(masses[k]).pin_memory() for k in ("edge", "pos", "mass")]
h_out = torch.empty((N, N), dtype=torch.float64).pin_memory()
plan = gappa_b200.KrdPlan(device=local, bins=spec["bins"])
plan.set_stream(torch.cuda.current_stream().cuda_stream)
plan.set_tree(tree["parent"], tree["child"], tree["bl"], tree["root"])
M = int(h[0].numel()); per = (M + world - 1) // world; lo, hi = min(rank * per, M), min((rank + 1) * per, M)
d_part = [torch.zeros(per, dtype=t.dtype, device="cuda") for t in h]
d_full = [torch.empty(per * world, dtype=t.dtype, device="cuda") for t in h]
plan.set_masses(h_off.numpy(), *h)
shard = plan.shard_doubles(world)
d_packed = torch.empty(shard, dtype=torch.float64, device="cuda"); d_gath = torch.empty(shard * world, dtype=torch.float64, device="cuda")
d_out = torch.empty((N, N), dtype=torch.float64, device="cuda")
def seg(name, fn, acc):
    torch.cuda.synchronize(); t0 = time.perf_counter(); fn(); torch.cuda.synchronize(); acc[name] = acc.get(name, 0) + (time.perf_counter() - t0) * 1e3
for it in range(4):
    acc = {}
    dist.barrier()
    def up():
        for part, full, src in zip(d_part, d_full, h):
            part[: hi - lo].copy_(src[lo:hi], non_blocking=True)
    seg("h2d", up, acc)
    def ag():
        for part, full in zip(d_part, d_full): dist.all_gather_into_tensor(full, part)
    seg("allgather_entries", ag, acc)
    seg("set_masses_device", lambda: plan.set_masses_device(h_off.numpy(), d_full[0], d_full[1], d_full[2]), acc)
    seg("compute_shard", lambda: plan.compute_shard(d_packed, rank, world, p=1.0), acc)
    seg("allgather_tiles", lambda: dist.all_gather_into_tensor(d_gath, d_packed), acc)
    seg("unpack", lambda: plan.unpack_shards(d_gath, d_out, world), acc)
    seg("d2h", lambda: h_out.copy_(d_out, non_blocking=True) if rank == 0 else None, acc)
    if rank == 0: print(it, {k: round(v, 3) for k, v in acc.items()}, "sum", round(sum(acc.values()), 3), flush=True)
dist.destroy_process_group()
