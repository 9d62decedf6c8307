"""Collective micro-benchmark sweep (the reference's edu.iu.benchmark.JobLauncher, BenchmarkMapper.java:96, :136):
every native collective of the symmetric workspace against its NCCL counterpart, 8 B .. 1 GiB.
    torchrun --nproc-per-node N scripts/coll_sweep.py [max_bytes] > profiles/coll_sweep_nN.md
Time = CUDA events around `iters` back-to-back calls, max over ranks.  busbw follows the usual convention
(allreduce 2(N-1)/N x bytes / t, allgather / reduce-scatter (N-1)/N x bytes / t, others bytes / t)."""
import os, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
import torch.distributed as dist
from harp_b200.parallel.symm import SymmetricWorkspace

rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr); dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
max_bytes = int(sys.argv[1]) if len(sys.argv) > 1 else (1 << 30)
ws = SymmetricWorkspace(dev, 2 * max_bytes + (64 << 20))


def timed(fn, iters):
    for _ in range(3):
        fn()
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record(); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / iters], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item()) * 1e3          # microseconds


if rank == 0:
    print(f"# Collective sweep, {world} x B200, backend={ws.backend}, NVLS={'yes' if ws.has_nvls else 'no'}\n")
    print("| op | bytes | native us | NVLS us | NCCL us | native busbw GB/s | NVLS busbw GB/s | NCCL busbw GB/s |")
    print("|---|---|---|---|---|---|---|---|")
sizes = [8, 64, 1024, 16 << 10, 256 << 10, 1 << 20, 5160000, 16 << 20, 64 << 20, 256 << 20, 1 << 30]
for nbytes in [s for s in sizes if s <= max_bytes]:
    q = world * 64
    n = max(world * 4, nbytes // 4 // q * q if nbytes // 4 >= q else nbytes // 4 // (world * 4) * (world * 4))       # floats
    ws.reset()
    t = ws.alloc(n, torch.float32); t2 = ws.alloc(n, torch.float32)
    u = torch.zeros(n, device=dev); shard = torch.zeros(n // world, device=dev)
    iters = 50 if nbytes <= (1 << 20) else (20 if nbytes <= (64 << 20) else 5)
    pred = (rank - 1) % world
    ops = {
        "allreduce": (lambda: ws.allreduce_(t, "sum", deterministic=True), (lambda: ws.nvls_allreduce_(t)) if ws.has_nvls else None,
                      lambda: dist.all_reduce(u), 2.0 * (world - 1) / world),
        "reduce_scatter (regroup)": (lambda: ws.reduce_scatter_(t), None, lambda: dist.reduce_scatter_tensor(shard, u), (world - 1) / world),
        "allgather": (lambda: ws.allgather_(t), (lambda: ws.nvls_allgather_(t)) if ws.has_nvls else None,
                      lambda: dist.all_gather_into_tensor(u, shard), (world - 1) / world),
        "broadcast": (lambda: ws.broadcast_(t, 0), (lambda: ws.nvls_broadcast_(t, 0)) if ws.has_nvls else None,
                      lambda: dist.broadcast(u, 0), 1.0),
        "reduce": (lambda: ws.reduce_(t, 0), None, lambda: dist.reduce(u, 0), 1.0),
        "rotate": (lambda: ws.rotate_(t2, t, pred), None,
                   lambda: [r.wait() for r in dist.batch_isend_irecv([dist.P2POp(dist.isend, u, (rank + 1) % world),
                                                                      dist.P2POp(dist.irecv, shard if False else u.clone(), pred)])], 1.0),
        "barrier": (lambda: ws.barrier(), None, lambda: dist.barrier(), 0.0),
    }
    rows = n // 64 if (n >= 64 * world and n % 64 == 0) else 0
    if rows:
        loc = t.view(rows, 64); nown = (rows + world - 1) // world
        gl = ws.alloc(nown * 64, torch.float32).view(nown, 64)
        ops["push (K18)"] = (lambda: ws.push_rows_(loc, gl), None, None, (world - 1) / world)
        ops["pull (K18)"] = (lambda: ws.pull_rows_(loc, gl), None, None, (world - 1) / world)
        ops["regroup rows (K2)"] = (lambda: ws.regroup_rows_(loc), None, None, (world - 1) / world)
    for name, (nat, nvls, nccl, f) in ops.items():
        if name == "barrier" and nbytes != 8:
            continue
        a = timed(nat, iters)
        b = timed(nvls, iters) if nvls else None
        c = timed(nccl, iters) if nccl else None
        if rank == 0:
            bw = lambda us: f"{f * n * 4 / us / 1e3:.1f}" if (us and f) else ""
            print(f"| {name} | {n * 4} | {a:.1f} | {'' if b is None else f'{b:.1f}'} | {'' if c is None else f'{c:.1f}'} | {bw(a)} | {bw(b)} | {bw(c)} |", flush=True)
dist.barrier()
dist.destroy_process_group()
