"""The "NCCL(+cuBLAS) build of the same algorithms" arm of bench.py (``--impl nccl_cublas``).

Nothing from this repo's kernels runs here: distances / Gram matrices come from cuBLAS through
``torch.addmm`` / ``torch.mm`` (TF32 allowed, the same product precision as the tcgen05 kernels), the arg-min
and the scatter-add are stock PyTorch ops, and the cross-GPU step is ``dist.all_reduce`` (NCCL).  It exists so
that BASELINE.md §3's middle column is measured on the same box, by the same timing harness, on the same
synthetic data as the product arm.
"""
from __future__ import annotations

import json

import torch
import torch.distributed as dist


def run(args, rank: int, world: int, dev: torch.device, sampler_cls) -> int:
    out = measure(args, rank, world, dev, sampler_cls, args.workload, args.steps, args.warmup)
    if rank == 0:
        print(json.dumps(out))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def measure(args, rank: int, world: int, dev: torch.device, sampler_cls, workload: str, steps: int, warmup: int):
    """Times ``steps`` iterations of the library-only arm; returns the result dict on rank 0 (None elsewhere)."""
    torch.backends.cuda.matmul.allow_tf32 = True

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    if workload == "kmeans":
        n, d, k = args.points_per_gpu, args.dim, args.k
        g = torch.Generator(device=dev).manual_seed(1234 + rank)
        X = torch.rand(n, d, device=dev, generator=g)
        C = torch.rand(k, d, device=dev, generator=torch.Generator(device=dev).manual_seed(99))
        acc = torch.zeros(k * d + k, device=dev)
        sums, counts = acc[: k * d].view(k, d), acc[k * d:]
        ones = torch.ones(n, device=dev)
        best = torch.empty(n, dtype=torch.int64, device=dev)
        # 8192 x 10 000 fp32 distances per GEMM: small enough that arg-min reads them back out of L2-sized traffic
        # instead of streaming a 10 GB block through HBM (the first version of this arm used 262 144 rows)
        chunk = 8192

        def step():
            cn = (C * C).sum(1)
            Ct = C.t().contiguous()
            for s in range(0, n, chunk):
                dist2 = torch.addmm(cn, X[s:s + chunk], Ct, alpha=-2.0)       # 1-D bias: the cuBLASLt epilogue path
                best[s:s + chunk] = dist2.argmin(1)
            acc.zero_()
            sums.index_add_(0, best, X)
            counts.index_add_(0, best, ones)
            if world > 1:
                dist.all_reduce(acc)
            nz = counts > 0
            C.copy_(torch.where(nz[:, None], sums / counts.clamp(min=1.0)[:, None], torch.zeros_like(sums)))
        metric = {"metric": "kmeans_allreduce_s_per_iter", "unit": "s/iter", "higher_is_better": False, "scaling": "weak",
                  "dtype": "tf32 (cuBLAS, allow_tf32)", "config": {"model": "K-means allreduce (BASELINE config 2)",
                  "global_batch": n * world, "points_per_gpu": n, "dim": d, "centroids": k, "parallelism": f"dp{world}",
                  "comm": "nccl", "l2": "inputs larger than L2"}}
        flops = 2.0 * n * world * k * d
    elif workload == "gram":
        d = 1024
        n = min(50_000_000 // world, 6_250_000)
        X = torch.randn(n, d, device=dev)
        G = torch.empty(d, d, device=dev)

        def step():
            torch.mm(X.t(), X, out=G)
            s = X.sum(0, dtype=torch.float64)
            if world > 1:
                dist.all_reduce(G)
                dist.all_reduce(s)
        metric = {"metric": "gram_s_per_pass", "unit": "s/pass", "higher_is_better": False, "scaling": "weak", "dtype": "tf32 (cuBLAS, allow_tf32)",
                  "config": {"model": "covariance/PCA/linreg partial results + allreduce (BASELINE config 5)",
                             "rows_per_gpu": n, "dim": d, "parallelism": f"dp{world}", "l2": "X exceeds L2"}}
        flops = 2.0 * n * world * d * d
    else:
        return {"impl": "nccl_cublas", "unavailable": f"no library-only formulation of workload {workload!r}"} if rank == 0 else None

    for _ in range(warmup):
        step()
    barrier()
    sampler = sampler_cls(dev.index) if rank == 0 else None
    if sampler:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(steps):
        step()
    e1.record()
    barrier()
    t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item()) / steps
    clocks = sampler.stop() if sampler else None
    if rank == 0:
        out = {"impl": "nccl_cublas", "value": ms / 1e3, "n_gpus": world, "steps": steps, "warmup": warmup,
               "ms_per_step": ms, "vs_baseline": None, "data": "synthetic", "gpu_launches": 0, "clocks": clocks,
               "tflops_tf32_aggregate": flops / (ms / 1e3) / 1e12}
        out.update(metric)
        return out
    return None
