"""Secondary bench.py workloads (BASELINE configs 3, 4, 5): ``--workload sgd | lda | gram``.

Same contract as the K-means default: W untimed warm-up steps, K timed steps between barriers +
synchronises, CUDA-event time, max over ranks, ONE JSON line from rank 0.  Strong scaling: the
named global problem is split over the N GPUs.
"""
from __future__ import annotations

import json
import math
import time

import os

import torch
import torch.distributed as dist

from . import _native


class _Steps:
    def __init__(self, steps, warmup):
        self.steps, self.warmup = int(steps), int(warmup)


def _finish(args, rank, world, dev, ms_local, out, sampler=None):
    t = torch.tensor([ms_local], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item()) / args.steps
    out.update({"n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "impl": "harp_b200",
                "data": "synthetic", "vs_baseline": None, "gpu_launches": _native.launches(),
                "clocks": sampler.stop() if sampler is not None else None})
    return ms, out


def run_workload(args, rank, world, dev, sampler_cls=None):
    """``bench.py --workload sgd|lda|gram``: measures one workload and prints its JSON line on rank 0."""
    symm = None
    if world > 1:
        from .parallel.symm import SymmetricWorkspace
        try:
            symm = SymmetricWorkspace(dev, bytes_=(3 << 29) if args.workload != "lda10k" else (12 << 30))
        except Exception:  # noqa: BLE001
            symm = None
    out = measure_workload(args.workload, args.steps, args.warmup, rank, world, dev, symm, sampler_cls)
    if rank == 0:
        print(json.dumps(out))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def measure_workload(workload, steps, warmup, rank, world, dev, symm, sampler_cls=None):
    """W warm-up steps, then ``steps`` timed steps between barriers + synchronises, CUDA-event time, max over ranks,
    clocks sampled during the timed region.  Returns the result dict on rank 0 (None elsewhere)."""
    args = _Steps(steps, warmup)
    args.workload = workload
    result = None

    def start_sampler():
        if sampler_cls is None or rank != 0:
            return None
        smp = sampler_cls(dev.index if dev.index is not None else 0)
        smp.start()
        return smp

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    if args.workload == "sgd":
        from .models.mfsgd import MFSGDEngine
        rows, cols, nnz, rk = 480_000, 17_700, 100_000_000, 128
        n_local = nnz // world
        g = torch.Generator(device=dev).manual_seed(11 + rank)
        rows_local = (rows + world - 1) // world
        r = (torch.rand(n_local, device=dev, generator=g) ** 2 * rows_local).long() % rows_local
        c = (torch.rand(n_local, device=dev, generator=g) ** 2 * cols).long() % cols
        v = torch.rand(n_local, device=dev, generator=g) * 4 + 1
        eng = MFSGDEngine(r, c, v, rows_local, cols, rk, 0.001, 0.01, symm=symm, seed=3)
        del r, c, v
        for _ in range(args.warmup):
            eng.epoch()
        eng.synchronize(); barrier()
        _native.reset_launches()
        smp = start_sampler()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier(); e0.record()
        for _ in range(args.steps):
            eng.epoch()
        e1.record(); eng.synchronize(); barrier()
        ms, out = _finish(args, rank, world, dev, e0.elapsed_time(e1), {}, smp)
        if rank == 0:
            out.update({"metric": "mfsgd_updates_per_s", "value": nnz / (ms / 1e3), "unit": "updates/s",
                        "higher_is_better": True, "scaling": "strong", "dtype": "fp32",
                        "config": {"model": "MF-SGD model rotation (BASELINE config 3)", "rows": rows, "cols": cols,
                                   "ratings": nnz, "rank": rk, "parallelism": f"rows dp{world} + H rotation",
                                   "l2": "W (245 MB) + ratings (1.2 GB) exceed L2"}})
            result = out
    elif args.workload in ("lda", "lda10k"):
        from .models.lda import LDAEngine
        D, V, K, L = 1_000_000, 100_000, 1000, 100
        if args.workload == "lda10k":        # the shape of the reference's published LDA runs: K = 10 000, V = 1 M
            V, K = 1_000_000, 10_000
        d_local = D // world
        g = torch.Generator(device=dev).manual_seed(5 + rank)
        docs = torch.arange(d_local, device=dev).repeat_interleave(L)
        words = (torch.rand(d_local * L, device=dev, generator=g) ** 3 * V).long() % V
        eng = LDAEngine(docs, words, d_local, V, K, 0.01, 0.01, symm=symm, seed=1)
        for _ in range(args.warmup):
            eng.iterate()
        eng.synchronize(); barrier()
        _native.reset_launches()
        smp = start_sampler()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier(); e0.record()
        for _ in range(args.steps):
            eng.iterate()
        e1.record(); eng.synchronize(); barrier()
        ms, out = _finish(args, rank, world, dev, e0.elapsed_time(e1), {}, smp)
        if rank == 0:
            out.update({"metric": "lda_cgs_tokens_per_s", "value": D * L / (ms / 1e3), "unit": "tokens/s",
                        "higher_is_better": True, "scaling": "strong", "dtype": "int32 counts / fp32 probabilities",
                        "sampler": eng.sampler,
                        "config": {"model": "LDA-CGS model rotation (BASELINE config 4)" if K == 1000 else
                                   "LDA-CGS model rotation, K = 10 000 topics, 1 M vocabulary (clueweb-shaped, lda-cgs.md:36-40)",
                                   "docs": D, "vocab": V,
                                   "topics": K, "tokens": D * L, "parallelism": f"docs dp{world} + word-topic rotation",
                                   "l2": "doc-topic table (4 GB) exceeds L2"}})
            result = out
    elif args.workload == "ccd":
        # MF-CCD++ on the MF-SGD matrix (BASELINE config 3 shape), r = 120 as in the reference's clueweb run
        # (website/content/docs/applications/mf.md:63): one step = one outer iteration = 120 features x (H sweep + W sweep)
        from .models.mfccd import MFCCDEngine
        nr, nc, nnz, r = 480_000, 17_700, 100_000_000, 120
        g = torch.Generator(device=dev).manual_seed(17)          # the SAME rating set on every rank, then split by owner
        rows = (torch.rand(nnz, device=dev, generator=g) ** 2 * nr).long() % nr
        cols = (torch.rand(nnz, device=dev, generator=g) ** 2 * nc).long() % nc
        vals = torch.rand(nnz, device=dev, generator=g) * 4 + 1
        mr, mc = rows % world == rank, cols % world == rank
        eng = MFCCDEngine(rows[mr], cols[mr], vals[mr], rows[mc], cols[mc], vals[mc], nr, nc, r, 0.1, symm=symm, seed=1)
        del rows, cols, vals, mr, mc
        torch.cuda.empty_cache()
        e_first = None
        for _ in range(args.warmup):
            e = eng.iterate()
            e_first = e if e_first is None else e_first
        barrier()
        _native.reset_launches()
        smp = start_sampler()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier(); e0.record()
        for _ in range(args.steps):
            e_last = eng.iterate()
        e1.record(); barrier()
        ms, out = _finish(args, rank, world, dev, e0.elapsed_time(e1), {}, smp)
        if rank == 0:
            out.update({"metric": "mfccd_rating_visits_per_s", "value": 2.0 * r * nnz / (ms / 1e3), "unit": "rating visits/s",
                        "higher_is_better": True, "scaling": "strong", "dtype": "fp32",
                        "s_per_iteration": ms / 1e3, "train_rmse_first_last": [e_first, e_last],
                        "config": {"model": "MF-CCD++ (W sharded by rows, H by columns, rank-one vectors all-gathered over NVLink)",
                                   "rows": nr, "cols": nc, "ratings": nnz, "r": r, "parallelism": f"rows/cols dp{world}",
                                   "l2": "two residual orders + index lists (1.6 GB per rank-share) exceed L2"}})
            result = out
    else:  # gram
        from .ops.gram import gram, moments
        n, d = 50_000_000 // world, 1024
        n = min(n, 6_250_000)
        X = torch.randn(n, d, device=dev)
        from .ops.gram import FusedGram
        s = symm.alloc(d, torch.float64) if symm is not None else None
        fused = None
        if world == 1 or symm is not None:          # Gram + partial-result allreduce in one launch
            nb = FusedGram.state_bytes(d)
            if symm is not None:
                st = symm.alloc(nb, torch.uint8)
                fused = FusedGram(st, [p + symm.offset_of(st) for p in symm.peer_ptrs], rank, d)
            else:
                st = torch.empty(nb, dtype=torch.uint8, device=dev)
                fused = FusedGram(st, [st.data_ptr()], 0, d)
            barrier()
        G = torch.empty(d, d, device=dev) if fused is None else None
        fuse_stats = os.environ.get("HB_GRAM_FUSED_STATS", "0") == "1"   # 1: moments inside the Gram launch (A/B; not faster under the power cap)

        def step():
            if fused is not None and fuse_stats:
                _, m = fused(X, want_moments=True)     # Gram + allreduce + column moments: one pass over X
            elif fused is not None:
                fused(X)
                m = moments(X)
            else:
                gram(X, out=G)
                m = moments(X)
            if world > 1:
                if symm is not None:
                    s.copy_(m.sum); symm.allreduce_sum_(s)
                else:
                    dist.all_reduce(G); dist.all_reduce(m.sum)
        for _ in range(args.warmup):
            step()
        barrier()
        _native.reset_launches()
        smp = start_sampler()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier(); e0.record()
        for _ in range(args.steps):
            step()
        e1.record(); barrier()
        ms, out = _finish(args, rank, world, dev, e0.elapsed_time(e1), {}, smp)
        if rank == 0:
            out.update({"metric": "gram_s_per_pass", "value": ms / 1e3, "unit": "s/pass", "higher_is_better": False,
                        "scaling": "weak", "dtype": "tf32",
                        "config": {"model": "covariance/PCA/linreg partial results + allreduce (BASELINE config 5)",
                                   "rows_per_gpu": n, "dim": d, "parallelism": f"dp{world}",
                                   "l2": "X (25.6 GB per GPU) exceeds L2"},
                        "tflops_tf32_aggregate": 2.0 * n * world * d * d / (ms / 1e3) / 1e12})
            result = out
    return result
