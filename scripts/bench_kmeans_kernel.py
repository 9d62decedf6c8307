"""Kernel-level timing of the K-means step (CUDA events, L2 flushed between iterations)."""
import json
import sys
import time
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch

from harp_b200 import _native
from harp_b200.ops.kmeans import kmeans_step


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 2_000_000
    d = int(sys.argv[2]) if len(sys.argv) > 2 else 128
    k = int(sys.argv[3]) if len(sys.argv) > 3 else 10_000
    iters = int(sys.argv[4]) if len(sys.argv) > 4 else 5
    variant = sys.argv[5] if len(sys.argv) > 5 else "auto"
    torch.manual_seed(0)
    X = torch.rand(n, d, device="cuda")
    C = X[torch.randperm(n, device="cuda")[:k]].clone()
    sums = torch.empty(k, d, device="cuda"); counts = torch.empty(k, device="cuda")
    ws = torch.empty(1 << 20, dtype=torch.uint8, device="cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    lib = _native.lib()
    for _ in range(3):
        kmeans_step(X, C, sums, counts, workspace=ws, variant=variant)
    torch.cuda.synchronize()
    times = []
    for _ in range(iters):
        lib.hb_l2_flush(_native.ptr(flush), flush.numel(), _native.stream_ptr())
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
        kmeans_step(X, C, sums, counts, workspace=ws, variant=variant)
        e1.record(); torch.cuda.synchronize()
        times.append(e0.elapsed_time(e1))
    ms = sorted(times)[len(times) // 2]
    flops = 2.0 * n * k * d
    out = {"variant": variant, "n": n, "d": d, "k": k, "ms_median": ms, "ms_all": times, "tflops_tf32": flops / ms / 1e9,
           "points_per_s": n / ms * 1e3, "counts_sum": float(counts.sum())}
    # torch baseline: fp32 matmul (TF32 allowed) + argmin + index_add
    torch.backends.cuda.matmul.allow_tf32 = True
    def torch_step():
        cn = (C * C).sum(1)
        best = torch.empty(n, dtype=torch.int64, device="cuda")
        for s in range(0, n, 262144):
            dist = torch.addmm(cn[None, :], X[s:s + 262144], C.t(), alpha=-2.0)
            best[s:s + 262144] = dist.argmin(1)
        sums.zero_().index_add_(0, best, X); counts.zero_().index_add_(0, best, torch.ones(n, device="cuda"))
    for _ in range(2):
        torch_step()
    torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record(); torch_step(); e1.record(); torch.cuda.synchronize()
    out["torch_tf32_ms"] = e0.elapsed_time(e1)
    print(json.dumps(out))
    Path("gpurun_out").mkdir(exist_ok=True)
    with open("gpurun_out/kmeans_kernel_bench.jsonl", "a") as f:
        f.write(json.dumps(out) + "\n")


if __name__ == "__main__":
    main()
