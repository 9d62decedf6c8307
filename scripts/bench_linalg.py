"""Times the K21 (TSQR) and K22 (implicit-ALS step 4) kernels on one GPU and prints one JSON line each.

    python scripts/bench_linalg.py [--rows 50000000] [--d 64] [--ratings 50000000] [--factors 64]

TSQR: X is 50 M x 64 fp32 (12.8 GB, larger than L2), timed with CUDA events after warm-up; the library arm is
torch.linalg.qr(mode="r") (cuSOLVER geqrf) on a 1/10 slice scaled to the full row count, and the Gram + Cholesky
route (TF32 tcgen05 Gram kernel of this repo) is timed alongside as the fast-but-squared-condition alternative.
ALS: synthetic implicit ratings with a heavy-tailed item popularity; one half-iteration over all users."""
import argparse
import json
import time

import numpy as np
import torch

from harp_b200.ops import gram as KG, linalg as KL


def timed(fn, steps, warmup=3):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
    ev[0].record()
    for i in range(steps):
        fn()
        ev[i + 1].record()
    torch.cuda.synchronize()
    ts = sorted(ev[i].elapsed_time(ev[i + 1]) for i in range(steps))
    return ts[len(ts) // 2]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=50_000_000)
    ap.add_argument("--d", type=int, default=64)
    ap.add_argument("--ratings", type=int, default=50_000_000)
    ap.add_argument("--users", type=int, default=2_000_000)
    ap.add_argument("--items", type=int, default=200_000)
    ap.add_argument("--factors", type=int, default=64)
    ap.add_argument("--steps", type=int, default=10)
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    g = torch.Generator(device=dev).manual_seed(0)

    X = torch.randn(a.rows, a.d, device=dev, generator=g)
    ms = timed(lambda: KL.tsqr_r(X), a.steps)
    R = KL.tsqr_r(X).double()
    sub = X[: a.rows // 10]
    ms_lib = timed(lambda: torch.linalg.qr(sub, mode="r"), 3, 1) * 10
    ms_gram = timed(lambda: torch.linalg.cholesky(KG.gram(X).double()), a.steps) if a.d % 32 == 0 else None
    G = KG.gram(X[:1_000_000], force_simt=True).double()
    R1 = KL.tsqr_r(X[:1_000_000]).double()
    err = float((R1.t() @ R1 - G).abs().max() / G.abs().max())
    print(json.dumps({"kernel": "tsqr", "rows": a.rows, "d": a.d, "ms": round(ms, 3),
                      "gbytes_per_s": round(a.rows * a.d * 4 / ms / 1e6, 1),
                      "library_geqrf_ms_scaled": round(ms_lib, 1), "gram_cholesky_ms": ms_gram and round(ms_gram, 3),
                      "rtr_vs_gram_rel_err_1m": err, "r_diag_min": float(R.diag().abs().min())}))
    del X, sub
    torch.cuda.empty_cache()

    rng = np.random.default_rng(0)
    users = torch.from_numpy(rng.integers(0, a.users, a.ratings)).to(dev)
    pop = rng.zipf(1.3, a.ratings) % a.items
    items = torch.from_numpy(pop.astype(np.int64)).to(dev)
    vals = torch.from_numpy(rng.integers(1, 6, a.ratings).astype(np.float32)).to(dev)
    order = torch.argsort(users)
    users, items, vals = users[order], items[order], vals[order]
    ptr = torch.zeros(a.users + 1, dtype=torch.int64, device=dev)
    ptr[1:] = torch.cumsum(torch.bincount(users, minlength=a.users), 0)
    F = KL.als_pad(a.factors)
    Y = torch.zeros(a.items, F, device=dev); Y[:, : a.factors] = torch.rand(a.items, a.factors, device=dev, generator=g) * 0.1
    out = torch.empty(a.users, F, device=dev)
    side = KL.ALSSide(ptr, items, vals)
    ms = timed(lambda: KL.als_solve(side, Y, out, a.factors, 0.1, 40.0), a.steps)
    flops = 2.0 * a.ratings * F * F + a.users * (F ** 3 / 3.0 + 2 * F * F)
    # library arm: the same normal matrices through torch (segment GEMM via index_add) + batched cholesky_solve, on a slice
    nu = min(a.users, 5_000)
    e = int(ptr[nu].item())
    def lib_arm():
        Yu = Y[items[:e]]
        A = torch.zeros(nu, F, F, device=dev)
        A.index_add_(0, users[:e], (Yu * (40.0 * vals[:e])[:, None])[:, :, None] * Yu[:, None, :])
        A += (Y.t() @ Y + 0.1 * torch.eye(F, device=dev))
        b = torch.zeros(nu, F, device=dev).index_add_(0, users[:e], Yu * (1 + 40.0 * vals[:e])[:, None])
        return torch.cholesky_solve(b.unsqueeze(-1), torch.linalg.cholesky(A))
    ms_lib = timed(lib_arm, 3, 1) * (a.users / nu)
    print(json.dumps({"kernel": "als_step4", "users": a.users, "items": a.items, "ratings": a.ratings, "factors": a.factors,
                      "ms": round(ms, 3), "gflops": round(flops / ms / 1e6, 1), "ratings_per_s": round(a.ratings / ms * 1e3),
                      "library_index_add_cholesky_ms_scaled": round(ms_lib, 1), "not_spd": int(side.info.item())}))


if __name__ == "__main__":
    main()
