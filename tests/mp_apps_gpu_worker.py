"""torchrun worker (NCCL, one GPU per rank): the applications whose hot loops got device kernels in round 2 — PageRank
(dense push), subgraph counting (compressed exchange and rotation), WDA-MDS (fused pass), MF-CCD (sweep kernel) — end to
end through their mappers on GPUs, against the same closed-form answers as the CPU worker (tests/mp_apps_worker.py)."""
import math
import os
import sys
import tempfile
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
import torch
import torch.distributed as dist

from harp_b200 import _native
from harp_b200.apps import contrib as AC, graph as AG, mf as AM
from harp_b200.io import formats as F, multifile
from harp_b200.mapper import init_distributed


def run(mapper, files, conf, out=None):
    rank, world = dist.get_rank(), dist.get_world_size()
    mine = multifile.shard_files(sorted(files), rank, world)
    return mapper.run(multifile.records_for(mine), conf, out)


def main():
    rank, world, dev = init_distributed("nccl")
    assert dev.type == "cuda"
    base = sys.argv[1]
    gen = rank == 0
    l0 = _native.launches()

    # ---------------- PageRank: dense rank vector + push kernel ----------------
    n = 12
    g = {0: [1, 2], 1: [2], 2: [0], 3: [2, 0], 4: [], 5: [4, 6], 6: [5], 7: [0], 8: [9], 9: [8, 10], 10: [11], 11: [3]}
    if gen:
        os.makedirs(f"{base}/pr", exist_ok=True)
        for k in range(3):
            with open(f"{base}/pr/part{k}", "w") as f:
                for v in range(k, n, 3):
                    f.write(" ".join(str(x) for x in [v] + g[v]) + "\n")
    dist.barrier()
    m = AC.PageRankMapper(); run(m, [f"{base}/pr/part{k}" for k in range(3)], {"num_urls": n, "num_iterations": 30})
    P = np.zeros((n, n))
    for s, ts in g.items():
        if ts:
            for t in ts: P[t, s] += 1 / len(ts)
        else:
            P[:, s] += 1 / n
    r = np.full(n, 1 / n)
    for _ in range(30): r = 0.85 * P @ r + 0.15 / n
    np.testing.assert_allclose(m.ranks.numpy(), r, rtol=1e-9)

    # ---------------- subgraph counting ----------------
    nv = 14
    rs = np.random.default_rng(3)
    und = {(min(a, b), max(a, b)) for a, b in rs.integers(0, nv, (40, 2)) if a != b}
    adj = {v: sorted({b for a, b in und if a == v} | {a for a, b in und if b == v}) for v in range(nv)}
    if gen:
        os.makedirs(f"{base}/sc", exist_ok=True)
        for k in range(2):
            F.write_adjacency(f"{base}/sc/g{k}", {v: adj[v] for v in range(k, nv, 2)})
        F.write_template(f"{base}/sc/u3.template", 3, [(0, 1), (1, 2)])
        F.write_template(f"{base}/sc/u4.template", 4, [(0, 1), (0, 2), (0, 3)])
    dist.barrier()
    exact_path = sum(len(adj[v]) * (len(adj[v]) - 1) // 2 for v in range(nv))
    exact_star = sum(math.comb(len(adj[v]), 3) for v in range(nv))
    for tmpl, exact in (("u3", exact_path), ("u4", exact_star)):
        res = {}
        for mode in ("exchange", "rotation"):
            m = AG.SubgraphMapper(); run(m, [f"{base}/sc/g0", f"{base}/sc/g1"], {"template_file": f"{base}/sc/{tmpl}.template", "num_iterations": 200, "comm": mode, "seed": 1})
            res[mode] = m.count
            assert abs(m.count - exact) < 0.15 * exact + 1, (tmpl, mode, m.count, exact)
        assert abs(res["exchange"] - res["rotation"]) < 1e-4 * max(1.0, exact)       # same colourings, same DP (fp32 tables)

    # ---------------- WDA-MDS ----------------
    pts = np.random.default_rng(5).standard_normal((30, 3))
    Dm = np.linalg.norm(pts[:, None] - pts[None], axis=2)
    if gen:
        os.makedirs(f"{base}/mds", exist_ok=True)
        for k, idx in enumerate(np.array_split(np.arange(30), 4)):
            F.write_dense_csv(f"{base}/mds/p{k}.csv", np.c_[idx, Dm[idx]])
    dist.barrier()
    m = AG.WDAMDSMapper(); run(m, [f"{base}/mds/p{k}.csv" for k in range(4)], {"n": 30, "target_dim": 3, "max_iter": 60, "alpha": 0.7, "cg_iter": 15})
    assert m.stress < 2e-3, m.stress_history[-3:]
    De = torch.cdist(m.X, m.X).cpu().numpy()
    assert np.abs(De - Dm).max() < 0.1

    # ---------------- MF-CCD through its mapper ----------------
    rs = np.random.default_rng(11)
    nr, nc, rk = 120, 80, 4
    Wt, Ht = rs.standard_normal((nr, rk)) / 2, rs.standard_normal((nc, rk)) / 2
    key = np.unique(rs.integers(0, nr * nc, 4000))
    rr, cc = key // nc, key % nc
    vv = (Wt[rr] * Ht[cc]).sum(1) + 3
    if gen:
        os.makedirs(f"{base}/ccd", exist_ok=True)
        for k, idx in enumerate(np.array_split(np.arange(key.size), 4)):
            with open(f"{base}/ccd/p{k}", "w") as f:
                for i in idx:
                    f.write(f"{rr[i]} {cc[i]} {vv[i]}\n")
    dist.barrier()
    m = AM.CCDCollectiveMapper(); run(m, [f"{base}/ccd/p{k}" for k in range(4)], {"r": 6, "lambda": 0.05, "num_iterations": 6})
    assert m.objective_history[-1] < 0.5 * m.objective_history[0], m.objective_history
    assert _native.launches() > l0 + 100                    # the device kernels did the work
    dist.barrier()
    if rank == 0:
        print("APPS_GPU_OK", world, flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
