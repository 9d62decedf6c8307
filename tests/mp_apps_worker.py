"""torchrun worker (gloo, CPU): every secondary application end to end on tiny generated inputs."""
import itertools
import math
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
GPU = os.environ.get("HARP_APPS_GPU", "0") == "1"          # same worker on GPUs (NCCL): tests/test_symm_gpu.py
if not GPU:
    os.environ["HARP_FORCE_CPU"] = "1"
import numpy as np
import torch
import torch.distributed as dist

from harp_b200.apps import contrib as AC, graph as AG, lda as AL, mf as AM
from harp_b200.io import datagen, formats as F, multifile
from harp_b200.mapper import init_distributed


def run(mapper, files, conf, out=None):
    rank, world = dist.get_rank(), dist.get_world_size()
    mine = multifile.shard_files(sorted(files), rank, world)
    return mapper.run(multifile.records_for(mine), conf, out)


def main():
    rank, world, _ = init_distributed("nccl" if GPU else "gloo")
    base = sys.argv[1]
    gen = rank == 0

    # ---------------- PageRank ----------------
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
    m = AC.PageRankMapper(); run(m, [f"{base}/pr/part{k}" for k in range(3)], {"num_urls": n, "num_iterations": 30, "tables": "dense"})
    np.testing.assert_allclose(m.ranks.numpy(), r, rtol=1e-9)            # dense rank vector + push op == key-value tables

    # ---------------- iterative SVM ----------------
    rng = np.random.default_rng(0)
    if gen:
        os.makedirs(f"{base}/svm", exist_ok=True)
        for k in range(4):
            y = rng.choice([-1.0, 1.0], 60)
            X = rng.standard_normal((60, 5)) + y[:, None] * 1.5
            F.write_libsvm(f"{base}/svm/p{k}", X, y)
    dist.barrier()
    m = AC.IterativeSVMMapper(); run(m, [f"{base}/svm/p{k}" for k in range(4)], {"num_iterations": 3, "num_features": 5})
    assert m.train_accuracy > 0.95 and m.history[-1] > 0, (m.train_accuracy, m.history)

    # ---------------- MLR ----------------
    if gen:
        os.makedirs(f"{base}/mlr", exist_ok=True)
        C = rng.standard_normal((4, 10)) * 3
        for k in range(4):
            y = rng.integers(0, 4, 80)
            F.write_libsvm(f"{base}/mlr/p{k}", C[y] + rng.standard_normal((80, 10)), y)
    dist.barrier()
    m = AC.MLRMapper(); run(m, [f"{base}/mlr/p{k}" for k in range(4)], {"num_topics": 4, "num_terms": 10, "num_iterations": 60, "alpha": 0.5})
    assert m.macro_f1 > 0.9, m.macro_f1

    # ---------------- LDA-CVB (static + dynamic) ----------------
    if gen:
        os.makedirs(f"{base}/cvb", exist_ok=True)
        docs = datagen.synth_corpus(40, 30, 3, 40, seed=1)
        for k, idx in enumerate(np.array_split(np.arange(40), 4)):
            with open(f"{base}/cvb/p{k}", "w") as f:
                for i in idx:
                    ids, cnt = np.unique(docs[i], return_counts=True)
                    f.write(f"doc{i} " + " ".join(f"{a}:{b}" for a, b in zip(ids, cnt)) + "\n")
    dist.barrier()
    lls = []
    for dyn in (False, True):
        m = AC.LDACVBMapper(); run(m, [f"{base}/cvb/p{k}" for k in range(4)], {"num_topics": 3, "num_terms": 30, "num_iterations": 8, "dynamic": dyn})
        assert m.likelihood_history[-1] > m.likelihood_history[0], m.likelihood_history
        assert m.alpha_vector.shape == (3,) and bool((m.alpha_vector > 0).all())
        assert float((m.alpha_vector - 0.1).abs().max()) > 1e-3          # the prior was re-estimated
        lls.append(m.likelihood_history)
    np.testing.assert_allclose(lls[0], lls[1], rtol=1e-9)            # push/pull variant == allreduce variant

    # ---------------- NN + random forest ----------------
    if gen:
        os.makedirs(f"{base}/nn", exist_ok=True)
        C = rng.standard_normal((3, 6)) * 3
        for k in range(5):
            y = rng.integers(0, 3, 90)
            F.write_dense_csv(f"{base}/nn/p{k}.csv", np.c_[C[y] + rng.standard_normal((90, 6)), y])
    dist.barrier()
    m = AC.NNMapper(); run(m, [f"{base}/nn/p{k}.csv" for k in range(4)], {"num_features": 6, "num_classes": 3, "epochs": 15, "sync_iter": 3, "lr": 0.2})
    assert m.train_accuracy > 0.93, m.train_accuracy
    m = AC.RandomForestMapper(); run(m, [f"{base}/nn/p{k}.csv" for k in range(4)], {"num_classes": 3, "num_trees": 4, "test_file": f"{base}/nn/p4.csv"})
    if rank == 0:
        assert m.test_accuracy > 0.9, m.test_accuracy

    # ---------------- subgraph counting: allgather and rotation (SAHAD) modes vs brute force ----------------
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
        for mode in ("allgather", "rotation"):
            m = AG.SubgraphMapper(); run(m, [f"{base}/sc/g0", f"{base}/sc/g1"], {"template_file": f"{base}/sc/{tmpl}.template", "num_iterations": 300, "comm": mode, "seed": 1})
            res[mode] = m.count
            assert abs(m.count - exact) < 0.12 * exact + 1, (tmpl, mode, m.count, exact)
        assert abs(res["allgather"] - res["rotation"]) < (1e-4 if GPU else 1e-6) * max(1.0, exact)       # same colourings, same DP

    # ---------------- WDA-MDS ----------------
    pts = np.random.default_rng(5).standard_normal((30, 3))
    Dm = np.linalg.norm(pts[:, None] - pts[None], axis=2)
    if gen:
        os.makedirs(f"{base}/mds", exist_ok=True)
        for k, idx in enumerate(np.array_split(np.arange(30), 4)):
            F.write_dense_csv(f"{base}/mds/p{k}.csv", np.c_[idx, Dm[idx]])
    dist.barrier()
    m = AG.WDAMDSMapper(); run(m, [f"{base}/mds/p{k}.csv" for k in range(4)], {"n": 30, "target_dim": 3, "max_iter": 60, "alpha": 0.7, "cg_iter": 15})
    assert m.stress < 1e-3, m.stress_history[-3:]
    De = torch.cdist(m.X, m.X).cpu().numpy()
    assert np.abs(De - Dm).max() < 0.1

    # ---------------- K-means with the SGX cost model ----------------
    if gen:
        datagen.generate_kmeans_data(400, 6, 4, f"{base}/sgx/data", seed=0)
        datagen.generate_centroids(4, 6, f"{base}/sgx/cen", seed=1)
    dist.barrier()
    m = AG.SGXSimuKMeansMapper()
    run(m, [f"{base}/sgx/data/data_{k}" for k in range(4)], {"num_centroids": 4, "vector_size": 6, "num_iterations": 3, "centroid_file": f"{base}/sgx/cen", "enclave_total_mb": 0.001})
    assert m.sgx_overhead_ms["init"] > 50 and m.sgx_overhead_ms["comm"] > 0 and m.sgx_overhead_ms["compute"] > 0
    assert all(a >= b - 1e-6 for a, b in zip(m.objective_history, m.objective_history[1:]))

    dist.barrier()
    if rank == 0:
        print("APPS_OK", world, flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
