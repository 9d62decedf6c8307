"""torchrun worker (GPU, NCCL): the rotation engines with the NVLink symmetric workspace."""
import math
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
import torch.distributed as dist


def main():
    rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); lr = int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(lr); dev = torch.device("cuda", lr)
    dist.init_process_group("nccl", device_id=dev)
    from harp_b200.parallel.symm import SymmetricWorkspace
    from harp_b200.models.mfsgd import MFSGDEngine
    from harp_b200.models.lda import LDAEngine
    ws = SymmetricWorkspace(dev, 512 << 20)
    in_use0 = ws.bytes_in_use()
    # ---- MF-SGD ----
    g = torch.Generator(device=dev).manual_seed(1 + rank)
    nr, nc, nnz, rk = 20000, 3000, 600000, 64
    Wt = torch.randn(nr, 8, device=dev, generator=g) / math.sqrt(8)
    Ht = torch.randn(nc, 8, device=dev, generator=torch.Generator(device=dev).manual_seed(77)) / math.sqrt(8)
    r = torch.randint(0, nr, (nnz,), device=dev, generator=g); c = torch.randint(0, nc, (nnz,), device=dev, generator=g)
    v = (Wt[r] * Ht[c]).sum(1) * 2 + 3
    # a world of one (the driver's single-GPU box) runs the same rotation protocol as a ring of one rank
    one = dict(fused=True, loopback=True) if world == 1 else {}
    eng = MFSGDEngine(r, c, v, nr, nc, rk, 0.01, 0.02, symm=ws, seed=5, **one)
    assert eng.fused and eng.S == 2, "the default path must be the whole-epoch kernel with two slices in flight"
    e0 = eng.rmse(r, c, v)
    from harp_b200 import _native
    l0 = _native.launches()
    for it in range(12):
        n = eng.epoch()
        assert int(n.item()) == nnz, (int(n.item()), nnz)
    eng.synchronize()
    assert _native.launches() - l0 == 12, "one launch per epoch expected"
    e1 = eng.rmse(r, c, v)
    assert e1 < 0.5 * e0, (e0, e1)
    assert all(r % world == 0 for r in eng.rot)
    ids = torch.cat([x[0] for x in eng.home_H().values()])
    assert bool((ids % world == rank).all())                       # shards are home again
    # ---- LDA ----
    D, V, K, L = 3000, 2000, 64, 50
    docs = torch.arange(D, device=dev).repeat_interleave(L)
    words = (torch.rand(D * L, device=dev, generator=g) ** 2 * V).long() % V
    lda = LDAEngine(docs, words, D, V, K, 0.1, 0.01, symm=ws, seed=2, **one)
    assert int(lda.topic_sums.sum()) == D * L * world
    ll0 = lda.log_likelihood()
    for it in range(8):
        n = lda.iterate()
        assert int(n.item()) == D * L
    lda.synchronize()
    ll1 = lda.log_likelihood()
    assert ll1 > ll0, (ll0, ll1)
    assert int(lda.topic_sums.sum()) == D * L * world and int(lda.n_dk.sum()) == D * L
    tot = torch.zeros(1, dtype=torch.int64, device=dev)
    for k in range(lda.S):
        tot += lda.Wbuf[k][lda.cur[k]].sum()
    dist.all_reduce(tot)
    assert int(tot.item()) == D * L * world
    # randomised rotation orders (RotationUtil.createRotationOrder), a single slice, other cluster sizes,
    # and the per-panel launch path as the A/B arm
    import random
    from harp_b200.dymoro import create_rotation_order
    orders = create_rotation_order(random.Random(7), 8, world)          # identical on every rank
    for kw in (dict(orders=orders), dict(num_slices=1, cluster=2), dict(fused=False), dict(orders=orders, fused=False)):
        eng2 = MFSGDEngine(r, c, v, nr, nc, rk, 0.01, 0.02, symm=ws, seed=5, **kw)
        f0 = eng2.rmse(r, c, v)
        for it in range(6):
            assert int(eng2.epoch().item()) == nnz
        eng2.synchronize()
        f1 = eng2.rmse(r, c, v)
        assert f1 < 0.7 * f0, (kw, f0, f1)
        ids = torch.cat([x[0] for x in eng2.home_H().values()])
        home = eng2.sched[0].shard_at[rank]
        assert bool((ids % world == home).all()), kw
    for kw in (dict(num_slices=1), dict(fused=False), dict(cluster=2)):
        lda2 = LDAEngine(docs, words, D, V, K, 0.1, 0.01, symm=ws, seed=2, **kw)
        m0 = lda2.log_likelihood()
        for it in range(4):
            assert int(lda2.iterate().item()) == D * L
        lda2.synchronize()
        assert lda2.log_likelihood() > m0 and int(lda2.topic_sums.sum()) == D * L * world, kw
    # ---- MF-CCD++: sharded factors, rank-one vectors all-gathered over NVLink; equals the single-process engine ----
    from harp_b200.models.mfccd import MFCCDEngine
    gc = torch.Generator(device="cpu").manual_seed(9)
    cr, cc_, cnnz = 3000, 500, 80000
    key = torch.unique(torch.randint(0, cr * cc_, (cnnz,), generator=gc))
    crow, ccol = (key // cc_).to(dev), (key % cc_).to(dev)
    cval = (torch.rand(key.numel(), generator=gc) * 4 + 1).to(dev)
    mr_, mc_ = crow % world == rank, ccol % world == rank
    ccd = MFCCDEngine(crow[mr_], ccol[mr_], cval[mr_], crow[mc_], ccol[mc_], cval[mc_], cr, cc_, 5, 0.05, symm=ws, seed=4)
    hist = [ccd.iterate() for _ in range(3)]
    assert hist[2] < hist[0], hist
    Wc, Hc = ccd.full_factors()
    pred = (Wc[:, crow] * Hc[:, ccol]).sum(0)
    rm = float(((cval - pred) ** 2).mean().sqrt())
    assert abs(rm - hist[2]) < 1e-3 * hist[2], (rm, hist)          # the local residuals ARE the model's residuals
    # ---- implicit ALS, 4-step protocol on the K22 kernel: owners solve their rows, factors meet by allreduce ----
    from harp_b200 import collective as C
    from harp_b200.daal import models as DM
    ga = torch.Generator(device="cpu").manual_seed(21)
    U_, I_, f_ = 400, 250, 10
    keya = torch.unique(torch.randint(0, U_ * I_, (12000,), generator=ga))
    ar, ac = keya // I_, keya % I_
    av = torch.ones(keya.numel(), dtype=torch.float64)
    Xc, Yc = DM.als_implicit(ar, ac, av, U_, I_, f_, 4, 0.1, 20.0)                      # single process, CPU, fp64
    mine_ = torch.arange(keya.numel()) % world == rank                                  # every rank reads a slice of the ratings
    comm_a = C.Comm(device=dev, symm=ws)
    Xg, Yg = DM.als_implicit(ar[mine_].to(dev), ac[mine_].to(dev), av[mine_].to(dev), U_, I_, f_, 4, 0.1, 20.0, comm=comm_a)
    assert float(((Xc @ Yc.t()) - (Xg @ Yg.t()).cpu()).abs().max()) < 5e-3
    # engines hand their slab blocks back: a long-lived process can build engines again and again
    if world > 1:
        from harp_b200.models.kmeans import KMeansEngine
        before = ws.bytes_in_use()
        for rep_ in range(3):
            e_s = MFSGDEngine(r, c, v, nr, nc, rk, 0.01, 0.02, symm=ws, seed=5)
            e_l = LDAEngine(docs, words, D, V, K, 0.1, 0.01, symm=ws, seed=2)
            e_k = KMeansEngine(torch.rand(20000, 64, device=dev), 300, comm="fused", symm=ws)
            e_s.epoch(); e_l.iterate(); e_k.step()
            e_k.close(); e_l.close(); e_s.close()
            assert ws.bytes_in_use() == before, (rep_, ws.bytes_in_use(), before)
        ccd.close()
    dist.barrier()
    if rank == 0:
        print(f"ENGINES_OK world={world} sgd_rmse {e0:.4f}->{e1:.4f} lda_ll {ll0:.4e}->{ll1:.4e}", flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
