"""CI-sized launch of every kernel family, one or two iterations each: the program `scripts/sanitize.sh` puts under
compute-sanitizer (memcheck / racecheck / synccheck / initcheck).  Shapes are tiny on purpose: the tools slow a kernel
down by one to two orders of magnitude.  Every section also checks its result, so a plain run is a smoke test too.

    PYTHONPATH=. python scripts/sanitize_targets.py [section ...]      # default: all sections
"""
import math
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
import torch


def sec_kmeans():
    from harp_b200.models.kmeans import KMeansEngine
    from harp_b200.ops.kmeans import kmeans_step
    g = torch.Generator(device="cuda").manual_seed(0)
    X = torch.rand(3000, 64, device="cuda", generator=g)
    C = torch.rand(100, 64, device="cuda", generator=g)
    ref = torch.cdist(X.double(), C.double()).argmin(1)
    for variant, simt in (("auto", False), ("single", False), ("pair_noext", False), ("auto", True)):
        res = kmeans_step(X, C, want_assign=True, want_obj=True, force_simt=simt, variant=variant)
        assert float((res.assign.long() != ref).double().mean()) < 0.05
        assert float(res.counts.sum()) == 3000
    for prec in ("tf32", "fp16x3", "bf16x3"):
        eng = KMeansEngine(X, 100, precision=prec)
        eng.set_centroids(C)
        for _ in range(2):
            eng.step()
        torch.cuda.synchronize()
        assert bool(torch.isfinite(eng.C).all())


def sec_gram():
    from harp_b200.ops import gram as G
    g = torch.Generator(device="cuda").manual_seed(1)
    X = torch.randn(5000, 64, device="cuda", generator=g) + 3
    ref = X.double().t() @ X.double()
    assert float((G.gram(X).double() - ref).abs().max() / ref.abs().max()) < 5e-3
    X2 = torch.randn(3000, 20, device="cuda", generator=g)
    assert float((G.gram(X2).double() - X2.double().t() @ X2.double()).abs().max()) < 1e-2
    m = G.moments(X)
    assert float((m.sum.double() - X.double().sum(0)).abs().max()) < 1e-2
    Y = torch.randn(5000, 2, device="cuda", generator=g)
    out = G.xty(X, Y)
    assert float((out[1:] - Y.double().t() @ X.double()).abs().max()) < 1e-6 * 5000
    cg = G.centered_gram(X, X.mean(0))
    Xc = X.double() - X.double().mean(0)
    assert float((cg.double() - Xc.t() @ Xc).abs().max() / 5000) < 5e-3


def _ratings(nr, nc, nnz, seed):
    g = torch.Generator().manual_seed(seed)
    Wt = torch.randn(nr, 8, generator=g) / math.sqrt(8)
    Ht = torch.randn(nc, 8, generator=g) / math.sqrt(8)
    r = torch.randint(0, nr, (nnz,), generator=g)
    c = torch.randint(0, nc, (nnz,), generator=g)
    v = (Wt[r] * Ht[c]).sum(1) * 2 + 3
    return r.cuda(), c.cuda(), v.float().cuda()


def sec_mf():
    from harp_b200.models.mfsgd import MFSGDEngine
    from harp_b200.ops import mf
    r, c, v = _ratings(4000, 800, 80000, 2)
    for kw in (dict(fused=False), dict(fused=True, cluster=1), dict(fused=True, cluster=4, loopback=True),
               dict(fused=True, cluster=8, loopback=True)):
        eng = MFSGDEngine(r, c, v, 4000, 800, 64, 0.01, 0.02, seed=5, **kw)
        e0 = eng.rmse(r, c, v)
        for _ in range(2):
            assert int(eng.epoch().item()) == r.numel()
        eng.synchronize()
        assert eng.rmse(r, c, v) < e0
    # rank 200 takes the non-quad cell
    eng = MFSGDEngine(r, c, v, 4000, 800, 200, 0.005, 0.02, seed=5, fused=True, cluster=2, loopback=True)
    assert int(eng.epoch().item()) == r.numel()
    eng.synchronize()
    # CCD rank-one update
    g = torch.Generator().manual_seed(0)
    rr = torch.sort(torch.randint(0, 300, (5000,), generator=g)).values
    cc = torch.randint(0, 200, (5000,), generator=g)
    ptr = torch.zeros(301, dtype=torch.int32); ptr[1:] = torch.cumsum(torch.bincount(rr, minlength=300), 0)
    resid = torch.randn(5000, generator=g).cuda(); Wf = torch.randn(300, generator=g).cuda()
    mf.ccd_update(ptr.cuda(), cc.int().cuda(), resid, None, Wf, torch.randn(200, generator=g).cuda(), 0.1)
    torch.cuda.synchronize()
    assert bool(torch.isfinite(Wf).all())


def sec_lda():
    from harp_b200.io.datagen import synth_corpus
    from harp_b200.models.lda import LDAEngine
    docs = synth_corpus(300, 200, 8, 50, 0)
    d = torch.cat([torch.full((len(x),), i) for i, x in enumerate(docs)]).cuda()
    w = torch.cat([torch.from_numpy(x) for x in docs]).cuda()
    for kw in (dict(sampler="dense", fused=False), dict(sampler="dense", fused=True, cluster=1),
               dict(sampler="dense", fused=True, cluster=4, loopback=True),
               dict(sampler="tokenlist", cluster=1), dict(sampler="tokenlist", cluster=4, loopback=True)):
        for K in (16, 128):
            eng = LDAEngine(d, w, 300, 200, K, 0.1, 0.01, seed=1, **kw)
            l0 = eng.log_likelihood()
            for _ in range(2):
                assert int(eng.iterate().item()) == d.numel()
            eng.synchronize()
            assert eng.log_likelihood() > l0 - abs(l0) * 0.02


def sec_linalg():
    from harp_b200.ops import linalg as KL
    g = torch.Generator(device="cuda").manual_seed(3)
    for n, d in ((3000, 64), (2000, 32), (5000, 16), (1500, 20)):
        X = torch.randn(n, d, device="cuda", generator=g)
        R = KL.tsqr_r(X).double()
        G = X.double().t() @ X.double()
        assert float((R.t() @ R - G).abs().max() / G.abs().max()) < 2e-5
    for f in (10, 32, 64):
        M = torch.randn(50, f, f + 3, device="cuda", generator=g)
        A = M @ M.transpose(1, 2) + 0.5 * torch.eye(f, device="cuda")
        B = torch.randn(50, f, device="cuda", generator=g)
        Xs, L, bad = KL.potrs_batched(A, B, want_factor=True)
        assert bad == 0 and float((torch.einsum("bij,bj->bi", A, Xs) - B).abs().max()) < 1e-2
    rng = np.random.default_rng(0)
    key = np.unique(rng.integers(0, 300 * 200, 6000))
    r, c = torch.from_numpy(key // 200).cuda(), torch.from_numpy(key % 200).cuda()
    ptr = torch.zeros(301, dtype=torch.int64, device="cuda"); ptr[1:] = torch.cumsum(torch.bincount(r, minlength=300), 0)
    for f in (10, 32, 64):
        F = KL.als_pad(f)
        Y = torch.zeros(200, F, device="cuda"); Y[:, :f] = torch.rand(200, f, device="cuda", generator=g)
        out = torch.empty(300, F, device="cuda")
        side = KL.ALSSide(ptr, c, torch.ones(key.size, device="cuda"))
        KL.als_solve(side, Y, out, f, 0.1, 20.0)
        assert int(side.info.item()) == 0 and bool(torch.isfinite(out).all())


def sec_subgraph():
    from harp_b200.apps.graph import SubTemplate, colour_coding_pass
    g = torch.Generator().manual_seed(2)
    n, deg, k = 3000, 5, 6
    src = torch.arange(n).repeat_interleave(deg)
    dst = torch.randint(0, n, (n * deg,), generator=g)
    key = torch.unique(torch.cat([src, dst]) * n + torch.cat([dst, src]))
    a, b = key // n, key % n
    a, b = a[a != b], b[a != b]
    ptr = torch.zeros(n + 1, dtype=torch.int64); ptr[1:] = torch.cumsum(torch.bincount(a, minlength=n), 0)
    idx = b.to(torch.int32)
    edges = [(0, 1), (0, 2), (2, 3), (2, 4), (4, 5)]
    adj_t = {i: [] for i in range(k)}
    for x, y in edges:
        adj_t[x].append(y); adj_t[y].append(x)
    tmpl = SubTemplate(0, frozenset(range(k)), adj_t)
    colours = torch.randint(0, k, (n,), generator=g)
    want = colour_coding_pass(ptr, idx, colours, tmpl, k)[:, 0]
    got = colour_coding_pass(ptr.cuda(), idx.cuda(), colours.cuda(), tmpl, k)[:, 0]
    assert float((got.double().cpu() - want).abs().max()) <= 1e-4 * float(want.abs().max()) + 1e-3


def sec_coll():
    """World of one: the flag protocols and row kernels against the rank's own slab."""
    import os
    import torch.distributed as dist
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1"); os.environ.setdefault("MASTER_PORT", "29541")
    os.environ.setdefault("RANK", "0"); os.environ.setdefault("WORLD_SIZE", "1"); os.environ.setdefault("LOCAL_RANK", "0")
    dev = torch.device("cuda", 0)
    dist.init_process_group("nccl", device_id=dev)
    from harp_b200.parallel.symm import SymmetricWorkspace
    ws = SymmetricWorkspace(dev, 8 << 20)
    t = ws.alloc(5000, torch.float32)
    src = torch.rand(5000, device=dev)
    for op in ("sum", "max"):
        t.copy_(src); ws.allreduce_(t, op); torch.cuda.synchronize()
        assert torch.equal(t, src)
    t.copy_(src); ws.reduce_scatter_(t); ws.allgather_(t); ws.broadcast_(t, root=0); ws.reduce_(t, root=0)
    b = ws.alloc(5000, torch.float32); ws.rotate_(b, t, 0); ws.barrier(); torch.cuda.synchronize()
    assert torch.equal(b, src)
    ws.check()
    dist.destroy_process_group()


SECTIONS = {"kmeans": sec_kmeans, "gram": sec_gram, "mf": sec_mf, "lda": sec_lda, "linalg": sec_linalg, "subgraph": sec_subgraph, "coll": sec_coll}

if __name__ == "__main__":
    names = sys.argv[1:] or list(SECTIONS)
    torch.cuda.set_device(0)
    for nme in names:
        SECTIONS[nme]()
        torch.cuda.synchronize()
        print(f"section {nme}: ok", flush=True)
    print("SANITIZE_TARGETS_OK")
