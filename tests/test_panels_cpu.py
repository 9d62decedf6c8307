"""CPU checks of the rating/token panel builder and of the rotation engines' slice layouts
(the block schedule the SGD / LDA kernels walk; reference: dymoro/Scheduler.java:95-237, RowColSplit)."""
import math

import pytest
import torch

from harp_b200.ops import mf


def _ratings(nr=500, nc=120, nnz=20000, seed=0, skew=True):
    g = torch.Generator().manual_seed(seed)
    if skew:
        r = (torch.rand(nnz, generator=g) ** 2 * nr).long() % nr
        c = (torch.rand(nnz, generator=g) ** 2 * nc).long() % nc
    else:
        r = torch.randint(0, nr, (nnz,), generator=g)
        c = torch.randint(0, nc, (nnz,), generator=g)
    v = torch.rand(nnz, generator=g) * 4 + 1
    return r, c, v


@pytest.mark.parametrize("B,C,balance,lpt", [(1, None, True, True), (8, None, True, True), (8, None, False, False),
                                             (8, 16, True, True), (5, 7, True, False), (16, None, True, True)])
def test_panel_holds_every_rating_once_in_a_consistent_block(B, C, balance, lpt):
    r, c, v = _ratings()
    p = mf.build_panel(r, c, v, B, max_run=8, balance=balance, C=C, lpt=lpt)
    Cn = p.ncolblocks
    assert Cn == (C or B) and p.blk_seg_ptr.numel() == B * Cn + 1
    assert p.nnz == r.numel() and int(p.blk_seg_ptr[-1]) == p.seg_col.numel() == p.seg_ptr.numel() - 1
    seg_len = (p.seg_ptr[1:] - p.seg_ptr[:-1]).long()
    assert int(seg_len.min()) >= 1 and int(seg_len.max()) <= 8                     # max_run respected
    seg_colv = (p.seg_col & 0x7FFFFFFF).long()
    cols_of_rating = torch.repeat_interleave(seg_colv, seg_len)
    # the multiset of (row, col, value) triples is preserved
    key_in = (r * 1000 + c).double() * 10 + v.double()
    key_out = (p.r_row.long() * 1000 + cols_of_rating).double() * 10 + p.r_val.double()
    torch.testing.assert_close(torch.sort(key_in).values, torch.sort(key_out).values, rtol=0, atol=1e-4)
    # block structure: a row lives in exactly one row-block, a non-shared column in exactly one column-block
    seg_blk = torch.searchsorted(p.blk_seg_ptr.long(), torch.arange(p.seg_col.numel()), right=True) - 1
    rb_of_rating = torch.repeat_interleave(seg_blk // Cn, seg_len)
    row_to_rb = {}
    for row, rb in zip(p.r_row.tolist(), rb_of_rating.tolist()):
        assert row_to_rb.setdefault(row, rb) == rb
    shared = p.seg_col < 0
    col_to_cb = {}
    for col, cb, sh in zip(seg_colv.tolist(), (seg_blk % Cn).tolist(), shared.tolist()):
        if not sh:
            assert col_to_cb.setdefault(col, cb) == cb
    if lpt:      # longest run first inside every block
        for b in range(B * Cn):
            s0, s1 = int(p.blk_seg_ptr[b]), int(p.blk_seg_ptr[b + 1])
            ln = seg_len[s0:s1]
            assert bool((ln[:-1] >= ln[1:]).all())


def test_serpentine_assignment_balances_skewed_rows():
    r, c, v = _ratings(nr=2000, nc=300, nnz=200000, seed=1)
    B = 16
    for balance, bound in ((True, 1.05), (False, 10.0)):
        p = mf.build_panel(r, c, v, B, balance=balance)
        per_block = (p.seg_ptr[p.blk_seg_ptr[1:].long()] - p.seg_ptr[p.blk_seg_ptr[:-1].long()]).view(B, B).sum(1).double()
        ratio = float(per_block.max() / per_block.mean())
        assert ratio < bound, (balance, ratio)
        if balance:
            balanced = ratio
        else:
            assert ratio > balanced          # the id-modulo layout is measurably worse on skewed data


def test_cpu_sgd_pass_matches_plain_sequential_sgd():
    """B = 1, one run per column in column order: the pass is plain sequential SGD in (column, input) order."""
    g = torch.Generator().manual_seed(2)
    nr, nc, rank = 60, 9, 8
    r = torch.randperm(nr, generator=g)
    c = torch.randint(0, nc, (nr,), generator=g)
    v = torch.rand(nr, generator=g) * 4 + 1
    W = torch.rand(nr, rank, generator=g, dtype=torch.float64) * 0.3
    H = torch.rand(nc, rank, generator=g, dtype=torch.float64) * 0.3
    W1, H1 = W.clone(), H.clone()
    p = mf.build_panel(r, c, v, 1, max_run=0, lpt=False)
    n = mf.sgd_pass(p, W1, H1, 0.05, 0.01)
    assert int(n) == nr
    W2, H2 = W.clone(), H.clone()
    for col in range(nc):
        for k in torch.nonzero(c == col).flatten().tolist():
            w, h = W2[r[k]].clone(), H2[col].clone()
            e = float(w @ h) - float(v[k])
            W2[r[k]] = w - 0.05 * (e * h + 0.01 * w)
            H2[col] = h - 0.05 * (e * w + 0.01 * h)
    torch.testing.assert_close(W1, W2, rtol=1e-6, atol=1e-7)     # ratings are float32, the factors float64
    torch.testing.assert_close(H1, H2, rtol=1e-6, atol=1e-7)


@pytest.mark.parametrize("S", [1, 2])
def test_engines_single_process_both_slice_layouts(S):
    from harp_b200.models.lda import LDAEngine
    from harp_b200.models.mfsgd import MFSGDEngine
    g = torch.Generator().manual_seed(4)
    nr, nc, nnz, rk = 80, 30, 1500, 6
    Wt, Ht = torch.randn(nr, 3, generator=g), torch.randn(nc, 3, generator=g)
    r = torch.randint(0, nr, (nnz,), generator=g); c = torch.randint(0, nc, (nnz,), generator=g)
    v = (Wt[r] * Ht[c]).sum(1) + 3
    eng = MFSGDEngine(r, c, v, nr, nc, rk, 0.02, 0.01, blocks=3, seed=1, num_slices=S)
    assert eng.S == S and len(eng.panels[0]) == S and sum(p.nnz for p in eng.panels[0]) == nnz
    e0 = eng.rmse(r, c, v)
    for _ in range(15):
        assert int(eng.epoch().item()) == nnz
    assert eng.rmse(r, c, v) < 0.6 * e0
    ids = torch.cat([x[0] for x in eng.home_H().values()])
    assert sorted(ids.tolist()) == list(range(nc))
    D, V, K, L = 40, 50, 4, 12
    docs = torch.arange(D).repeat_interleave(L)
    words = torch.randint(0, V, (D * L,), generator=g)
    lda = LDAEngine(docs, words, D, V, K, 0.1, 0.01, blocks=2, seed=3, num_slices=S)
    assert lda.S == S and int(lda.topic_sums.sum()) == D * L
    ll0 = lda.log_likelihood()
    for _ in range(6):
        assert int(lda.iterate().item()) == D * L
    assert lda.log_likelihood() > ll0 and int(lda.n_dk.sum()) == D * L


@pytest.mark.parametrize("dtype,B,C", [(torch.float64, 1, None), (torch.float64, 6, None), (torch.float32, 5, 9)])
def test_host_cpp_sgd_pass_equals_the_python_loop(dtype, B, C):
    """csrc/host/mf_host.cpp::hh_sgd_pass against the Python restatement it replaces on the CPU path."""
    from harp_b200 import _native
    if not _native.have_host_lib():
        pytest.skip("host library not built")
    r, c, v = _ratings(nr=200, nc=50, nnz=3000, seed=7)
    p = mf.build_panel(r, c, v, B, C=C)
    g = torch.Generator().manual_seed(0)
    W = (torch.rand(200, 8, generator=g, dtype=torch.float64) * 0.3).to(dtype)
    H = (torch.rand(50, 8, generator=g, dtype=torch.float64) * 0.3).to(dtype)
    W1, H1, W2, H2 = W.clone(), H.clone(), W.clone(), H.clone()
    n1 = mf._cpu_sgd_pass(p, W1, H1, 0.02, 0.05)
    n2 = mf._host_sgd_pass(p, W2, H2, 0.02, 0.05)
    assert n1 == n2 == 3000
    tol = dict(rtol=1e-12, atol=1e-13) if dtype == torch.float64 else dict(rtol=2e-4, atol=2e-5)
    torch.testing.assert_close(W2, W1, **tol)          # fp32 tables: the C++ pass accumulates in double
    torch.testing.assert_close(H2, H1, **tol)


@pytest.mark.parametrize("dk", [torch.int32, torch.int16])
def test_host_cpp_lda_pass_keeps_the_tables_consistent(dk):
    from harp_b200 import _native
    from harp_b200.ops import lda as L
    if not _native.have_host_lib():
        pytest.skip("host library not built")
    g = torch.Generator().manual_seed(3)
    D, V, K, n = 60, 40, 7, 2500
    docs = torch.randint(0, D, (n,), generator=g); words = torch.randint(0, V, (n,), generator=g)
    z = torch.randint(0, K, (n,), generator=g)
    panel = L.build_token_panel(docs, words, z, 3)
    n_dk = torch.zeros(D, K, dtype=torch.int32); n_wk = torch.zeros(V, K, dtype=torch.int32)
    L.count_into(panel.tok_doc, panel.tok_z, K, n_dk); L.count_into(panel.tok_word, panel.tok_z, K, n_wk)
    n_dk = n_dk.to(dk)
    commons = (1.0 / (n_wk.sum(0).double() + 0.01 * V)).float()
    before = panel.tok_z.clone()
    for it in range(3):
        assert L._host_pass(panel, n_dk, n_wk, commons, 0.1, 0.01, 11 + it) == n
    assert not torch.equal(before, panel.tok_z)
    nd = torch.zeros(D, K, dtype=torch.int32); nw = torch.zeros(V, K, dtype=torch.int32)
    L.count_into(panel.tok_doc, panel.tok_z, K, nd); L.count_into(panel.tok_word, panel.tok_z, K, nw)
    assert torch.equal(nd, n_dk.int()) and torch.equal(nw, n_wk) and int(n_dk.min()) >= 0


def test_ccd_engine_equals_the_reference_loops():
    """MFCCDEngine (two residual orders, pending rank-one term applied on the fly, one pass per side and feature)
    against a direct restatement of CCDMPTask.doColCCD / doRowCCD on ONE residual array, inner = 2."""
    import math
    import torch
    from harp_b200.models.mfccd import MFCCDEngine
    g = torch.Generator().manual_seed(0)
    nr, nc, r, lam = 60, 40, 4, 0.1
    key = torch.unique(torch.randint(0, nr * nc, (900,), generator=g))
    rows, cols = key // nc, key % nc
    vals = torch.rand(key.numel(), generator=g, dtype=torch.float64) * 4 + 1
    eng = MFCCDEngine(rows, cols, vals, rows, cols, vals, nr, nc, r, lam, seed=3, inner=2)
    W, H, res = eng.W_loc.clone(), eng.H_loc.clone(), vals.clone()
    for it in range(3):
        for f in range(r):
            for _ in range(2):
                for j in range(nc):
                    m = cols == j
                    if m.any():
                        wt, ht = W[f, rows[m]], H[f, j].clone()
                        s = ((res[m] + ht * wt) * wt).sum() / (lam * float(m.sum()) + (wt * wt).sum())
                        res[m] -= (s - ht) * wt
                        H[f, j] = s
                for i in range(nr):
                    m = rows == i
                    if m.any():
                        ht, wt = H[f, cols[m]], W[f, i].clone()
                        z = ((res[m] + wt * ht) * ht).sum() / (lam * float(m.sum()) + (ht * ht).sum())
                        res[m] -= (z - wt) * ht
                        W[f, i] = z
        assert abs(eng.iterate() - math.sqrt(float((res ** 2).mean()))) < 1e-12
    assert torch.allclose(eng.W_loc, W, atol=1e-12) and torch.allclose(eng.H_loc, H, atol=1e-12)
