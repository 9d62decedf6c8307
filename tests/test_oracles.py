"""The product's CPU paths against the independent NumPy oracles of oracle/ (written from the Java sources)."""
import math

import numpy as np
import torch


def test_sgd_cpu_path_matches_the_sequential_oracle():
    """One worker, one block: the host twin of the SGD pass visits the ratings in the panel's stored order; the oracle
    replays exactly that order rating by rating."""
    from oracle import mf as O
    from harp_b200.ops import mf
    rs = np.random.default_rng(0)
    nr, nc, r, nnz = 40, 30, 8, 500
    key = np.unique(rs.integers(0, nr * nc, nnz))
    rows, cols = key // nc, key % nc
    vals = rs.random(key.size) * 4 + 1
    W = rs.random((nr, r)) * 0.3
    H = rs.random((nc, r)) * 0.3
    panel = mf.build_panel(torch.from_numpy(rows), torch.from_numpy(cols), torch.from_numpy(vals), 1, max_run=0, hot_spread=False,
                           balance=False, lpt=False)
    Wt, Ht = torch.from_numpy(W.copy()), torch.from_numpy(H.copy())
    n = mf.sgd_pass(panel, Wt, Ht, 0.01, 0.05)
    assert int(n.item()) == key.size
    order_cols = np.repeat(panel.seg_col.numpy() & 0x7FFFFFFF, np.diff(panel.seg_ptr.numpy()))
    Wo, Ho = W.copy(), H.copy()
    O.sgd_pass(Wo, Ho, panel.r_row.numpy(), order_cols, panel.r_val.numpy().astype(np.float64), 0.01, 0.05)
    np.testing.assert_allclose(Wt.numpy(), Wo, rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(Ht.numpy(), Ho, rtol=1e-5, atol=1e-6)


def test_ccd_engine_matches_the_oracle():
    from oracle import mf as O
    from harp_b200.models.mfccd import MFCCDEngine
    rs = np.random.default_rng(1)
    nr, nc, r, lam = 50, 35, 3, 0.1
    key = np.unique(rs.integers(0, nr * nc, 800))
    rows, cols = key // nc, key % nc
    vals = rs.random(key.size) * 4 + 1
    tr, tc, tv = torch.from_numpy(rows), torch.from_numpy(cols), torch.from_numpy(vals)
    eng = MFCCDEngine(tr, tc, tv, tr, tc, tv, nr, nc, r, lam, seed=5)
    W, H, res = eng.W_loc.numpy().copy(), eng.H_loc.numpy().copy(), vals.copy()
    for _ in range(3):
        assert abs(eng.iterate() - O.ccd_iteration(W, H, res, rows, cols, lam)) < 1e-12
    np.testing.assert_allclose(eng.W_loc.numpy(), W, atol=1e-12)
    np.testing.assert_allclose(eng.H_loc.numpy(), H, atol=1e-12)


def test_lda_engine_likelihood_and_conditional_match_the_oracle():
    """Book-keeping and likelihood of the CPU engine against the oracle's tables rebuilt from the engine's assignments;
    the oracle's sampler, run from the same start, reaches the same likelihood level."""
    from oracle import lda as O
    from harp_b200.io.datagen import synth_corpus
    from harp_b200.models.lda import LDAEngine
    D, V, K, a, b = 60, 50, 5, 0.1, 0.01
    docs_l = synth_corpus(D, V, K, 30, 0)
    d = torch.cat([torch.full((len(x),), i) for i, x in enumerate(docs_l)])
    w = torch.cat([torch.from_numpy(x) for x in docs_l])
    eng = LDAEngine(d, w, D, V, K, a, b, seed=1, blocks=1)

    n_dk = eng.n_dk.numpy().astype(np.float64)
    ids, r = eng.home_word_rows()
    n_wk = np.zeros((V, K))
    n_wk[ids.numpy()] = r.numpy()
    n_k = n_wk.sum(0)
    assert n_dk.sum() == d.numel() == n_wk.sum()
    ll_engine = eng.log_likelihood()
    assert abs(ll_engine - O.log_likelihood(n_dk, n_wk, n_k, a, b)) < 1e-6 * abs(ll_engine)
    # same start, oracle sampler vs engine sampler: both climb to the same level (sampling noise apart)
    rng = np.random.default_rng(0)
    zs = np.empty(d.numel(), dtype=np.int64)
    # rebuild an assignment vector consistent with the engine's document-topic table
    pos = 0
    dn, wn = d.numpy(), w.numpy()
    order = np.argsort(dn, kind="stable")
    dn, wn = dn[order], wn[order]
    for doc in range(D):
        ks = np.repeat(np.arange(K), n_dk[doc].astype(int))
        zs[pos:pos + ks.size] = rng.permutation(ks); pos += ks.size
    o_dk = np.zeros((D, K)); o_wk = np.zeros((V, K)); o_k = np.zeros(K)
    np.add.at(o_dk, (dn, zs), 1); np.add.at(o_wk, (wn, zs), 1); np.add.at(o_k, zs, 1)
    for _ in range(15):
        eng.iterate()
        O.gibbs_pass(dn, wn, zs, o_dk, o_wk, o_k, a, b, rng)
    ll_e, ll_o = eng.log_likelihood(), O.log_likelihood(o_dk, o_wk, o_k, a, b)
    assert ll_e > ll_engine and abs(ll_e - ll_o) < 0.05 * abs(ll_o), (ll_engine, ll_e, ll_o)


def test_covariance_and_linear_regression_match_the_oracle():
    from oracle import stats as O
    from harp_b200.daal import linalg as L
    rs = np.random.default_rng(2)
    X = rs.standard_normal((400, 7)) * np.arange(1, 8) + 3
    Y = X @ rs.standard_normal((7, 2)) + 0.5 + 0.01 * rs.standard_normal((400, 2))
    cov, mean = O.covariance_from_partials(np.array_split(X, 3))
    res = L.covariance(torch.from_numpy(X))
    np.testing.assert_allclose(res.covariance.numpy(), cov, rtol=1e-9, atol=1e-9)
    np.testing.assert_allclose(res.mean.numpy(), mean, rtol=1e-12)
    beta = O.linear_regression_normal_equations(np.array_split(X, 3), np.array_split(Y, 3))
    model = L.linear_regression_normeq(torch.from_numpy(X), torch.from_numpy(Y))
    got = model.predict(torch.from_numpy(X)).numpy()
    np.testing.assert_allclose(got, np.c_[np.ones(400), X] @ beta, rtol=1e-8, atol=1e-8)
