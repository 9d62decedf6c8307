"""Harp-DAAL equivalents vs NumPy / SciPy / scikit-learn references (CPU; the same code runs on CUDA tensors)."""
import math

import numpy as np
import pytest
import torch

from harp_b200.daal import classify as C, linalg as L, models as M


@pytest.fixture(scope="module")
def X():
    g = torch.Generator().manual_seed(0)
    A = torch.randn(6, 6, generator=g, dtype=torch.float64)
    return torch.randn(500, 6, generator=g, dtype=torch.float64) @ A + torch.arange(6, dtype=torch.float64)


def test_moments_cov_pca(X):
    m = L.low_order_moments(X)
    np.testing.assert_allclose(m.mean.numpy(), X.numpy().mean(0), rtol=1e-10)
    np.testing.assert_allclose(m.variance.numpy(), X.numpy().var(0, ddof=1), rtol=1e-9)
    np.testing.assert_allclose(m.minimum.numpy(), X.numpy().min(0)); np.testing.assert_allclose(m.maximum.numpy(), X.numpy().max(0))
    c = L.covariance(X)
    np.testing.assert_allclose(c.covariance.numpy(), np.cov(X.numpy().T), rtol=1e-8, atol=1e-10)
    np.testing.assert_allclose(c.correlation.numpy(), np.corrcoef(X.numpy().T), rtol=1e-8, atol=1e-10)
    p = L.pca_correlation(X)
    w = np.sort(np.linalg.eigvalsh(np.corrcoef(X.numpy().T)))[::-1]
    np.testing.assert_allclose(p.eigenvalues.numpy(), w, rtol=1e-8)
    p2 = L.pca_svd(X)
    np.testing.assert_allclose(p2.eigenvalues.numpy(), w, rtol=1e-7)
    np.testing.assert_allclose(np.abs(p2.eigenvectors.numpy()), np.abs(p.eigenvectors.numpy()), atol=1e-6)
    # sparse (CSR) input takes the same path
    c2 = L.covariance(X.to_sparse_csr())
    np.testing.assert_allclose(c2.covariance.numpy(), c.covariance.numpy(), rtol=1e-10)


def test_qr_svd_cholesky_pivoted(X):
    Q, R = L.qr_distributed(X)
    np.testing.assert_allclose((Q @ R).numpy(), X.numpy(), atol=1e-9)
    np.testing.assert_allclose((Q.t() @ Q).numpy(), np.eye(6), atol=1e-9)
    U, S, Vh = L.svd_distributed(X)
    np.testing.assert_allclose(S.numpy(), np.linalg.svd(X.numpy(), compute_uv=False), rtol=1e-9)
    np.testing.assert_allclose((U * S @ Vh).numpy(), X.numpy(), atol=1e-8)
    A = X.t() @ X
    Lc = L.cholesky(A)
    np.testing.assert_allclose((Lc @ Lc.t()).numpy(), A.numpy(), rtol=1e-10)
    Qp, Rp, perm = L.pivoted_qr(X[:20])
    np.testing.assert_allclose((Qp @ Rp).numpy(), X[:20][:, perm].numpy(), atol=1e-8)
    d = np.abs(np.diag(Rp.numpy()))
    assert np.all(d[:-1] >= d[1:] - 1e-9)                          # pivoting orders the diagonal


def test_linear_and_ridge_regression(X):
    g = torch.Generator().manual_seed(1)
    beta = torch.randn(6, 2, generator=g, dtype=torch.float64)
    Y = X @ beta + torch.tensor([1.5, -2.0]) + 0.01 * torch.randn(500, 2, generator=g, dtype=torch.float64)
    A = np.c_[np.ones(500), X.numpy()]
    ref = np.linalg.lstsq(A, Y.numpy(), rcond=None)[0].T
    for fn in (L.linear_regression_normeq, L.linear_regression_qr):
        m = fn(X, Y)
        np.testing.assert_allclose(m.beta.numpy(), ref, rtol=1e-6, atol=1e-6)
    r = L.ridge_regression(X, Y, ridge=3.0)
    P = np.eye(7); P[0, 0] = 0
    ref_r = np.linalg.solve(A.T @ A + 3.0 * P, A.T @ Y.numpy()).T
    np.testing.assert_allclose(r.beta.numpy(), ref_r, rtol=1e-8)
    q = L.quality_metrics_regression(Y[:, 0], L.linear_regression_normeq(X, Y).predict(X)[:, 0])
    assert q["r2"] > 0.999


def test_normalisation_quantiles_sorting_outliers(X):
    z = L.zscore(X)
    np.testing.assert_allclose(z.mean(0).numpy(), 0, atol=1e-12); np.testing.assert_allclose(z.std(0).numpy(), 1, rtol=1e-12)
    mm = L.minmax(X, -1, 1)
    assert float(mm.min()) == -1 and abs(float(mm.max()) - 1) < 1e-12
    q = L.quantiles(X, (0.5,))
    xs = np.sort(X.numpy(), 0)
    np.testing.assert_allclose(q[:, 0].numpy(), xs[math.ceil(0.5 * 500) - 1])
    assert torch.equal(L.sort_columns(X), torch.sort(X, 0).values)
    Xo = X.clone(); Xo[0] += 100
    assert L.outlier_univariate(Xo)[0].min() == 0 and L.outlier_univariate(Xo)[1:].mean() > 0.95
    assert L.outlier_multivariate(Xo, 6.0)[0, 0] == 0
    assert L.outlier_bacon(Xo)[0, 0] == 0 and L.outlier_bacon(Xo)[1:].mean() > 0.9


def _blobs(n=300, seed=0, sep=4.0, classes=3, d=4):
    g = torch.Generator().manual_seed(seed)
    y = torch.randint(0, classes, (n,), generator=g)
    centers = torch.randn(classes, d, generator=g, dtype=torch.float64) * sep
    return centers[y] + torch.randn(n, d, generator=g, dtype=torch.float64), y


def test_naive_bayes_svm_knn_kernels():
    Xb, y = _blobs()
    Xc = (Xb - Xb.min()).round()                                      # count-like non-negative features
    nb = C.naive_bayes_train(Xc, y, 3)
    from sklearn.naive_bayes import MultinomialNB
    sk = MultinomialNB(alpha=1.0).fit(Xc.numpy(), y.numpy())
    np.testing.assert_allclose(nb.log_theta.numpy(), sk.feature_log_prob_, rtol=1e-10)
    assert (nb.predict(Xc) == torch.from_numpy(sk.predict(Xc.numpy()))).all()
    svm = C.svm_train_multiclass(Xb, y, C=1.0, kernel="linear")
    assert float((svm.predict(Xb) == y).double().mean()) > 0.97
    svm_r = C.svm_train_multiclass(Xb, y, C=1.0, kernel="rbf", sigma=2.0)
    assert float((svm_r.predict(Xb) == y).double().mean()) > 0.97
    from sklearn.svm import SVC
    yb = torch.where(y == 0, 1.0, -1.0)
    mine = C.svm_train_binary(Xb, yb, C=0.5, kernel="linear", tol=1e-4)
    ref = SVC(C=0.5, kernel="linear", tol=1e-5).fit(Xb.numpy(), yb.numpy())
    np.testing.assert_allclose(mine.decision(Xb).numpy(), ref.decision_function(Xb.numpy()), atol=2e-2)
    assert float((C.knn_predict(Xb, y, Xb, 5) == y).double().mean()) > 0.97
    from sklearn.metrics.pairwise import rbf_kernel
    np.testing.assert_allclose(C.kernel_rbf(Xb[:20], Xb[:30], 1.5).numpy(), rbf_kernel(Xb[:20].numpy(), Xb[:30].numpy(), gamma=1 / (2 * 1.5 ** 2)), rtol=1e-9)
    np.testing.assert_allclose(C.kernel_linear(Xb[:5].to_sparse_csr(), Xb[:7], 2.0, 1.0).numpy(), 2 * Xb[:5].numpy() @ Xb[:7].numpy().T + 1, rtol=1e-12)


def test_trees_forest_boosting():
    Xb, y = _blobs(400, 1, classes=2)
    ypm = torch.where(y == 0, -1.0, 1.0).double()
    s = C.stump_train(Xb, ypm)
    assert float((s.predict(Xb) == ypm).double().mean()) > 0.8
    t = C.DecisionTree("classification", max_depth=6).fit(Xb, y)
    assert float((t.predict(Xb) == y).double().mean()) > 0.97
    g = torch.Generator().manual_seed(2)
    yr = Xb[:, 0] * 2 + (Xb[:, 1] > 0).double() * 5 + 0.01 * torch.randn(400, generator=g, dtype=torch.float64)
    tr = C.DecisionTree("regression", max_depth=8).fit(Xb, yr)
    assert float(((tr.predict(Xb) - yr) ** 2).mean()) < 0.2 * float(yr.var())
    f = C.DecisionForest("classification", n_trees=8, max_depth=6).fit(Xb, y)
    assert float((f.predict(Xb) == y).double().mean()) > 0.97 and len(f.traverse()) == 8
    fr = C.DecisionForest("regression", n_trees=8, max_depth=8).fit(Xb, yr)
    assert float(((fr.predict(Xb) - yr) ** 2).mean()) < 0.3 * float(yr.var())
    for train in (C.adaboost_train, C.brownboost_train):
        m = train(Xb, ypm)
        assert float((m.predict(Xb) == ypm).double().mean()) > 0.93, train.__name__
    X3, y3 = _blobs(300, 3, classes=3)
    lb = C.logitboost_train(X3, y3, 3, 20)
    assert float((lb.predict(X3) == y3).double().mean()) > 0.95
    qm = L.quality_metrics_classification(y3, lb.predict(X3), 3)
    assert qm["accuracy"] > 0.95 and 0 <= qm["error_rate"] < 0.05


def test_kmeans_em_als_apriori_solvers_nn():
    Xb, y = _blobs(600, 4, classes=3, d=5)
    Cn, hist = M.kmeans_daal(Xb, Xb[:3].clone(), 10)
    assert all(a >= b - 1e-9 for a, b in zip(hist, hist[1:]))
    g = M.em_gmm(Xb, 3, 60, seed=1)
    assert abs(float(g.weights.sum()) - 1) < 1e-9 and g.log_likelihood > -600 * 12
    # implicit ALS: reconstruct preferences of a low-rank interaction matrix
    gen = torch.Generator().manual_seed(0)
    U, I = 40, 30
    P = (torch.rand(U, 3, generator=gen) @ torch.rand(3, I, generator=gen) > 0.9)
    r, c = P.nonzero(as_tuple=True)
    Xf, Yf = M.als_implicit(r, c, torch.ones(r.numel()), U, I, 8, 10, 0.1, 20.0)
    pred = Xf @ Yf.t()
    assert float(pred[P].mean()) > float(pred[~P].mean()) + 0.3
    sup, rules = M.apriori([[1, 2, 3], [1, 2], [1, 3], [2, 3], [1, 2, 3], [1, 2]], 0.5, 0.7)
    assert abs(sup[frozenset([1, 2])] - 4 / 6) < 1e-12 and ((2,), (1,), 0.8) in [(a, b, round(cf, 6)) for a, b, cf in rules]
    gen = torch.Generator().manual_seed(5)
    Xs = torch.randn(400, 3, generator=gen, dtype=torch.float64)
    ys = Xs @ torch.tensor([1.0, -2.0, 0.5], dtype=torch.float64) + 3.0
    obj = M.MSEObjective(Xs, ys)
    want = torch.tensor([3.0, 1.0, -2.0, 0.5], dtype=torch.float64)
    np.testing.assert_allclose(M.solver_lbfgs(obj, 50).numpy(), want.numpy(), atol=1e-6)
    np.testing.assert_allclose(M.solver_sgd(obj, 3000, 0.05, 16).numpy(), want.numpy(), atol=5e-2)
    np.testing.assert_allclose(M.solver_sgd(obj, 3000, 0.02, 16, momentum=0.9).numpy(), want.numpy(), atol=5e-2)
    np.testing.assert_allclose(M.solver_adagrad(obj, 3000, 0.5, 16).numpy(), want.numpy(), atol=5e-2)
    net = M.MLP([5, 16, 3], seed=0, dtype=torch.float64)
    h = net.train(Xb, y, epochs=30, batch_size=32, lr=0.1)
    assert h[-1] < h[0] and float((net.predict(Xb) == y).double().mean()) > 0.95


def test_initialize_create_job_runs_a_mapper(tmp_path):
    """``Initialize.createJob(...).waitForCompletion(true)`` — the launcher idiom of every Harp-DAAL app."""
    import os
    import numpy as np
    from harp_b200.daal.initialize import DataGenerator, Initialize
    from harp_b200.mapper import CollectiveMapper

    os.environ["HARP_FORCE_CPU"] = "1"
    files = DataGenerator.generate_dense_data_multi(40, 3, 2, str(tmp_path / "in"))
    assert len(files) == 2

    class CountRows(CollectiveMapper):
        def map_collective(self, reader, context):
            self.rows = 0
            while reader.next_key_value():
                self.rows += len(np.loadtxt(reader.get_current_value(), delimiter=",", ndmin=2))

    init = Initialize({}, ["1", "2", "1000", "3", str(tmp_path / "in"), str(tmp_path / "work"), "extra"])
    init.loadDistributedLibs()
    job = init.createJob("count", None, CountRows)
    assert init.getSysArgNum() == 6 and init.app_args() == ["extra"] and job.getConfiguration()["num_iterations"] == 3
    assert job.waitForCompletion(True) and job.isSuccessful() and job.mapper.rows == 40


def test_kmeans_init_methods_and_csr_kmeans_without_densifying():
    """daal_kmeans/allreducecsr: kmeans.init (deterministic / random / plusPlus, dense and CSR give the same centroids)
    and Lloyd passes over CSR rows through sparse products equal the dense run."""
    g = torch.Generator().manual_seed(0)
    Xd = torch.rand(300, 20, generator=g, dtype=torch.float64)
    Xd[Xd < 0.7] = 0
    Xs = Xd.to_sparse_csr()
    for method in ("deterministic", "random", "plusPlus"):
        Cd, Cs = M.kmeans_init(Xd, 6, method, seed=3), M.kmeans_init(Xs, 6, method, seed=3)
        assert Cd.shape == (6, 20) and torch.allclose(Cd, Cs), method
    assert torch.equal(M.kmeans_init(Xd, 4, "deterministic"), Xd[:4])
    C0 = M.kmeans_init(Xs, 6, "plusPlus", seed=1)
    Ca, ha = M.kmeans_daal(Xd, C0, 8)
    Cb, hb = M.kmeans_daal(Xs, C0, 8)
    np.testing.assert_allclose(ha, hb, rtol=1e-9)
    assert torch.allclose(Ca.double(), Cb, atol=1e-9)
    assert all(a >= b - 1e-9 for a, b in zip(hb, hb[1:]))
