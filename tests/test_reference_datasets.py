"""The DAAL-style algorithms on the sample data that ships with the reference (``datasets/daal_*``), the inputs of
its ``ml/daal/test_scripts/harp-daal-*.sh`` acceptance scripts.  Those scripts only check that the job completes;
here the results are compared with NumPy or with the ground-truth files that sit next to the data.  Skipped when
the reference tree is not mounted (nothing is copied into this repository)."""
import glob
import os

import numpy as np
import pytest
import torch

from harp_b200.daal import classify, linalg, models

BASE = os.environ.get("HARP_REFERENCE_DATASETS", "/root/reference/datasets")
pytestmark = pytest.mark.skipif(not os.path.isdir(os.path.join(BASE, "daal_cov")),
                                reason="reference data sets not available")


def _csv(path, ncols=None):
    with open(path) as f:
        first = f.readline()
    n = ncols or len([x for x in first.strip().split(",") if x])
    return np.loadtxt(path, delimiter=",", usecols=range(n), ndmin=2)


def _dir(rel, ncols=None):
    return [_csv(f, ncols) for f in sorted(glob.glob(os.path.join(BASE, rel, "*.csv")))]


def test_covariance_moments_pca_on_daal_cov_dense():
    parts = _dir("daal_cov/daal_cov_dense")
    assert len(parts) == 4
    X = np.concatenate(parts)
    Xt = torch.from_numpy(X)
    res = linalg.covariance(Xt)
    np.testing.assert_allclose(res.covariance.numpy(), np.cov(X, rowvar=False), rtol=1e-9, atol=1e-9)
    np.testing.assert_allclose(res.mean.numpy(), X.mean(0), rtol=1e-12, atol=1e-12)
    np.testing.assert_allclose(res.correlation.numpy(), np.corrcoef(X, rowvar=False), rtol=1e-9, atol=1e-9)
    mom = linalg.low_order_moments(Xt)
    np.testing.assert_allclose(mom.variance.numpy(), X.var(0, ddof=1), rtol=1e-9)
    np.testing.assert_allclose(mom.minimum.numpy(), X.min(0)); np.testing.assert_allclose(mom.maximum.numpy(), X.max(0))
    pca = linalg.pca_correlation(Xt)
    want = np.sort(np.linalg.eigvalsh(np.corrcoef(X, rowvar=False)))[::-1]
    np.testing.assert_allclose(pca.eigenvalues.numpy(), want, rtol=1e-8, atol=1e-10)


def test_linear_and_ridge_regression_against_the_ground_truth_file():
    tr = np.concatenate(_dir("daal_reg/train", 12))
    te = _csv(os.path.join(BASE, "daal_reg/test/linear_regression_test.csv"), 12)
    gt = _csv(os.path.join(BASE, "daal_reg/groundTruth/linear_regression_test_groundTruth.csv"), 2)
    X, Y = torch.from_numpy(tr[:, :10]), torch.from_numpy(tr[:, 10:])
    beta_np, *_ = np.linalg.lstsq(np.hstack([np.ones((len(tr), 1)), tr[:, :10]]), tr[:, 10:], rcond=None)
    for fit in (linalg.linear_regression_normeq, linalg.linear_regression_qr, linalg.ridge_regression):
        m = fit(X, Y)
        pred = m.predict(torch.from_numpy(te[:, :10])).numpy()
        assert np.abs(pred - gt).max() < 1e-2 * np.abs(gt).max()           # the file holds noisy targets
        if fit is not linalg.ridge_regression:
            np.testing.assert_allclose(pred, np.hstack([np.ones((len(te), 1)), te[:, :10]]) @ beta_np, rtol=1e-7, atol=1e-6)


def test_qr_and_svd_on_daal_qr_dense():
    parts = _dir("daal_qr/daal_qr_dense")
    X = torch.from_numpy(np.concatenate(parts))
    qr = linalg.qr_distributed(X)
    Q, R = (qr.Q, qr.R) if hasattr(qr, "Q") else qr
    torch.testing.assert_close(Q @ R, X, rtol=1e-9, atol=1e-9)
    torch.testing.assert_close(Q.t() @ Q, torch.eye(Q.shape[1], dtype=Q.dtype), rtol=1e-9, atol=1e-9)
    sv = linalg.svd_distributed(X)
    S = sv.S if hasattr(sv, "S") else sv[1]
    np.testing.assert_allclose(np.sort(S.numpy())[::-1], np.linalg.svd(X.numpy(), compute_uv=False), rtol=1e-9)


def test_naive_bayes_and_neural_network_reach_the_ground_truth_labels():
    tr = np.concatenate(_dir("daal_naive/densedistri/train", 21))
    te = _csv(glob.glob(os.path.join(BASE, "daal_naive/densedistri/test/*.csv"))[0], 21)
    gt = _csv(glob.glob(os.path.join(BASE, "daal_naive/densedistri/groundTruth/*.csv"))[0], 1)[:, 0]
    nb = classify.naive_bayes_train(torch.from_numpy(tr[:, :20]), torch.from_numpy(tr[:, 20]).long(), 20)
    assert float((nb.predict(torch.from_numpy(te[:, :20])).numpy() == gt).mean()) > 0.98
    tr = np.concatenate(_dir("daal_nn/train", 21))
    te = _csv(os.path.join(BASE, "daal_nn/test/neural_network_test.csv"), 20)
    gt = _csv(os.path.join(BASE, "daal_nn/groundTruth/neural_network_test_ground_truth.csv"), 1)[:, 0]
    net = models.MLP([20, 20, 40, 2], seed=1)
    net.train(torch.tensor(tr[:, :20], dtype=torch.float32), torch.tensor(tr[:, 20], dtype=torch.long), epochs=20,
              batch_size=50, lr=0.05)
    assert float((net.predict(torch.tensor(te, dtype=torch.float32)).numpy() == gt).mean()) > 0.97


def test_kmeans_on_daal_kmeans_dense_is_monotone_and_matches_the_oracle():
    from oracle import kmeans as oracle
    X = np.concatenate(_dir("daal_kmeans/densedistri"))
    init = X[:10].copy()
    got = models.kmeans_daal(torch.from_numpy(X), torch.from_numpy(init), 8, None, "allreduce")
    C = got.centroids if hasattr(got, "centroids") else got[0]
    want, hist = oracle.run(X, init, 8)
    np.testing.assert_allclose(C.numpy(), want, rtol=1e-9, atol=1e-9)
    assert all(a >= b - 1e-6 for a, b in zip(hist, hist[1:]))


def test_multiclass_svm_and_boosting_on_the_reference_samples():
    tr = np.concatenate(_dir("daal_svm/multidense/train"))
    te = np.concatenate(_dir("daal_svm/multidense/test"))
    m = classify.svm_train_multiclass(torch.from_numpy(tr[:, :-1]), torch.from_numpy(tr[:, -1]).long())
    acc = float((m.predict(torch.from_numpy(te[:, :-1])).numpy() == te[:, -1]).mean())
    assert acc > 0.85, acc                                   # scikit-learn SVC(kernel="linear", C=1): 0.873
    for name, fit in (("adaboost", classify.adaboost_train), ("brownboost", classify.brownboost_train)):
        tr = np.concatenate([_csv(f) for f in sorted(glob.glob(os.path.join(BASE, f"daal_{name}/train/*")))])
        te = _csv(glob.glob(os.path.join(BASE, f"daal_{name}/test/*.csv"))[0])
        pred = fit(torch.from_numpy(tr[:, :-1]), torch.from_numpy(tr[:, -1])).predict(torch.from_numpy(te[:, :-1]))
        assert float((pred.numpy().reshape(-1) == te[:, -1]).mean()) > 0.97, name


BATCH_RUNS = [
    ("svm", "daal_svm/multidense/train", ["20", "daal_svm/multidense/test/svm_multi_class_test_dense.csv"], "test_accuracy", 0.85),
    ("adaboost", "daal_adaboost/train", ["20", "daal_adaboost/test/adaboost_test.csv"], "test_accuracy", 0.97),
    ("brownboost", "daal_brownboost/train", ["20", "daal_brownboost/test/brownboost_test.csv"], "test_accuracy", 0.97),
    ("logitboost", "daal_logitboost/train", ["20", "5", "daal_logitboost/test/logitboost_test.csv"], "test_accuracy", 0.97),
    ("stump", "daal_stump/train", ["20", "daal_stump/test/stump_test.csv"], "test_accuracy", 0.97),
    ("dtree_cls_traverse", "daal_dtree/train", ["5"], "nodes", None),
    ("dforest_reg", "daal_dforest/regdensebatch/train", ["13", "daal_dforest/regdensebatch/test/df_regression_test.csv"], "test_rmse", None),
    ("em", "daal_em/batchdense/train", ["3", "2"], "weights", None),
    ("als_batch", "daal_als/implicitdense/train", ["5"], None, None),
    ("ar", "daal_ar/batchdense/train", ["0.01", "0.6"], "n_itemsets", None),
    ("cholesky", "daal_cholesky/train", ["5"], None, None),
    ("pivoted_qr", "daal_pivoted_qr/train", [], None, None),
    ("sorting", "daal_sorting/train", [], None, None),
    ("quantile", "daal_quantile/train", [], None, None),
    ("zscore", "daal_normalization/train", [], None, None),
    ("minmax", "daal_normalization/train", [], None, None),
    ("outlier_uni", "daal_outlier/train", [], None, None),
    ("outlier_multi", "daal_outlier/train", [], None, None),
    ("outlier_bacon", "daal_outlier/train", [], None, None),
    ("kernel_lin", "daal_kernelfunc/densebatch", [], None, None),
    ("kernel_rbf", "daal_kernelfunc/densebatch", [], None, None),
    ("mse", "daal_optimization_solvers/mse/train", ["3"], "gradient_norm", None),
    ("lbfgs", "daal_optimization_solvers/lbfgs/train", ["10"], None, None),
    ("sgd_moment", "daal_optimization_solvers/mse/train", ["3"], None, None),
    ("adagrad", "daal_optimization_solvers/mse/train", ["3"], None, None),
    ("quality_linreg", "daal_quality_metrics/linregmetrics/train", ["10"], "rms", None),
    ("quality_svm", "daal_quality_metrics/svmmultimetrics/train", ["20", "daal_quality_metrics/svmmultimetrics/test"], None, None),
    ("knn", "daal_knn/batchdense/train", ["5", "5", "daal_knn/batchdense/test"], "test_accuracy", 0.93),
    ("dforest_cls_traverse", "daal_dforest/clsdensebatch/train", ["3", "daal_dforest/clsdensebatch/test"], "test_accuracy", 0.7),
    ("dtree_reg", "daal_dtree/train", ["5"], None, None),
    ("cov_csr", "daal_cov/daal_cov_csr", [], "covariance", None),
    ("pca_csr", "daal_pca/daal_pca_csr", [], "eigenvalues", None),
    ("mom_csr", "daal_mom/daal_mom_csr", [], "variance", None),
    ("kmeans_csr", "daal_kmeans/csrdistri", ["10"], None, None),
    ("svm_csr", "daal_svm/multicsr/train", ["daal_svm/multicsr/trainTruth/svm_multi_class_train_labels.csv",
                                            "daal_svm/multicsr/test/svm_multi_class_test_csr.csv",
                                            "daal_svm/multicsr/testTruth/svm_multi_class_test_labels.csv"], "test_accuracy", 0.8),
    ("kernel_lin_csr", "daal_kernelfunc/csrbatch", [], None, None),
    ("kernel_rbf_csr", "daal_kernelfunc/csrbatch", [], None, None),
]


@pytest.mark.parametrize("algo,indir,app,key,minimum", BATCH_RUNS, ids=[r[0] for r in BATCH_RUNS])
def test_every_batch_launcher_runs_on_its_reference_sample(tmp_path, monkeypatch, capsys, algo, indir, app, key, minimum):
    """One CLI run per remaining ``*DaalLauncher`` of the reference (``ml/daal/test_scripts/harp-daal-*.sh`` only
    check that the job completes; accuracies are asserted where the sample has labels)."""
    from harp_b200.launch import daal as launcher
    monkeypatch.setenv("HARP_FORCE_CPU", "1")
    app = [os.path.join(BASE, x) if x.startswith("daal_") else x for x in app]
    work = tmp_path / "work"
    assert launcher.main([algo, "1", "2", "1000", "5", os.path.join(BASE, indir), str(work)] + app) == 0
    out = capsys.readouterr().out
    assert f"Harp-DAAL {algo} finished" in out and (work / f"{algo}.out").exists()
    if key is not None:
        assert f"{algo}.{key}" in out or key in (work / f"{algo}.out").read_text()
    if minimum is not None:
        line = [l for l in out.splitlines() if l.startswith(f"{algo}.{key}:")][0]
        assert float(line.split(":")[1]) > minimum, line
