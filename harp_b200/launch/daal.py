"""Harp-DAAL launchers: ``python -m harp_b200.launch.daal <algorithm> <num_mappers> <threads> <mem_MB>
<iterations> <inputDir> <workDir> [app args]`` — the six standard arguments of ``Initialize.loadSysArgs``
followed by the per-algorithm ones of the reference launchers:

  cov | pca | pca_svd | mom        <fileDim>                    (daal_cov, daal_pca, daal_mom)
  linreg | linreg_qr | ridge       <featureDim> <nDependent>    (daal_linreg, daal_ridgereg)
  naive                            <featureDim> <nClasses>      (daal_naive; last column = label)
  kmeans                           <featureDim> <nCentroids> [comm]   (daal_kmeans)
  qr | svd                         <fileDim>                    (daal_qr, daal_svd)
  nn                               <featureDim> <nClasses> <batch> <lr>   (daal_nn)

Batch algorithms (one mapper in the reference: the ``*Batch`` launchers) gather all input files on every worker and
compute on the master's copy; labels are the last column of the training files, ``[testFile]`` is a CSV of the same
width used for the reported accuracy / error:

  svm                              <featureDim> [testFile]          (daal_svm MultiClassDenseBatch)
  adaboost | brownboost | stump    <featureDim> [testFile]          (labels -1 / +1)
  logitboost                       <featureDim> <nClasses> [testFile]
  dtree_cls | dtree_reg | dforest_cls | dforest_reg   <featureDim> [testFile]   (+ ``_traverse``: node counts)
  knn                              <featureDim> <k> <testFile>
  em                               <fileDim> <nComponents>
  als | als_batch                  <nFactors>                       (dense implicit-feedback matrix CSV)
  ar                               <minSupport> <minConfidence>     (``transaction,item`` lines)
  cholesky | pivoted_qr | sorting | quantile | zscore | minmax      <fileDim>
  outlier_uni | outlier_multi | outlier_bacon                       <fileDim>
  kernel_lin | kernel_rbf          <fileDim>
  mse | sgd | sgd_mini | sgd_moment | lbfgs | adagrad               <featureDim>   (daal_optimization_solvers)
  quality_svm | quality_linreg     <featureDim> <testFile>          (daal_quality_metrics)

CSR inputs (DAAL's three-line CSR text files, one block of rows per file — the reference's ``csr*`` launchers):

  cov_csr | pca_csr | mom_csr                                       (daal_cov/csrdistri, daal_pca/corcsrdistr, daal_mom/csrdistri)
  kmeans_csr                       <nCentroids>                     (daal_kmeans/allreducecsr)
  svm_csr                          <trainLabels> [testCsr testLabels]   (daal_svm/MultiClassCSRBatch)
  kernel_lin_csr | kernel_rbf_csr                                   (daal_kernel_func/{Lin,Rbf}CSRBatch)

Input = comma-separated dense CSV files dealt to workers like ``MultiFileInputFormat``; results are
printed by rank 0 and written under ``<workDir>``.
"""
from __future__ import annotations

import os
import sys

import torch

from ..daal import classify, linalg, models
from ..daal.datasource import HarpDAALDataSource
from ..daal.initialize import Initialize, Service
from ..io import multifile
from ..mapper import CollectiveMapper, Context, KeyValReader
from .common import check_world, run_mapper, setup_logging, split_overrides


class DaalMapper(CollectiveMapper):
    def setup(self, context: Context) -> None:
        self.c = context.get_configuration()

    def map_collective(self, reader: KeyValReader, context: Context) -> None:
        files = []
        while reader.next_key_value():
            files.append(reader.get_current_value())
        c, algo, a = self.c, self.c["algorithm"], self.c["app_args"]
        ds = HarpDAALDataSource(c["num_threads"], c, self.device)
        comm = self.comm
        self.result = None
        if algo in ("cov", "pca", "pca_svd", "mom", "qr", "svd"):
            X = ds.create_dense_numeric_table(files, int(a[0]) if a else None)
            self.result = {"cov": linalg.covariance, "pca": linalg.pca_correlation, "pca_svd": linalg.pca_svd,
                           "mom": linalg.low_order_moments, "qr": linalg.qr_distributed, "svd": linalg.svd_distributed}[algo](X, comm)
        elif algo in ("linreg", "linreg_qr", "ridge"):
            X, Y = ds.create_dense_numeric_table_split(files, int(a[0]), int(a[1]))
            fn = {"linreg": linalg.linear_regression_normeq, "linreg_qr": linalg.linear_regression_qr,
                  "ridge": linalg.ridge_regression}[algo]
            self.result = fn(X, Y, comm)
        elif algo == "naive":
            X, Y = ds.create_dense_numeric_table_split(files, int(a[0]), 1)
            self.result = classify.naive_bayes_train(X, Y[:, 0].long(), int(a[1]), comm)
        elif algo == "kmeans":
            X = ds.create_dense_numeric_table(files, int(a[0]))
            k = int(a[1])
            g = torch.Generator().manual_seed(0)
            init = torch.rand(k, int(a[0]), generator=g, dtype=torch.float64).to(self.device) * 10
            self.result = models.kmeans_daal(X, init, c["num_iterations"], comm, a[2] if len(a) > 2 else "allreduce")
        elif algo == "nn":
            X, Y = ds.create_dense_numeric_table_split(files, int(a[0]), 1)
            net = models.MLP([int(a[0]), 32, int(a[1])], seed=0, device=self.device, dtype=X.dtype)
            hist = net.train(X, Y[:, 0].long(), c["num_iterations"], int(a[2]) if len(a) > 2 else 64,
                             float(a[3]) if len(a) > 3 else 0.1, comm)
            self.result = (net, hist)
        elif algo in ("cov_csr", "pca_csr", "mom_csr", "kmeans_csr"):
            X = ds.load_csr_numeric_table(files)                      # this worker's rows, sparse CSR
            if algo == "kmeans_csr":
                k = int(a[0])
                Xd = X.to_dense().double()
                g = torch.Generator().manual_seed(0)
                init = torch.rand(k, Xd.shape[1], generator=g, dtype=torch.float64).to(self.device) * float(Xd.abs().max() or 1.0)
                self.result = models.kmeans_daal(Xd, init, c["num_iterations"], comm, "allreduce")
            else:
                self.result = {"cov_csr": linalg.covariance, "pca_csr": linalg.pca_correlation,
                               "mom_csr": linalg.low_order_moments}[algo](X, comm)
        elif algo in ("svm_csr", "kernel_lin_csr", "kernel_rbf_csr"):
            every = self._all_files(files)
            X = ds.load_csr_numeric_table(every)
            if algo == "svm_csr":
                y = ds.create_dense_numeric_table([a[0]], 1)[:, 0].long()
                m = classify.svm_train_multiclass(X.to_dense(), y)
                out = {"n_pairs": len(m.pairs)}
                if len(a) > 2:
                    Xt = ds.load_csr_numeric_table([a[1]], X.shape[1])
                    yt = ds.create_dense_numeric_table([a[2]], 1)[:, 0].long()
                    out["test_accuracy"] = float((m.predict(Xt.to_dense()) == yt).double().mean())
                self.result = out
            else:
                self.result = (classify.kernel_linear if algo == "kernel_lin_csr" else classify.kernel_rbf)(X)
        elif algo in BATCH:
            self.result = self.run_batch(algo, a, files, ds)
        else:
            raise SystemExit(f"unknown algorithm {algo!r}")

    # ------------------------------------------------------------------------------------------------------
    def _all_files(self, files):
        """Batch algorithms see the whole input on every worker (the reference runs them with one mapper)."""
        every = self.comm.all_gather_object(sorted(files)) if not self.comm.is_the_only_worker() else [sorted(files)]
        return sorted(f for fs in every for f in fs)

    def run_batch(self, algo, a, files, ds):
        files = self._all_files(files)
        dev = self.device

        def table(ncols=None, paths=None):
            if paths:      # a test "file" may be a directory of part files
                paths = [q for p in paths for q in (multifile.list_input_files(p) if os.path.isdir(p) else [p])]
            return ds.create_dense_numeric_table(paths or files, ncols)

        def xy(d, paths=None, long_labels=True):
            T = table(d + 1, paths)
            y = T[:, d]
            return T[:, :d], (y.long() if long_labels else y)

        def test_path(i):
            return a[i] if len(a) > i else None

        def report(pred, truth):
            return {"test_accuracy": float((pred.reshape(-1).to(truth.dtype) == truth).double().mean()), "n_test": int(truth.numel())}

        if algo == "svm":
            d = int(a[0]); X, y = xy(d)
            m = classify.svm_train_multiclass(X, y)
            out = {"n_pairs": len(m.pairs), "n_support": sum(int(p[2].sv.shape[0]) for p in m.pairs)}
            if test_path(1):
                Xt, yt = xy(d, [test_path(1)]); out.update(report(m.predict(Xt), yt))
            return out
        if algo in ("adaboost", "brownboost", "stump", "logitboost"):
            d = int(a[0])
            if algo == "logitboost":
                X, y = xy(d); m = classify.logitboost_train(X, y, int(a[1])); ti = 2
            else:
                X, y = xy(d, long_labels=False)
                m = {"adaboost": classify.adaboost_train, "brownboost": classify.brownboost_train,
                     "stump": classify.stump_train}[algo](X, y); ti = 1
            out = {"model": type(m).__name__}
            if test_path(ti):
                Xt, yt = xy(d, [test_path(ti)], long_labels=(algo == "logitboost")); out.update(report(m.predict(Xt), yt))
            return out
        if algo.startswith(("dtree_", "dforest_")):
            kind, task = algo.split("_")[0], ("classification" if "_cls" in algo else "regression")
            d = int(a[0]); X, y = xy(d, long_labels=(task == "classification"))
            m = (classify.DecisionTree(task, max_depth=12) if kind == "dtree" else classify.DecisionForest(task, n_trees=10, max_depth=12))
            m.fit(X, y)
            out = {"task": task}
            if algo.endswith("_traverse"):
                out["nodes"] = m.traverse() if kind == "dforest" else classify.DecisionForest.traverse(type("F", (), {"trees": [m]})())
            if test_path(1):
                Xt, yt = xy(d, [test_path(1)], long_labels=(task == "classification"))
                p = m.predict(Xt)
                out.update(report(p, yt) if task == "classification" else {"test_rmse": float(((p - yt) ** 2).mean().sqrt())})
            return out
        if algo == "knn":
            d, k = int(a[0]), int(a[1]); X, y = xy(d); Xt, yt = xy(d, [a[2]])
            return report(classify.knn_predict(X, y, Xt, k), yt)
        if algo == "em":
            g = models.em_gmm(table(int(a[0])), int(a[1]))
            return {"weights": g.weights.tolist(), "means": g.means.tolist()}
        if algo in ("als", "als_batch"):
            R = table()
            r, c = torch.nonzero(R, as_tuple=True)
            U, V = models.als_implicit(r, c, R[r, c], R.shape[0], R.shape[1], int(a[0]) if a else 10, self.c["num_iterations"])[:2]
            return {"users_factors": U, "items_factors": V}
        if algo == "ar":
            tx = {}
            for f in files:
                with open(f) as fh:
                    for line in fh:
                        parts = [p for p in line.strip().split(",") if p]
                        if len(parts) >= 2:
                            tx.setdefault(int(parts[0]), []).append(int(parts[1]))
            sets, rules = models.apriori(list(tx.values()), float(a[0]) if a else 0.1, float(a[1]) if len(a) > 1 else 0.6)[:2]
            return {"n_transactions": len(tx), "n_itemsets": len(sets), "n_rules": len(rules)}
        if algo in ("cholesky", "pivoted_qr", "sorting", "quantile", "zscore", "minmax", "outlier_uni", "outlier_multi",
                    "outlier_bacon", "kernel_lin", "kernel_rbf"):
            X = table(int(a[0]) if a else None)
            fn = {"cholesky": linalg.cholesky, "pivoted_qr": linalg.pivoted_qr, "sorting": linalg.sort_columns,
                  "quantile": linalg.quantiles, "zscore": linalg.zscore, "minmax": linalg.minmax,
                  "outlier_uni": linalg.outlier_univariate, "outlier_multi": linalg.outlier_multivariate,
                  "outlier_bacon": linalg.outlier_bacon, "kernel_lin": classify.kernel_linear,
                  "kernel_rbf": classify.kernel_rbf}[algo]
            return fn(X)
        if algo in ("mse", "sgd", "sgd_mini", "sgd_moment", "lbfgs", "adagrad"):
            d = int(a[0]); X, y = xy(d, long_labels=False)
            obj = models.MSEObjective(X, y)
            n_it = max(int(self.c["num_iterations"]), 1)
            if algo == "mse":
                theta = torch.zeros(d + 1, dtype=torch.float64, device=dev)
                val, grad = obj.value_and_grad(theta) if hasattr(obj, "value_and_grad") else (obj.value(theta), obj.gradient(theta))
                return {"value_at_zero": float(val), "gradient_norm": float(torch.as_tensor(grad).norm())}
            theta = {"sgd": lambda: models.solver_sgd(obj, n_it), "sgd_mini": lambda: models.solver_sgd(obj, n_it, batch_size=8),
                     "sgd_moment": lambda: models.solver_sgd(obj, n_it, momentum=0.9),
                     "lbfgs": lambda: models.solver_lbfgs(obj, n_it), "adagrad": lambda: models.solver_adagrad(obj, n_it)}[algo]()
            return {"theta": theta}
        if algo in ("quality_svm", "quality_linreg"):
            d = int(a[0])
            if algo == "quality_svm":
                X, y = xy(d); Xt, yt = xy(d, [a[1]])
                m = classify.svm_train_multiclass(X, y)
                return linalg.quality_metrics_classification(yt, m.predict(Xt), int(y.max().item()) + 1)
            T = table(None)
            k = T.shape[1] - d
            m = linalg.linear_regression_normeq(T[:, :d], T[:, d:])
            Tt = table(None, [a[1]]) if len(a) > 1 else T
            return linalg.quality_metrics_regression(Tt[:, d:d + k], m.predict(Tt[:, :d]))
        raise SystemExit(f"unknown batch algorithm {algo!r}")


BATCH = {"svm", "adaboost", "brownboost", "stump", "logitboost", "knn", "em", "als", "als_batch", "ar", "cholesky",
         "pivoted_qr", "sorting", "quantile", "zscore", "minmax", "outlier_uni", "outlier_multi", "outlier_bacon",
         "kernel_lin", "kernel_rbf", "mse", "sgd", "sgd_mini", "sgd_moment", "lbfgs", "adagrad", "quality_svm",
         "quality_linreg"} | {f"{k}_{t}{s}" for k in ("dtree", "dforest") for t in ("cls", "reg") for s in ("", "_traverse")}


def main(argv=None) -> int:
    setup_logging()
    pos, over = split_overrides(sys.argv[1:] if argv is None else argv)
    if len(pos) < 7:
        print(__doc__, file=sys.stderr)
        return -1
    algo = pos[0]
    conf = {}
    init = Initialize(conf, pos[1:])
    init.load_distributed_libs()
    sysargs = init.load_sys_args()
    conf.update({"algorithm": algo, "app_args": init.app_args()})
    conf.update(over)
    rank, world = check_world(sysargs.num_mappers)
    m = DaalMapper()
    run_mapper(m, sysargs.input_dir, conf)
    if rank == 0:
        os.makedirs(sysargs.work_dir, exist_ok=True)
        res = m.result
        with open(os.path.join(sysargs.work_dir, f"{algo}.out"), "w") as f:
            f.write(repr(res)[:100000])
        named = []
        if hasattr(res, "_fields"):
            named = [(n, getattr(res, n)) for n in res._fields]
        elif isinstance(res, dict):
            named = list(res.items())
        elif isinstance(res, torch.Tensor):
            named = [("result", res)]
        elif isinstance(res, (tuple, list)):
            named = [(f"result{i}", v) for i, v in enumerate(res)]
        for name, v in named:
            if isinstance(v, torch.Tensor):
                Service.print_numeric_table(f"{algo}.{name}:", v.detach().cpu())
            elif isinstance(v, (int, float, str)):
                print(f"{algo}.{name}: {v}")
        print(f"Harp-DAAL {algo} finished")
    return 0


if __name__ == "__main__":
    sys.exit(main())
