"""Harp-DAAL launchers: ``python -m harp_b200.launch.daal <algorithm> <num_mappers> <threads> <mem_MB>
<iterations> <inputDir> <workDir> [app args]`` — the six standard arguments of ``Initialize.loadSysArgs``
followed by the per-algorithm ones of the reference launchers:

  cov | pca | pca_svd | mom        <fileDim>                    (daal_cov, daal_pca, daal_mom)
  linreg | linreg_qr | ridge       <featureDim> <nDependent>    (daal_linreg, daal_ridgereg)
  naive                            <featureDim> <nClasses>      (daal_naive; last column = label)
  kmeans                           <featureDim> <nCentroids> [comm]   (daal_kmeans)
  qr | svd                         <fileDim>                    (daal_qr, daal_svd)
  nn                               <featureDim> <nClasses> <batch> <lr>   (daal_nn)

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
        else:
            raise SystemExit(f"unknown algorithm {algo!r}")


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
        if hasattr(res, "_fields"):
            for name in res._fields:
                v = getattr(res, name)
                if isinstance(v, torch.Tensor):
                    Service.print_numeric_table(f"{algo}.{name}:", v.detach().cpu())
        print(f"Harp-DAAL {algo} finished")
    return 0


if __name__ == "__main__":
    sys.exit(main())
