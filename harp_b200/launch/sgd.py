"""``edu.iu.sgd.SGDLauncher`` (``ml/java/src/main/java/edu/iu/sgd/SGDLauncher.java:48-100``).

Usage: <input dir> <r> <lambda> <epsilon> <num of iterations> <training ratio> <num of map tasks>
       <num of threads per worker> <schedule ratio> <memory (MB)> <work dir> <test dir>
"""
from __future__ import annotations

import os
import sys

from ..apps.mf import SGDCollectiveMapper
from ..io import multifile
from .common import check_world, run_mapper, setup_logging, split_overrides


def main(argv=None) -> int:
    setup_logging()
    pos, over = split_overrides(sys.argv[1:] if argv is None else argv)
    if len(pos) < 12:
        print(__doc__, file=sys.stderr)
        return -1
    input_dir, r, lam, eps, n_iter, train_ratio, n_maps = pos[0], int(pos[1]), float(pos[2]), float(pos[3]), int(pos[4]), int(pos[5]), int(pos[6])
    work_dir, test_dir = pos[10], pos[11]
    rank, world = check_world(n_maps)
    conf = {"r": r, "lambda": lam, "epsilon": eps, "num_iterations": max(n_iter, 1), "train_ratio": train_ratio,
            "model_dir": os.path.join(work_dir, "model"),
            "test_files": multifile.list_input_files(test_dir) if os.path.isdir(test_dir) else [test_dir]}
    conf.update(over)
    m = SGDCollectiveMapper()
    run_mapper(m, input_dir, conf)
    if rank == 0:
        print(f"MF-SGD finished: test RMSE {m.test_rmse:.6f}; model in {conf['model_dir']}")
    return 0


if __name__ == "__main__":
    sys.exit(main())
