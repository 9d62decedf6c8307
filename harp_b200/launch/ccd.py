"""``edu.iu.ccd.CCDLauncher`` (``ml/java/src/main/java/edu/iu/ccd/CCDLauncher.java:52-63``).

Usage: <input dir> <r> <lambda> <num of iterations> <num of map tasks> <num of threads per worker>
       <memory (MB)> <work dir> <test dir>
"""
from __future__ import annotations

import os
import sys

from ..apps.mf import CCDCollectiveMapper
from ..io import multifile
from .common import check_world, run_mapper, setup_logging, split_overrides


def main(argv=None) -> int:
    setup_logging()
    pos, over = split_overrides(sys.argv[1:] if argv is None else argv)
    if len(pos) < 9:
        print(__doc__, file=sys.stderr)
        return -1
    input_dir, r, lam, n_iter, n_maps = pos[0], int(pos[1]), float(pos[2]), int(pos[3]), int(pos[4])
    work_dir, test_dir = pos[7], pos[8]
    rank, world = check_world(n_maps)
    conf = {"r": r, "lambda": lam, "num_iterations": max(n_iter, 1), "model_dir": os.path.join(work_dir, "model"),
            "test_files": multifile.list_input_files(test_dir) if os.path.isdir(test_dir) else [test_dir]}
    conf.update(over)
    m = CCDCollectiveMapper()
    run_mapper(m, input_dir, conf)
    if rank == 0:
        print(f"MF-CCD finished: test RMSE {m.test_rmse:.6f}")
    return 0


if __name__ == "__main__":
    sys.exit(main())
