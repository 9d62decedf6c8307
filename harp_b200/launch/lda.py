"""``edu.iu.lda.LDALauncher`` (``ml/java/src/main/java/edu/iu/lda/LDALauncher.java:50-110``).

Usage: <doc dir> <num of topics> <alpha> <beta> <num of iterations> <min training percentage>
       <max training percentage> <num of mappers> <num of threads per worker> <schedule ratio>
       <memory (MB)> <work dir> <print model>
"""
from __future__ import annotations

import os
import sys

from ..apps.lda import LDACollectiveMapper
from .common import check_world, run_mapper, setup_logging, split_overrides


def main(argv=None) -> int:
    setup_logging()
    pos, over = split_overrides(sys.argv[1:] if argv is None else argv)
    if len(pos) < 13:
        print(__doc__, file=sys.stderr)
        return -1
    doc_dir, K, alpha, beta, n_iter, lo, hi, n_maps = pos[0], int(pos[1]), float(pos[2]), float(pos[3]), int(pos[4]), int(pos[5]), int(pos[6]), int(pos[7])
    work_dir, print_model = pos[11], pos[12].lower() == "true"
    rank, world = check_world(n_maps)
    conf = {"num_topics": K, "alpha": alpha, "beta": beta, "num_iterations": max(n_iter, 1), "min_bound": lo,
            "max_bound": hi, "model_dir": os.path.join(work_dir, "model"), "print_model": print_model}
    conf.update(over)
    m = LDACollectiveMapper()
    run_mapper(m, doc_dir, conf)
    if rank == 0:
        print(f"LDA-CGS finished: log-likelihood {m.log_likelihood:.6e}")
    return 0


if __name__ == "__main__":
    sys.exit(main())
