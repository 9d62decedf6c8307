"""Launchers of the remaining ``contrib/`` tutorials and graph / embedding apps:

  pagerank  <input dir> <num urls> <num iterations> <num map tasks> <work dir>      (edu.iu.simplepagerank.HarpPageRank)
  svm       <input dir> <num iterations> <num map tasks> <work dir>                 (edu.iu.svm.IterativeSVM)
  mlr       <input dir> <num topics> <num terms> <num iterations> <alpha> <num map tasks> <work dir>   (edu.iu.mlr.MLRMapCollective)
  ldacvb    <input dir> <num topics> <num terms> <num iterations> <num map tasks> <work dir> [dynamic]  (edu.iu.lda.LdaMapCollective)
  nn        <input dir> <num features> <num classes> <epochs> <sync iters> <num map tasks> <work dir>   (edu.iu.NN.NNMapCollective)
  rf        <input dir> <test file> <num classes> <trees per worker> <num map tasks> <work dir>         (edu.iu.rf / randomforest / com.rf.fast)
  subgraph  <graph dir> <template file> <num colourings> <num map tasks> <work dir> [allgather|rotation] (edu.iu.subgraph.SCLauncher, edu.iu.sahad.*)
  mds       <distance dir> <n> <target dim> <max iter> <num map tasks> <work dir>    (edu.iu.wdamds.MDSLauncher)
  sgxkmeans <points dir> <centroid file> <num centroids> <vector size> <iterations> <num map tasks> <work dir>  (edu.iu.kmeans.sgxsimu)
"""
from __future__ import annotations

import os
import sys

from ..apps import contrib as AC, graph as AG
from .common import check_world, run_mapper, setup_logging, split_overrides


def main(argv=None) -> int:
    setup_logging()
    pos, over = split_overrides(sys.argv[1:] if argv is None else argv)
    if len(pos) < 4:
        print(__doc__, file=sys.stderr)
        return -1
    app, a = pos[0], pos[1:]
    try:
        if app == "pagerank":
            m, conf, maps, work = AC.PageRankMapper(), {"num_urls": int(a[1]), "num_iterations": int(a[2])}, int(a[3]), a[4]
        elif app == "svm":
            m, conf, maps, work = AC.IterativeSVMMapper(), {"num_iterations": int(a[1])}, int(a[2]), a[3]
        elif app == "mlr":
            m, conf, maps, work = AC.MLRMapper(), {"num_topics": int(a[1]), "num_terms": int(a[2]), "num_iterations": int(a[3]), "alpha": float(a[4])}, int(a[5]), a[6]
        elif app == "ldacvb":
            m, conf, maps, work = AC.LDACVBMapper(), {"num_topics": int(a[1]), "num_terms": int(a[2]), "num_iterations": int(a[3]),
                                                     "dynamic": len(a) > 6 and a[6] == "dynamic"}, int(a[4]), a[5]
        elif app == "nn":
            m, conf, maps, work = AC.NNMapper(), {"num_features": int(a[1]), "num_classes": int(a[2]), "epochs": int(a[3]), "sync_iter": int(a[4])}, int(a[5]), a[6]
        elif app == "rf":
            m, conf, maps, work = AC.RandomForestMapper(), {"test_file": a[1], "num_classes": int(a[2]), "num_trees": int(a[3])}, int(a[4]), a[5]
        elif app == "subgraph":
            m, conf, maps, work = AG.SubgraphMapper(), {"template_file": a[1], "num_iterations": int(a[2]), "comm": a[5] if len(a) > 5 else "allgather"}, int(a[3]), a[4]
        elif app == "mds":
            m, conf, maps, work = AG.WDAMDSMapper(), {"n": int(a[1]), "target_dim": int(a[2]), "max_iter": int(a[3])}, int(a[4]), a[5]
        elif app == "sgxkmeans":
            m, conf, maps, work = AG.SGXSimuKMeansMapper(), {"centroid_file": a[1], "num_centroids": int(a[2]), "vector_size": int(a[3]),
                                                              "num_iterations": int(a[4])}, int(a[5]), a[6]
        else:
            print(__doc__, file=sys.stderr)
            return -1
    except (IndexError, ValueError):
        print(__doc__, file=sys.stderr)
        return -1
    conf.update(over)
    rank, world = check_world(maps)
    run_mapper(m, a[0], conf, os.path.join(work, "out"))
    if rank == 0:
        summary = {k: getattr(m, k) for k in ("train_accuracy", "test_accuracy", "macro_f1", "count", "stress", "sgx_overhead_ms") if hasattr(m, k)}
        print(f"{app} finished: {summary}")
    return 0


if __name__ == "__main__":
    sys.exit(main())
