"""``edu.iu.kmeans.common.KmeansMapCollective`` (``contrib/.../kmeans/common/KmeansMapCollective.java:33-238``).

Usage: <numOfDataPoints> <num of Centroids> <size of vector> <number of map tasks>
       <number of iteration> <workDir> <localDir> <communication operation>
   communication operation = allreduce | regroup-allgather | broadcast-reduce | push-pull
"""
from __future__ import annotations

import os
import shutil
import sys

from ..apps.kmeans import ContribKMeansMapper
from ..io import datagen
from .common import check_world, master_first, run_mapper, setup_logging, split_overrides


def main(argv=None) -> int:
    setup_logging()
    pos, over = split_overrides(sys.argv[1:] if argv is None else argv)
    if len(pos) < 8:
        print(__doc__, file=sys.stderr)
        return -1
    n_points, n_cen, vec, n_maps, n_iter = (int(x) for x in pos[:5])
    work_dir, local_dir, op = pos[5], pos[6], pos[7]
    if op not in ContribKMeansMapper.VARIANTS:
        print(f"unknown communication operation {op!r}", file=sys.stderr)
        return -1
    rank, world = check_world(n_maps)
    data_dir = os.path.join(work_dir, "data")
    cen_dir = os.path.join(work_dir, "centroids")
    out_dir = os.path.join(work_dir, "out")

    def prepare():
        for d in (data_dir, cen_dir, out_dir):
            shutil.rmtree(d, ignore_errors=True)
        os.makedirs(cen_dir, exist_ok=True)
        seed = int(over.get("seed", 0))
        datagen.generate_points_text(n_points, vec, world, data_dir, seed=seed)
        # initial centroids = random rows in the data range (Utils.generateInitialCentroids)
        datagen.generate_centroids(n_cen, vec, os.path.join(cen_dir, "centroids_0"), seed=seed + 1, value_range=10.0)

    master_first(prepare)
    conf = {"num_centroids": n_cen, "vector_size": vec, "num_iterations": n_iter, "work_dir": work_dir,
            "centroid_file": os.path.join(cen_dir, "centroids_0"), "comm": op}
    conf.update(over)
    mapper = ContribKMeansMapper()
    run_mapper(mapper, data_dir, conf, out_dir, retries=3)
    if rank == 0:
        print(f"K-means ({op}) finished: MSE : {mapper.final_mse}")
        if getattr(mapper, "collective_launches", None) is not None:
            print(f"collective kernel launches: {mapper.collective_launches}")
    return 0


if __name__ == "__main__":
    sys.exit(main())
