"""``edu.iu.kmeans.regroupallgather.KMeansLauncher`` (``…/KMeansLauncher.java:40-226``).

    torchrun --nproc-per-node 2 --master-addr 127.0.0.1 -m harp_b200.launch.kmeans \\
        1000 10 100 5 2 2 10 /tmp/kmeans_work /tmp/kmeans_local

Usage: <num of data points> <num of centroids> <vector size> <num of point files per worker>
       <number of map tasks> <num threads> <number of iterations> <work dir> <local points dir>
       [regenerate data: true|false]
Options: --seed=<n>; --engine=device runs the iterations in ``models.kmeans.KMeansEngine`` (one fused launch per
iteration on GPUs) instead of the reference's table regroup/allgather structure (default ``--engine=table``); --resume=<centroid file> continues from a centroid dump of an earlier run (the reference
only writes such dumps, KMUtil.storeCentroids; nothing in it reads them back — SURVEY §5.4) and keeps the data.
"""
from __future__ import annotations

import os
import shutil
import sys

from ..apps.kmeans import KMeansRegroupAllgatherMapper
from ..io import datagen
from .common import check_world, master_first, run_mapper, setup_logging, split_overrides

USAGE = __doc__.split("Usage:")[1]


def main(argv=None) -> int:
    setup_logging()
    pos, over = split_overrides(sys.argv[1:] if argv is None else argv)
    if len(pos) < 9:
        print("Usage:" + USAGE, file=sys.stderr)
        return -1
    n_points, n_cen, vec, files_per_worker, n_maps, _threads, n_iter = (int(x) for x in pos[:7])
    work_dir, local_dir = pos[7], pos[8]
    regenerate = (pos[9].lower() == "true") if len(pos) > 9 else True
    resume = over.pop("resume", None)
    if resume is not None:
        regenerate = False
    rank, world = check_world(n_maps)
    n_files = world * files_per_worker
    if n_points // n_files == 0 or n_cen // world == 0:
        return -1
    n_iter = n_iter or 1
    data_dir = os.path.join(work_dir, "data")
    cen_dir = os.path.join(work_dir, "centroids")
    out_dir = os.path.join(work_dir, "out")

    resume_copy = os.path.join(work_dir, "resume_centroids")

    def prepare():
        if resume is not None:           # the dump usually lives under cen_dir, which is wiped next
            shutil.copyfile(resume, resume_copy)
        shutil.rmtree(cen_dir, ignore_errors=True)
        shutil.rmtree(out_dir, ignore_errors=True)
        os.makedirs(cen_dir, exist_ok=True)
        if regenerate or not os.path.isdir(data_dir):
            shutil.rmtree(data_dir, ignore_errors=True)
            os.makedirs(local_dir, exist_ok=True)
            datagen.generate_kmeans_data(n_points, vec, n_files, data_dir, seed=int(over.get("seed", 0)))
        if resume is not None:
            shutil.copyfile(resume_copy, os.path.join(cen_dir, "init_centroids"))
        else:
            datagen.generate_centroids(n_cen, vec, os.path.join(cen_dir, "init_centroids"),
                                       seed=int(over.get("seed", 0)) + 1)

    master_first(prepare)
    conf = {"num_centroids": n_cen, "vector_size": vec, "num_iterations": n_iter,
            "centroid_file": os.path.join(cen_dir, "init_centroids"),
            "out_file": os.path.join(cen_dir, "out", "output")}
    conf.update(over)
    mapper = KMeansRegroupAllgatherMapper()
    run_mapper(mapper, data_dir, conf, out_dir)
    if rank == 0:
        print(f"K-means (regroup-allgather) finished: {n_iter} iterations, objective "
              f"{mapper.objective_history[-1]:.6g}; centroids in {conf['out_file']}")
    return 0


if __name__ == "__main__":
    sys.exit(main())
