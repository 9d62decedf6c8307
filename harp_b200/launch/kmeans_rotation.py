"""``edu.iu.kmeans.rotation.KMeansLauncher`` — same positional arguments as
``harp_b200.launch.kmeans`` (``ml/java/.../kmeans/rotation/KMeansLauncher.java``)."""
from __future__ import annotations

import os
import shutil
import sys

from ..apps.kmeans import KMeansRotationMapper
from ..io import datagen
from .common import check_world, master_first, run_mapper, setup_logging, split_overrides


def main(argv=None) -> int:
    setup_logging()
    pos, over = split_overrides(sys.argv[1:] if argv is None else argv)
    if len(pos) < 9:
        print("Usage: <points> <centroids> <vector size> <files per worker> <map tasks> <threads> "
              "<iterations> <work dir> <local dir>", file=sys.stderr)
        return -1
    n_points, n_cen, vec, fpw, n_maps, _t, n_iter = (int(x) for x in pos[:7])
    work_dir = pos[7]
    rank, world = check_world(n_maps)
    data_dir = os.path.join(work_dir, "data")
    cen_dir = os.path.join(work_dir, "centroids")

    def prepare():
        shutil.rmtree(data_dir, ignore_errors=True)
        shutil.rmtree(cen_dir, ignore_errors=True)
        os.makedirs(cen_dir, exist_ok=True)
        datagen.generate_kmeans_data(n_points, vec, world * fpw, data_dir, seed=int(over.get("seed", 0)))
        datagen.generate_centroids(n_cen, vec, os.path.join(cen_dir, "init_centroids"), seed=int(over.get("seed", 0)) + 1)

    master_first(prepare)
    conf = {"num_centroids": n_cen, "vector_size": vec, "num_iterations": n_iter or 1,
            "centroid_file": os.path.join(cen_dir, "init_centroids"),
            "out_file": os.path.join(cen_dir, "out", "output")}
    mapper = KMeansRotationMapper()
    run_mapper(mapper, data_dir, conf)
    if rank == 0:
        print(f"K-means (rotation) finished; objective {mapper.objective_history[-1]:.6g}")
    return 0


if __name__ == "__main__":
    sys.exit(main())
