"""``edu.iu.benchmark.JobLauncher`` (``ml/java/.../benchmark/JobLauncher.java:40-80``).

Usage: <command> <number of bytes per partition> <number of partitions> <number of mappers> <number of iterations>
   command = allreduce | allgather | bcast | rotate | regroup      (append --nccl=1 for the NCCL baseline)
"""
from __future__ import annotations

import sys

from ..apps.examples import BenchmarkMapper
from .common import check_world, setup_logging, split_overrides


def main(argv=None) -> int:
    setup_logging()
    pos, over = split_overrides(sys.argv[1:] if argv is None else argv)
    if len(pos) < 5:
        print(__doc__, file=sys.stderr)
        return -1
    rank, world = check_world(int(pos[3]))
    conf = {"cmd": pos[0], "bytes_per_partition": int(pos[1]), "num_partitions": int(pos[2]), "num_iterations": int(pos[4])}
    conf.update(over)
    m = BenchmarkMapper()
    m.run([], conf)
    if rank == 0:
        gbs = m.bytes_moved / m.seconds_per_op / 1e9
        print(f"{pos[0]}: {m.seconds_per_op * 1e3:.3f} ms per op over {world} workers, table {m.bytes_moved / 1e6:.2f} MB "
              f"({gbs:.2f} GB/s algorithmic)")
    return 0


if __name__ == "__main__":
    sys.exit(main())
