"""``edu.iu.examples.ExamplesMain`` (``ml/java/.../examples/ExamplesMain.java:59-81``):

    torchrun --nproc-per-node 2 -m harp_b200.launch.examples -mappers 2 -elements 1000 -iterations 10 \\
        -operation allreduce -partitions 2 -data double -verify
"""
from __future__ import annotations

import argparse
import sys

from ..apps.examples import ExampleMapper
from .common import check_world, setup_logging


def main(argv=None) -> int:
    setup_logging()
    ap = argparse.ArgumentParser(prefix_chars="-")
    ap.add_argument("-mappers", type=int, default=2)
    ap.add_argument("-elements", type=int, default=1000)
    ap.add_argument("-iterations", type=int, default=100)
    ap.add_argument("-operation", default="allreduce")
    ap.add_argument("-partitions", type=int, default=1)
    ap.add_argument("-data", default="int")
    ap.add_argument("-verify", action="store_true")
    a = ap.parse_args(sys.argv[1:] if argv is None else argv)
    rank, world = check_world(a.mappers)
    m = ExampleMapper()
    ctx = m.run([], {"operation": a.operation, "elements": a.elements, "partitions": a.partitions,
                     "iterations": a.iterations, "data": a.data, "verify": a.verify})
    if rank == 0:
        print(f"Harp Example Completed! operation={a.operation} verified={m.verified} seconds={m.elapsed:.4f}")
    return 0 if m.verified else 1


if __name__ == "__main__":
    sys.exit(main())
