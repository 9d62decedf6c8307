"""Shared launcher plumbing: one process per GPU under ``torchrun`` replaces ``hadoop jar … Launcher``.

Launchers keep the reference's positional argument order; ``<number of map tasks>`` must equal the
torchrun world size (or 1 for a single process), ``<num threads>`` / ``<mem>`` are accepted and
ignored (CTAs per SM and HBM replace them).  ``--key=value`` pairs after the positionals override
configuration entries (the ``-D k=v`` of Hadoop's ToolRunner).
"""
from __future__ import annotations

import logging
import os
import sys
import time
from typing import Any, Dict, List, Sequence, Tuple

import torch.distributed as dist

from ..io import multifile
from ..mapper import CollectiveMapper, Context, init_distributed


def split_overrides(argv: Sequence[str]) -> Tuple[List[str], Dict[str, str]]:
    pos, over = [], {}
    for a in argv:
        if a.startswith("--") and "=" in a:
            k, v = a[2:].split("=", 1)
            over[k.replace("-", "_")] = v
        elif a.startswith("-D") and "=" in a:
            k, v = a[2:].split("=", 1)
            over[k] = v
        else:
            pos.append(a)
    return pos, over


def check_world(num_map_tasks: int) -> Tuple[int, int]:
    rank, world, _ = init_distributed()
    if num_map_tasks != world:
        if rank == 0:
            print(f"[harp_b200] <number of map tasks>={num_map_tasks} but the job runs {world} process(es); "
                  f"using {world} workers", file=sys.stderr)
    return rank, world


def master_first(fn, *args, **kwargs):
    """Runs ``fn`` on rank 0 only, then a barrier (data generation before the job)."""
    rank = dist.get_rank() if dist.is_initialized() else 0
    out = None
    if rank == 0:
        out = fn(*args, **kwargs)
    if dist.is_initialized():
        dist.barrier()
    return out


def run_mapper(mapper: CollectiveMapper, input_dir: str, conf: Dict[str, Any], output_dir=None,
               retries: int = 1) -> Context:
    """Deals the input files like MultiFileInputFormat and runs the mapper; whole-job retry like
    the reference launchers (``contrib/.../kmeans/common/KmeansMapCollective.java:129-177``)."""
    rank = dist.get_rank() if dist.is_initialized() else 0
    world = dist.get_world_size() if dist.is_initialized() else 1
    files = multifile.list_input_files(input_dir) if input_dir else []
    mine = multifile.shard_files(files, rank, world)
    last = None
    for attempt in range(max(1, retries)):
        try:
            t0 = time.time()
            ctx = mapper.run(multifile.records_for(mine), conf, output_dir)
            if rank == 0:
                logging.getLogger("harp_b200").info("job finished in %.3f s", time.time() - t0)
            return ctx
        except IOError as e:  # noqa: PERF203
            last = e
    raise last


def setup_logging() -> None:
    logging.basicConfig(level=os.environ.get("HARP_LOG", "WARNING"),
                        format="%(asctime)s %(name)s %(levelname)s %(message)s")
