"""``HarpSession`` — the NumPy-facing session API of the reference's ``python/`` directory, working.

The reference ships a py4j stub (``python/harp_session.py``, ``python/collectives/collectives.py``,
``python/harp_constants.py``): ``HarpSession(name).com.barrier(ctx, op)`` is ``pass`` and ``broadcast`` ends in a Java
method that prints and returns false.  Here the same surface runs on this package's collectives: arrays are
``numpy.ndarray`` (or torch tensors), updated in place, on whatever process group ``torchrun`` started.

    s = HarpSession("demo")
    s.com.broadcast("demo", "bcast-1", arr, Type.DOUBLE, PartitioningMode.HOMOGENEOUS, 0, False)
    s.com.allreduce("demo", "sum-1", arr, Type.DOUBLE)
"""
from __future__ import annotations

from enum import Enum
from typing import Optional, Union

import numpy as np
import torch

from . import collective as C
from .mapper import init_distributed
from .partition import ArrayCombiner, Operation, Partition, Table

ArrayLike = Union[np.ndarray, torch.Tensor]


class Type(Enum):                      # python/harp_constants.py (= io/DataType.java tags 1..6)
    BYTE = 1
    SHORT = 2
    INT = 3
    FLOAT = 4
    LONG = 5
    DOUBLE = 6


class PartitioningMode(Enum):          # python/harp_constants.py
    HETEROGENEOUS = 1
    HOMOGENEOUS = 2


_TORCH_DTYPE = {Type.BYTE: torch.uint8, Type.SHORT: torch.int16, Type.INT: torch.int32, Type.FLOAT: torch.float32,
                Type.LONG: torch.int64, Type.DOUBLE: torch.float64}


class Collectives:
    """``python/collectives/collectives.py`` plus the ops the stub never got to."""

    def __init__(self, ctx: "HarpContext"):
        self._ctx = ctx

    def _as_tensor(self, data: ArrayLike, data_type: Optional[Type]) -> torch.Tensor:
        t = torch.from_numpy(data) if isinstance(data, np.ndarray) else data
        if data_type is not None and t.dtype != _TORCH_DTYPE[data_type]:
            raise TypeError(f"array is {t.dtype}, declared {data_type.name}")
        return t

    def _table(self, t: torch.Tensor, op: Operation) -> Table:
        table = Table(0, ArrayCombiner(op))
        table.add_partition(Partition(0, t.view(-1)))
        return table

    def barrier(self, ctx_name: str, op_name: str) -> bool:
        return C.barrier(self._ctx.comm)

    def broadcast(self, ctx_name: str, op_name: str, data: ArrayLike, data_type: Optional[Type] = None,
                  partition_mode: PartitioningMode = PartitioningMode.HOMOGENEOUS, bcast_worker_id: int = 0,
                  use_mst_bcast: bool = False) -> bool:
        """Every worker's ``data`` becomes the root's (in place)."""
        t = self._as_tensor(data, data_type)
        comm = self._ctx.comm
        table = Table(0, ArrayCombiner(Operation.SUM))
        if comm.rank == bcast_worker_id:
            table.add_partition(Partition(0, t.view(-1)))
        ok = C.broadcast(comm, table, bcast_worker_id, use_mst_bcast)
        if comm.rank != bcast_worker_id:
            got = table.get_partition(0).get()
            got = got.get() if hasattr(got, "get") and not isinstance(got, torch.Tensor) else got
            t.view(-1).copy_(got.to(t.device))
        return ok

    def allreduce(self, ctx_name: str, op_name: str, data: ArrayLike, data_type: Optional[Type] = None,
                  op: Operation = Operation.SUM) -> bool:
        t = self._as_tensor(data, data_type)
        return C.allreduce(self._ctx.comm, self._table(t, op))

    def reduce(self, ctx_name: str, op_name: str, data: ArrayLike, data_type: Optional[Type] = None,
               reduce_worker_id: int = 0, op: Operation = Operation.SUM) -> bool:
        """The root's ``data`` becomes the combination of everyone's; other workers keep theirs."""
        t = self._as_tensor(data, data_type)
        table = self._table(t.clone() if self._ctx.comm.rank != reduce_worker_id else t, op)
        return C.reduce(self._ctx.comm, table, reduce_worker_id)

    def allgather(self, ctx_name: str, op_name: str, data: ArrayLike, data_type: Optional[Type] = None):
        """Returns the workers' arrays stacked in worker order (``[num_workers, ...]``)."""
        t = self._as_tensor(data, data_type)
        comm = self._ctx.comm
        table = Table(0, ArrayCombiner(Operation.SUM))
        table.add_partition(Partition(comm.rank, t.reshape(-1).clone()))
        C.allgather(comm, table)
        rows = []
        for r in range(comm.world):
            p = table.get_partition(r).get()
            rows.append((p.get() if not isinstance(p, torch.Tensor) else p).view(t.shape))
        out = torch.stack(rows)
        return out.numpy() if isinstance(data, np.ndarray) else out


class HarpContext:
    """``python/context/harp_context.py``: name, this worker's id, the collectives."""

    def __init__(self, name: str):
        self._name = name
        rank, world, device = init_distributed()
        self.comm = C.Comm(device=torch.device("cpu") if device.type != "cuda" else device)
        self._com = Collectives(self)

    @property
    def name(self) -> str:
        return self._name

    @property
    def self_id(self) -> int:
        return self.comm.rank

    @property
    def num_workers(self) -> int:
        return self.comm.world

    @property
    def com(self) -> Collectives:
        return self._com


class HarpSession:
    """``python/harp_session.py``."""

    def __init__(self, name: str):
        self.ctx = HarpContext(name)

    @property
    def name(self) -> str:
        return self.ctx.name

    @property
    def com(self) -> Collectives:
        return self.ctx.com
