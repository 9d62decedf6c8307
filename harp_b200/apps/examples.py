"""``edu.iu.examples`` (one mapper per collective with a self-checking ``-verify`` mode,
``ml/java/src/main/java/edu/iu/examples/{AllReduce,AllGather,Reduce,BCast,Rotate}.java``) and the collective
micro-benchmark ``edu.iu.benchmark`` (``BenchmarkMapper.java:96-136``)."""
from __future__ import annotations

import time


import torch

from ..mapper import CollectiveMapper, Context, KeyValReader
from ..partition import ArrayCombiner, DoubleArrPlus, Operation, Partition, Table

_DTYPES = {"int": torch.int32, "long": torch.int64, "float": torch.float32, "double": torch.float64,
           "short": torch.int16, "byte": torch.uint8}


class ExampleMapper(CollectiveMapper):
    """conf: ``operation`` (allreduce|allgather|reduce|bcast|rotate), ``elements``, ``partitions``, ``iterations``,
    ``data`` (int|long|float|double|short|byte), ``verify``."""

    def setup(self, context: Context) -> None:
        c = context.get_configuration()
        self.op = c.get("operation", "allreduce")
        self.elements = int(c.get("elements", 1000))
        self.partitions = int(c.get("partitions", 1))
        self.iterations = int(c.get("iterations", 100))
        self.dtype = _DTYPES[c.get("data", "int")]
        self.verify = bool(c.get("verify", False))

    def _table(self, value: float, ids=None) -> Table:
        t = Table(0, ArrayCombiner(Operation.SUM))
        for pid in (range(self.partitions) if ids is None else ids):
            t.add_partition(Partition(pid, torch.full((self.elements,), value, device=self.device).to(self.dtype)))
        return t

    def map_collective(self, reader: KeyValReader, context: Context) -> None:
        N, me = self.get_num_workers(), self.get_self_id()
        self.verified = True
        t0 = time.perf_counter()
        if self.op == "allreduce":
            iters = min(self.iterations, 8) if self.verify else self.iterations
            t = self._table(1)
            for it in range(iters):
                self.allreduce("examples", f"allreduce-{it}", t)
            if self.verify:                      # allreduce of ones must equal N^iter (AllReduce.java:43-71)
                want = float(N ** iters)
                for p in t.get_partitions():
                    self.verified &= bool((p.get().double() == want).all())
        elif self.op == "allgather":
            for it in range(self.iterations):
                t = self._table(me, ids=[me * self.partitions + j for j in range(self.partitions)])
                self.allgather("examples", f"allgather-{it}", t)
                if self.verify:
                    self.verified &= t.get_num_partitions() == N * self.partitions
                    for p in t.get_partitions():
                        self.verified &= bool((p.get().double() == float(p.id() // self.partitions)).all())
        elif self.op == "reduce":
            for it in range(self.iterations):
                t = self._table(1)
                self.reduce("examples", f"reduce-{it}", t, self.get_master_id())
                if self.verify:
                    if self.is_master():
                        for p in t.get_partitions():
                            self.verified &= bool((p.get().double() == float(N)).all())
                    else:
                        self.verified &= t.is_empty()
        elif self.op == "bcast":
            for it in range(self.iterations):
                t = self._table(it + 1) if self.is_master() else Table(0, ArrayCombiner(Operation.SUM))
                self.broadcast("examples", f"bcast-{it}", t, self.get_master_id(), it % 2 == 1)
                if self.verify:
                    self.verified &= t.get_num_partitions() == self.partitions
                    for p in t.get_partitions():
                        self.verified &= bool((p.get().double() == float(it + 1)).all())
        elif self.op == "rotate":
            t = self._table(me, ids=[me])
            for it in range(self.iterations):
                self.rotate("examples", f"rotate-{it}", t, None)
                if self.verify:                  # after it+1 ring steps we hold the block of worker me-(it+1)
                    want = (me - (it + 1)) % N
                    self.verified &= t.get_partition_ids() == [want]
                    self.verified &= bool((t.get_partition(want).get().double() == float(want)).all())
        else:
            raise ValueError(f"unknown operation {self.op!r}")
        self.elapsed = time.perf_counter() - t0
        ok = all(self.comm.all_gather_object(bool(self.verified)))
        self.verified = ok
        if self.is_master():
            context.write("operation", self.op)
            context.write("verified", ok)
            context.write("seconds", self.elapsed)


class BenchmarkMapper(CollectiveMapper):
    """conf: ``cmd`` (allreduce|allgather|bcast|rotate|regroup), ``bytes_per_partition``, ``num_partitions``,
    ``num_iterations``.  Payloads are float64 arrays like the reference's ``DoubleArray`` partitions.  On a GPU
    job the table is slab-backed inside the symmetric workspace when it fits, so ``allreduce`` exercises the
    in-kernel NVLink path; ``nccl=1`` forces the NCCL baseline for comparison."""

    def setup(self, context: Context) -> None:
        c = context.get_configuration()
        self.cmd = c.get("cmd", "allreduce")
        self.bytes_pp = int(c.get("bytes_per_partition", 1 << 20))
        self.nparts = int(c.get("num_partitions", 1))
        self.iters = int(c.get("num_iterations", 10))
        self.force_nccl = bool(int(c.get("nccl", 0)))

    def map_collective(self, reader: KeyValReader, context: Context) -> None:
        N, me, dev = self.get_num_workers(), self.get_self_id(), self.device
        n = max(1, self.bytes_pp // 8)
        symm = None if self.force_nccl else self.symm
        def make(ids):
            if symm is not None and len(ids) * n * 8 <= symm.data_bytes // 2:
                symm.reset()
                slab = symm.alloc(len(ids) * n, torch.float64).view(len(ids), n)
                slab.fill_(1.0)
            else:
                slab = torch.ones(len(ids), n, dtype=torch.float64, device=dev)
            return Table.dense(0, DoubleArrPlus(), slab, ids)
        if self.force_nccl:
            self.comm.symm = None
        times = []
        for it in range(self.iters + 2):
            if self.cmd in ("allreduce", "bcast", "regroup"):
                t = make(list(range(self.nparts)))
            else:
                t = make([me * self.nparts + j for j in range(self.nparts)])
            if dev.type == "cuda":
                torch.cuda.synchronize()
            self.barrier("bench", f"pre-{it}")
            t0 = time.perf_counter()
            if self.cmd == "allreduce":
                self.allreduce("bench", f"allreduce-{it}", t)
            elif self.cmd == "allgather":
                self.allgather("bench", f"allgather-{it}", t)
            elif self.cmd == "bcast":
                if not self.is_master():
                    t = Table(0, DoubleArrPlus())
                self.broadcast("bench", f"bcast-{it}", t, self.get_master_id(), False)
            elif self.cmd == "rotate":
                self.rotate("bench", f"rotate-{it}", t, None)
            elif self.cmd == "regroup":
                self.regroup("bench", f"regroup-{it}", t, None)
            else:
                raise ValueError(self.cmd)
            if dev.type == "cuda":
                torch.cuda.synchronize()
            dt = time.perf_counter() - t0
            if it >= 2:
                times.append(dt)
        worst = max(self.comm.all_gather_object(sum(times) / max(len(times), 1)))
        self.seconds_per_op = worst
        self.bytes_moved = self.nparts * n * 8
        if self.is_master():
            context.write("cmd", self.cmd)
            context.write("ms_per_op", worst * 1e3)
            context.write("MB_per_table", self.bytes_moved / 1e6)
