"""Table / Partition / Combiner / Partitioner — Harp's data model on torch tensors.

Reference semantics (citations relative to
``core/harp-collective/src/main/java/edu/iu/harp``):

* ``Table.addPartition`` inserts when the id is absent and otherwise *combines in
  place* with the table's combiner — the reduction rule of every collective
  (``partition/Table.java:116-128``).
* ``Partitioner.getWorkerID(pid) = |pid % numWorkers|`` with Java remainder
  semantics for negative ids (``partition/Partitioner.java:36-43``).
* element-wise combiners SUM/MINUS/MULTIPLY/MAX/MIN on equal-length arrays, a
  length mismatch yields ``COMBINE_FAILED`` (``combiner/DoubleArrCombiner.java:23-68``).

B200-first differences: payloads are torch tensors (CPU or CUDA); a *dense*
table keeps all partitions as views into one contiguous slab so that the
NVLink collectives can move / reduce it with a single kernel.
"""
from __future__ import annotations

import enum
from typing import Any, Callable, Dict, Generic, Iterable, Iterator, List, Optional, Sequence, TypeVar

import torch

P = TypeVar("P")


class PartitionStatus(enum.Enum):
    ADDED = 0
    ADD_FAILED = 1
    COMBINED = 2
    COMBINE_FAILED = 3


class Operation(enum.Enum):
    """``combiner/Operation.java:3-9``."""
    SUM = 0
    MINUS = 1
    MULTIPLY = 2
    MAX = 3
    MIN = 4


class Partition(Generic[P]):
    """``(id, payload)`` — ``partition/Partition.java:32-58``."""

    __slots__ = ("_id", "_data", "_owner")

    def __init__(self, partition_id: int, data: P):
        self._id = int(partition_id)
        self._data = data
        self._owner = None          # the slab-backed Table this partition is a row view of (Table.dense)

    def id(self) -> int:
        return self._id

    def get(self) -> P:
        return self._data

    def set(self, data: P) -> None:
        self._data = data
        if self._owner is not None:
            # the payload no longer is the slab row: the table must not take the flat-buffer collective paths
            self._owner.slab = None
            self._owner = None

    def num_bytes(self) -> int:
        d = self._data
        if isinstance(d, torch.Tensor):
            return 4 + d.numel() * d.element_size()
        if hasattr(d, "get_num_write_bytes"):
            return 4 + int(d.get_num_write_bytes())
        return 4

    def release(self) -> None:
        d = self._data
        if hasattr(d, "release"):
            d.release()
        self._data = None

    free = release

    def __repr__(self) -> str:
        return f"Partition(id={self._id}, data={type(self._data).__name__})"


class PartitionCombiner(Generic[P]):
    """``combine(cur, new) -> PartitionStatus`` — ``partition/PartitionCombiner.java:25-37``."""

    #: set by array combiners so the collective layer can pick a fused device reduction
    op: Optional[Operation] = None

    def combine(self, cur: P, new: P) -> PartitionStatus:  # pragma: no cover - interface
        raise NotImplementedError


class ArrayCombiner(PartitionCombiner[torch.Tensor]):
    """Element-wise tensor combiner (``combiner/DoubleArrCombiner.java:23-68`` and its
    Byte/Int/Long/Short twins — the dtype is carried by the tensor)."""

    def __init__(self, op: Operation = Operation.SUM):
        self.op = op

    def combine(self, cur: torch.Tensor, new: torch.Tensor) -> PartitionStatus:
        if cur.numel() != new.numel():
            return PartitionStatus.COMBINE_FAILED
        new = new.to(device=cur.device, dtype=cur.dtype).view_as(cur)
        if self.op is Operation.SUM:
            cur.add_(new)
        elif self.op is Operation.MINUS:
            cur.sub_(new)
        elif self.op is Operation.MULTIPLY:
            cur.mul_(new)
        elif self.op is Operation.MAX:
            torch.maximum(cur, new, out=cur)
        elif self.op is Operation.MIN:
            torch.minimum(cur, new, out=cur)
        else:  # pragma: no cover
            return PartitionStatus.COMBINE_FAILED
        return PartitionStatus.COMBINED


# Reference names kept as aliases (``combiner/*ArrCombiner.java``, ``example/*ArrPlus.java``).
ByteArrCombiner = ShortArrCombiner = IntArrCombiner = LongArrCombiner = FloatArrCombiner = DoubleArrCombiner = ArrayCombiner


def DoubleArrPlus() -> ArrayCombiner:
    """``example/DoubleArrPlus.java:26-49``."""
    return ArrayCombiner(Operation.SUM)


IntArrPlus = LongArrPlus = FloatArrPlus = DoubleArrPlus


def DoubleArrMax() -> ArrayCombiner:
    return ArrayCombiner(Operation.MAX)


LongArrMax = IntArrMax = DoubleArrMax


class FnCombiner(PartitionCombiner[Any]):
    """Wraps ``fn(cur, new) -> merged | None`` (None = merged in place) for custom payloads
    (string concatenation, log-sum-exp, set union … SURVEY §9.1 item 8)."""

    def __init__(self, fn: Callable[[Any, Any], Any]):
        self.fn = fn
        self._slot = None  # set by Table to write back replaced payloads

    def combine(self, cur, new) -> PartitionStatus:
        out = self.fn(cur, new)
        self._slot = out
        return PartitionStatus.COMBINED


class NeverCombine(PartitionCombiner[Any]):
    """ids must never collide (``ml/java/.../lda/TopicCountListCombiner.java:25-30``)."""

    def combine(self, cur, new) -> PartitionStatus:
        return PartitionStatus.COMBINE_FAILED


class LogSumExpCombiner(PartitionCombiner[torch.Tensor]):
    """log-add of two log-domain arrays (``contrib/.../lda/DoubleArrLogPlus.java``)."""

    def combine(self, cur, new) -> PartitionStatus:
        if cur.numel() != new.numel():
            return PartitionStatus.COMBINE_FAILED
        torch.logaddexp(cur, new.to(cur.device, cur.dtype).view_as(cur), out=cur)
        return PartitionStatus.COMBINED


class Partitioner:
    """``partition/Partitioner.java:28-43``."""

    def __init__(self, num_workers: int):
        self.num_workers = int(num_workers)

    def get_num_workers(self) -> int:
        return self.num_workers

    def get_worker_id(self, partition_id: int) -> int:
        pid = int(partition_id)
        # Java's % truncates toward zero; abs() of that remainder
        r = abs(pid) % self.num_workers
        return r

    getWorkerID = get_worker_id

    def worker_ids(self, partition_ids: torch.Tensor) -> torch.Tensor:
        """Vectorised owner computation."""
        return partition_ids.abs().remainder(self.num_workers)


class MappedPartitioner(Partitioner):
    """Explicit ``pid -> worker`` map with the modulo rule as fallback."""

    def __init__(self, num_workers: int, mapping: Dict[int, int]):
        super().__init__(num_workers)
        self.mapping = dict(mapping)

    def get_worker_id(self, partition_id: int) -> int:
        w = self.mapping.get(int(partition_id))
        return super().get_worker_id(partition_id) if w is None else w

    getWorkerID = get_worker_id


class RandomPartitioner(Partitioner):
    """Seeded random owner per id (``ml/java/.../sgd/RandomPartitioner.java:29-37``): every
    worker derives the same owner for the same id from the shared seed."""

    def __init__(self, num_workers: int, seed: int):
        super().__init__(num_workers)
        self.seed = int(seed)

    def get_worker_id(self, partition_id: int) -> int:
        # splitmix64 of (seed, id): stateless, identical on all workers
        x = (self.seed * 0x9E3779B97F4A7C15 + (int(partition_id) & 0xFFFFFFFFFFFFFFFF)) & 0xFFFFFFFFFFFFFFFF
        x ^= x >> 30
        x = (x * 0xBF58476D1CE4E5B9) & 0xFFFFFFFFFFFFFFFF
        x ^= x >> 27
        x = (x * 0x94D049BB133111EB) & 0xFFFFFFFFFFFFFFFF
        x ^= x >> 31
        return int(x % self.num_workers)

    getWorkerID = get_worker_id

    def worker_ids(self, partition_ids: torch.Tensor) -> torch.Tensor:
        """Vectorised splitmix64 (two's-complement int64 arithmetic wraps like the scalar version)."""
        def lsr(v, n):      # logical shift right on int64
            return (v >> n) & ((1 << (64 - n)) - 1)
        def mul(v, c):      # wrapping multiply by an unsigned 64-bit constant
            c = c - (1 << 64) if c >= (1 << 63) else c
            return v * c
        seed_mix = (self.seed * 0x9E3779B97F4A7C15) & 0xFFFFFFFFFFFFFFFF
        seed_mix = seed_mix - (1 << 64) if seed_mix >= (1 << 63) else seed_mix
        x = partition_ids.to(torch.int64) + seed_mix
        x = x ^ lsr(x, 30)
        x = mul(x, 0xBF58476D1CE4E5B9)
        x = x ^ lsr(x, 27)
        x = mul(x, 0x94D049BB133111EB)
        x = x ^ lsr(x, 31)
        # unsigned modulo of a value stored as signed int64
        n = self.num_workers
        lo = (x & 0x7FFFFFFFFFFFFFFF) % n
        hi = torch.where(x < 0, torch.full_like(x, (1 << 63) % n), torch.zeros_like(x))
        return (lo + hi) % n


class PartitionFunction:
    """Post-regroup per-partition callback (``partition/PartitionFunction.java:25-27``)."""

    def apply(self, partition: Partition) -> None:  # pragma: no cover - interface
        raise NotImplementedError


class Table(Generic[P]):
    """Per-worker ``partitionID -> Partition`` map with a table-level combiner
    (``partition/Table.java:45-193``)."""

    def __init__(self, table_id: int, combiner: PartitionCombiner, size_hint: int = 0):
        self._table_id = int(table_id)
        self._combiner = combiner
        self._partitions: Dict[int, Partition[P]] = {}
        #: contiguous backing slab when built by :meth:`dense` (None otherwise)
        self.slab: Optional[torch.Tensor] = None

    # -- construction helpers -------------------------------------------------
    @classmethod
    def dense(cls, table_id: int, combiner: PartitionCombiner, slab: torch.Tensor,
              partition_ids: Optional[Sequence[int]] = None) -> "Table[torch.Tensor]":
        """Table whose partitions are row-views of ``slab[num_partitions, ...]``."""
        t = cls(table_id, combiner)
        ids = list(range(slab.shape[0])) if partition_ids is None else list(partition_ids)
        assert len(ids) == slab.shape[0]
        for i, pid in enumerate(ids):
            part = Partition(pid, slab[i])
            part._owner = t
            t._partitions[int(pid)] = part
        t.slab = slab
        return t

    def is_dense(self) -> bool:
        return self.slab is not None and len(self._partitions) == self.slab.shape[0]

    # -- reference API ---------------------------------------------------------
    def get_table_id(self) -> int:
        return self._table_id

    def get_combiner(self) -> PartitionCombiner:
        return self._combiner

    def get_num_partitions(self) -> int:
        return len(self._partitions)

    def get_partition_ids(self) -> List[int]:
        return list(self._partitions.keys())

    def get_partitions(self) -> List[Partition[P]]:
        return list(self._partitions.values())

    def add_partition(self, partition: Optional[Partition[P]]) -> PartitionStatus:
        if partition is None:
            return PartitionStatus.ADD_FAILED
        cur = self._partitions.get(partition.id())
        if cur is None:
            return self.insert_partition(partition)
        st = self._combiner.combine(cur.get(), partition.get())
        if isinstance(self._combiner, FnCombiner) and self._combiner._slot is not None:
            cur.set(self._combiner._slot)
            self._combiner._slot = None
        return st

    def insert_partition(self, partition: Partition[P]) -> PartitionStatus:
        self._partitions[partition.id()] = partition
        if self.slab is not None:
            self.slab = None  # no longer guaranteed to be slab-backed
        return PartitionStatus.ADDED

    def get_partition(self, partition_id: int) -> Optional[Partition[P]]:
        return self._partitions.get(int(partition_id))

    def remove_partition(self, partition_id: int) -> Optional[Partition[P]]:
        p = self._partitions.pop(int(partition_id), None)
        if p is not None:
            self.slab = None
        return p

    def is_empty(self) -> bool:
        return not self._partitions

    def release(self) -> None:
        for p in self._partitions.values():
            p.release()
        self._partitions.clear()
        self.slab = None

    def free(self) -> None:
        self.release()

    # -- camelCase aliases (reference spelling) -------------------------------------
    getTableID = get_table_id
    getCombiner = get_combiner
    getNumPartitions = get_num_partitions
    getPartitionIDs = get_partition_ids
    getPartitions = get_partitions
    addPartition = add_partition
    insertPartition = insert_partition
    getPartition = get_partition
    removePartition = remove_partition
    isEmpty = is_empty

    def __iter__(self) -> Iterator[Partition[P]]:
        return iter(list(self._partitions.values()))

    def __len__(self) -> int:
        return len(self._partitions)

    def __contains__(self, pid: int) -> bool:
        return int(pid) in self._partitions

    def __repr__(self) -> str:
        return f"Table(id={self._table_id}, partitions={sorted(self._partitions)})"


def add_partitions_to_table(table: Table, partitions: Iterable[Partition]) -> bool:
    """``partition/PartitionUtil.java:92-130`` — insert unseen ids, combine seen ones."""
    ok = True
    for p in partitions:
        st = table.add_partition(p)
        if st in (PartitionStatus.ADD_FAILED, PartitionStatus.COMBINE_FAILED):
            ok = False
    return ok
