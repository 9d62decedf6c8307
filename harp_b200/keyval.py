"""Key-value tables (``core/harp-collective/.../keyval``).

Reference: a KV table is a ``Table`` whose partitions are open hash maps;
``Int2IntKVTable`` uses ``partitionID = key`` (``keyval/Int2IntKVTable.java:146-148``),
``Key2ValKVTable`` uses ``partitionID = key.hashCode()`` (``Key2ValKVTable.java:100-222``);
colliding keys are merged with a ``ValCombiner`` / ``Type{Int,Long,Double}Combiner``
(sum by default), and two partitions with the same id are merged key by key
(``Int2IntKVPartition.java:67-165``).

B200-first representation: a partition is a *sorted COO pair* ``(keys, vals)`` of
tensors; putting a batch of pairs is ``cat → sort → segment-reduce`` (device
friendly, no per-key hashing), and merging two partitions is the same operation.
Scalar ``add_key_val`` / ``get_val`` keep the reference API.
"""
from __future__ import annotations

import io
import struct
from typing import Any, Callable, Dict, Hashable, Iterator, List, Optional, Tuple

import torch

from .partition import Partition, PartitionCombiner, PartitionStatus, Table
from .resource import Writable, register_writable


class TypeCombiner:
    """``TypeIntCombiner`` / ``TypeLongCombiner`` / ``TypeDoubleCombiner``: scalar merge rule."""

    def __init__(self, op: str = "sum"):
        assert op in ("sum", "max", "min", "prod", "last")
        self.op = op

    def combine(self, cur, new):
        if self.op == "sum":
            return cur + new
        if self.op == "max":
            return max(cur, new)
        if self.op == "min":
            return min(cur, new)
        if self.op == "prod":
            return cur * new
        return new


TypeIntCombiner = TypeLongCombiner = TypeDoubleCombiner = TypeCombiner

_SEG = {"sum": "sum", "max": "amax", "min": "amin", "prod": "prod"}


def segment_combine(keys: torch.Tensor, vals: torch.Tensor, op: str = "sum") -> Tuple[torch.Tensor, torch.Tensor]:
    """Sort by key and merge equal keys — the device-side equivalent of hash-map put-with-combine."""
    if keys.numel() == 0:
        return keys, vals
    order = torch.argsort(keys, stable=True)
    k = keys[order]
    v = vals[order]
    uniq, inv = torch.unique_consecutive(k, return_inverse=True)
    if uniq.numel() == k.numel():
        return k, v
    if op == "last":
        # last occurrence wins
        idx = torch.zeros(uniq.numel(), dtype=torch.long, device=k.device)
        idx.scatter_(0, inv, torch.arange(k.numel(), device=k.device))  # later writes win for sorted stable order
        return uniq, v[idx]
    out_shape = (uniq.numel(),) + tuple(v.shape[1:])
    if op == "sum":
        out = torch.zeros(out_shape, dtype=v.dtype, device=v.device).index_add_(0, inv, v)
    else:
        out = torch.zeros(out_shape, dtype=v.dtype, device=v.device)
        idx = inv.view(-1, *([1] * (v.dim() - 1))).expand_as(v)
        out = out.scatter_reduce(0, idx, v, reduce=_SEG[op], include_self=False)
    return uniq, out


@register_writable
class KVPartition(Writable):
    """Sorted-COO map with typed keys/values (replaces ``Int2IntKVPartition`` & friends)."""

    key_dtype = torch.int64
    val_dtype = torch.float64
    default_return_val = 0

    def __init__(self, key_dtype=None, val_dtype=None, device="cpu"):
        if key_dtype is not None:
            self.key_dtype = key_dtype
        if val_dtype is not None:
            self.val_dtype = val_dtype
        self.device = torch.device(device)
        self.keys = torch.empty(0, dtype=self.key_dtype, device=self.device)
        self.vals = torch.empty(0, dtype=self.val_dtype, device=self.device)
        self._pending: List[Tuple[Any, Any]] = []

    def initialize(self) -> None:
        self.clear()

    def clear(self) -> None:
        self.keys = torch.empty(0, dtype=self.key_dtype, device=self.device)
        self.vals = torch.empty(0, dtype=self.val_dtype, device=self.device)
        self._pending = []

    # -- mutation --------------------------------------------------------------------
    def put_key_val(self, key, val, combiner: Optional[TypeCombiner] = None) -> None:
        self._pending.append((key, val))
        if len(self._pending) >= 4096:
            self._flush(combiner)

    def put_batch(self, keys: torch.Tensor, vals: torch.Tensor, combiner: Optional[TypeCombiner] = None) -> None:
        self._flush(combiner)
        op = combiner.op if combiner is not None else "sum"
        k = torch.cat([self.keys, keys.to(self.device, self.key_dtype)])
        v = torch.cat([self.vals, vals.to(self.device, self.val_dtype)])
        self.keys, self.vals = segment_combine(k, v, op)

    def _flush(self, combiner: Optional[TypeCombiner] = None) -> None:
        if not self._pending:
            return
        ks = torch.tensor([p[0] for p in self._pending], dtype=self.key_dtype, device=self.device)
        vs = torch.tensor([p[1] for p in self._pending], dtype=self.val_dtype, device=self.device)
        self._pending = []
        op = combiner.op if combiner is not None else "sum"
        k = torch.cat([self.keys, ks])
        v = torch.cat([self.vals, vs])
        self.keys, self.vals = segment_combine(k, v, op)

    def merge(self, other: "KVPartition", combiner: Optional[TypeCombiner] = None) -> None:
        other._flush(combiner)
        self.put_batch(other.keys, other.vals, combiner)

    # -- queries ----------------------------------------------------------------------
    def get_val(self, key):
        self._flush()
        if self.keys.numel() == 0:
            return self.default_return_val
        i = int(torch.searchsorted(self.keys, torch.tensor(key, dtype=self.key_dtype, device=self.device)))
        if i < self.keys.numel() and int(self.keys[i]) == int(key):
            return self.vals[i].item()
        return self.default_return_val

    def contains_key(self, key) -> bool:
        self._flush()
        if self.keys.numel() == 0:
            return False
        i = int(torch.searchsorted(self.keys, torch.tensor(key, dtype=self.key_dtype, device=self.device)))
        return i < self.keys.numel() and int(self.keys[i]) == int(key)

    def size(self) -> int:
        self._flush()
        return int(self.keys.numel())

    def is_empty(self) -> bool:
        return self.size() == 0

    def items(self) -> Iterator[Tuple[int, Any]]:
        self._flush()
        return zip(self.keys.tolist(), self.vals.tolist())

    def to(self, device) -> "KVPartition":
        self._flush()
        self.device = torch.device(device)
        self.keys = self.keys.to(self.device)
        self.vals = self.vals.to(self.device)
        return self

    # -- Writable ---------------------------------------------------------------------
    def write(self, out) -> None:
        self._flush()
        k = self.keys.cpu().contiguous().numpy()
        v = self.vals.cpu().contiguous().numpy()
        out.write(struct.pack("<q", k.shape[0]))
        out.write(str(k.dtype).encode().ljust(16))
        out.write(str(v.dtype).encode().ljust(16))
        out.write(k.tobytes())
        out.write(v.tobytes())

    def read(self, inp) -> None:
        import numpy as np

        (n,) = struct.unpack("<q", inp.read(8))
        kd = np.dtype(inp.read(16).decode().strip())
        vd = np.dtype(inp.read(16).decode().strip())
        self.keys = torch.from_numpy(np.frombuffer(inp.read(n * kd.itemsize), dtype=kd).copy()).to(self.device)
        self.vals = torch.from_numpy(np.frombuffer(inp.read(n * vd.itemsize), dtype=vd).copy()).to(self.device)
        self.key_dtype, self.val_dtype = self.keys.dtype, self.vals.dtype
        self._pending = []

    def get_num_write_bytes(self) -> int:
        self._flush()
        return 40 + self.keys.numel() * self.keys.element_size() + self.vals.numel() * self.vals.element_size()


class KVPartitionCombiner(PartitionCombiner[KVPartition]):
    """Merges two KV partitions key by key (``Int2IntKVTable.java:36-58``)."""

    def __init__(self, val_combiner: TypeCombiner):
        self.val_combiner = val_combiner

    def combine(self, cur: KVPartition, new: KVPartition) -> PartitionStatus:
        cur.merge(new, self.val_combiner)
        return PartitionStatus.COMBINED


class KVTable(Table[KVPartition]):
    """Base of the typed KV tables (``keyval/KVTable.java``)."""

    key_dtype = torch.int64
    val_dtype = torch.float64

    def __init__(self, table_id: int, combiner: Optional[TypeCombiner] = None, device="cpu",
                 num_partitions_hint: int = 0):
        self.val_combiner = combiner if combiner is not None else TypeCombiner("sum")
        super().__init__(table_id, KVPartitionCombiner(self.val_combiner))
        self.device = torch.device(device)
        #: when > 0, keys are folded onto this many partitions (key % hint) instead of 1 per key
        self.num_partitions_hint = int(num_partitions_hint)

    def get_kv_partition_id(self, key) -> int:
        k = int(key)
        return k % self.num_partitions_hint if self.num_partitions_hint > 0 else k

    def _get_or_create(self, key) -> KVPartition:
        pid = self.get_kv_partition_id(key)
        p = self.get_partition(pid)
        if p is None:
            part = KVPartition(self.key_dtype, self.val_dtype, self.device)
            self.insert_partition(Partition(pid, part))
            return part
        return p.get()

    def add_key_val(self, key, val) -> None:
        self._get_or_create(key).put_key_val(key, val, self.val_combiner)

    def add_batch(self, keys: torch.Tensor, vals: torch.Tensor) -> None:
        """Vectorised insert of many pairs."""
        keys = keys.to(self.device)
        vals = vals.to(self.device)
        if self.num_partitions_hint > 0:
            pids = keys.to(torch.int64).abs().remainder(self.num_partitions_hint)
        else:
            pids = keys.to(torch.int64)
        for pid in torch.unique(pids).tolist():
            m = pids == pid
            p = self.get_partition(pid)
            if p is None:
                part = KVPartition(self.key_dtype, self.val_dtype, self.device)
                self.insert_partition(Partition(pid, part))
            else:
                part = p.get()
            part.put_batch(keys[m], vals[m], self.val_combiner)

    def get_val(self, key):
        p = self.get_partition(self.get_kv_partition_id(key))
        return p.get().get_val(key) if p is not None else KVPartition.default_return_val

    def remove_val(self, key):
        p = self.get_partition(self.get_kv_partition_id(key))
        if p is None:
            return None
        part = p.get()
        part._flush()
        m = part.keys != int(key)
        old = part.get_val(key)
        part.keys, part.vals = part.keys[m], part.vals[m]
        return old

    def items(self) -> Iterator[Tuple[int, Any]]:
        for p in self.get_partitions():
            yield from p.get().items()

    def num_keys(self) -> int:
        return sum(p.get().size() for p in self.get_partitions())

    addKeyVal = add_key_val
    getVal = get_val
    removeVal = remove_val


class Int2IntKVTable(KVTable):
    key_dtype, val_dtype = torch.int32, torch.int32


class Int2LongKVTable(KVTable):
    key_dtype, val_dtype = torch.int32, torch.int64


class Long2IntKVTable(KVTable):
    key_dtype, val_dtype = torch.int64, torch.int32


class Long2DoubleKVTable(KVTable):
    key_dtype, val_dtype = torch.int64, torch.float64


class Int2DoubleKVTable(KVTable):
    key_dtype, val_dtype = torch.int32, torch.float64


# ---------------------------------------------------------------------------
# Generic object-keyed tables (Int2ValKVTable / Key2ValKVTable)
# ---------------------------------------------------------------------------
@register_writable
class ObjKVPartition(Writable):
    """``Key2ValKVPartition`` / ``Int2ValKVPartition``: python-dict backed, arbitrary values."""

    def __init__(self):
        self.map: Dict[Hashable, Any] = {}

    def initialize(self) -> None:
        self.map = {}

    def clear(self) -> None:
        self.map = {}

    def put_key_val(self, key, val, combiner: Optional[Callable[[Any, Any], Any]]) -> None:
        if key in self.map and combiner is not None:
            self.map[key] = combiner(self.map[key], val)
        else:
            self.map[key] = val

    def get_val(self, key):
        return self.map.get(key)

    def remove_val(self, key):
        return self.map.pop(key, None)

    def size(self) -> int:
        return len(self.map)

    def items(self):
        return self.map.items()

    def write(self, out) -> None:
        import pickle

        data = pickle.dumps(self.map)
        out.write(struct.pack("<q", len(data)))
        out.write(data)

    def read(self, inp) -> None:
        import pickle

        (n,) = struct.unpack("<q", inp.read(8))
        self.map = pickle.loads(inp.read(n))


class ObjKVPartitionCombiner(PartitionCombiner[ObjKVPartition]):
    def __init__(self, val_combiner: Optional[Callable[[Any, Any], Any]]):
        self.val_combiner = val_combiner

    def combine(self, cur: ObjKVPartition, new: ObjKVPartition) -> PartitionStatus:
        for k, v in new.items():
            cur.put_key_val(k, v, self.val_combiner)
        return PartitionStatus.COMBINED


def _stable_hash(key: Hashable) -> int:
    """Java-like deterministic hash (python's str hash is salted per process)."""
    if isinstance(key, int):
        return key & 0x7FFFFFFF
    if isinstance(key, str):
        h = 0
        for ch in key:
            h = (31 * h + ord(ch)) & 0xFFFFFFFF  # String.hashCode
        return h & 0x7FFFFFFF
    if isinstance(key, tuple):
        h = 1
        for x in key:
            h = (31 * h + _stable_hash(x)) & 0xFFFFFFFF
        return h & 0x7FFFFFFF
    return _stable_hash(repr(key))


class Key2ValKVTable(Table[ObjKVPartition]):
    """``keyval/Key2ValKVTable.java:100-222`` — ``partitionID = hash(key) % numPartitions``."""

    def __init__(self, table_id: int, val_combiner: Optional[Callable[[Any, Any], Any]] = None,
                 num_partitions: int = 64):
        super().__init__(table_id, ObjKVPartitionCombiner(val_combiner))
        self.val_combiner = val_combiner
        self.num_partitions = int(num_partitions)

    def get_kv_partition_id(self, key) -> int:
        return _stable_hash(key) % self.num_partitions

    def add_key_val(self, key, val) -> None:
        pid = self.get_kv_partition_id(key)
        p = self.get_partition(pid)
        if p is None:
            part = ObjKVPartition()
            self.insert_partition(Partition(pid, part))
        else:
            part = p.get()
        part.put_key_val(key, val, self.val_combiner)

    def get_val(self, key):
        p = self.get_partition(self.get_kv_partition_id(key))
        return p.get().get_val(key) if p is not None else None

    def remove_val(self, key):
        p = self.get_partition(self.get_kv_partition_id(key))
        return p.get().remove_val(key) if p is not None else None

    def items(self):
        for p in self.get_partitions():
            yield from p.get().items()

    addKeyVal = add_key_val
    getVal = get_val
    removeVal = remove_val


class Int2ValKVTable(Key2ValKVTable):
    """``keyval/Int2ValKVTable.java`` — int keys, arbitrary values, ``partitionID = key``-style."""

    def get_kv_partition_id(self, key) -> int:
        return int(key) % self.num_partitions
