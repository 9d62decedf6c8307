"""Typed arrays, Writable payloads and the resource pools.

Reference (``core/harp-collective/src/main/java/edu/iu/harp/resource``):
``Array.java:23-75`` (``(array, start, size)`` triple), ``DoubleArray.java:43-104``
(``create(len, approximate)`` pulls from a pool, ``release`` returns, ``free`` drops),
``ArrayPool.java:73-228`` (power-of-two size classes when ``approximate``),
``WritablePool.java:76-159``, ``ResourcePool.java:19-104`` (singleton, ``clean``/``log``),
``Writable.java:30-153`` (user-defined serialisable payloads, pooled by class).

Here an array is a torch tensor view handed out by a pool of torch storages
(device memory ultimately comes from PyTorch's caching allocator; pinned host
staging buffers come from the C++ pool in ``csrc/host/pool.cpp``).  The wire
format of the Java transport is not reproduced (nothing is persisted in it), but
``Writable`` keeps ``write``/``read`` so custom payloads can cross ranks through
the generic (object) collective path.
"""
from __future__ import annotations

import io
import pickle
import struct
import threading
from collections import defaultdict
from typing import Any, Dict, List, Optional, Tuple, Type

import torch

# DataType tags of the reference wire format (io/DataType.java:24-33) — kept for introspection
DATA_TYPE_TAGS = {torch.uint8: 1, torch.int16: 2, torch.int32: 3, torch.float32: 4,
                  torch.int64: 5, torch.float64: 6}
WRITABLE_TAG = 7


def _round_pow2(n: int) -> int:
    """ArrayPool.java:73-84 — smallest power of two >= n (n >= 1)."""
    if n <= 1:
        return 1
    return 1 << (int(n) - 1).bit_length()


class ArrayPool:
    """Free-lists of tensors keyed by (device, size class) for one dtype (``ArrayPool.java``)."""

    def __init__(self, dtype: torch.dtype):
        self.dtype = dtype
        self._free: Dict[Tuple[str, int], List[torch.Tensor]] = defaultdict(list)
        self._in_use: Dict[int, Tuple[str, int]] = {}
        self._lock = threading.Lock()

    def get_array(self, size: int, approximate: bool, device="cpu", pinned: bool = False) -> torch.Tensor:
        cap = _round_pow2(size) if approximate else int(size)
        key = (f"{device}{'/pinned' if pinned else ''}", cap)
        with self._lock:
            lst = self._free.get(key)
            buf = lst.pop() if lst else None
        if buf is None:
            buf = torch.empty(cap, dtype=self.dtype, device=device, pin_memory=pinned and str(device) == "cpu")
        with self._lock:
            self._in_use[buf.data_ptr() if cap else id(buf)] = key
        return buf

    def release_array(self, buf: torch.Tensor) -> bool:
        k = buf.data_ptr() if buf.numel() else id(buf)
        with self._lock:
            key = self._in_use.pop(k, None)
            if key is None:
                return False
            self._free[key].append(buf)
        return True

    def free_array(self, buf: torch.Tensor) -> bool:
        k = buf.data_ptr() if buf.numel() else id(buf)
        with self._lock:
            return self._in_use.pop(k, None) is not None

    def clean(self) -> None:
        with self._lock:
            self._free.clear()

    def stats(self) -> Dict[str, int]:
        with self._lock:
            cached = sum(len(v) for v in self._free.values())
            cached_elems = sum(k[1] * len(v) for k, v in self._free.items())
            return {"in_use": len(self._in_use), "cached": cached, "cached_elems": cached_elems}


class Writable:
    """User-defined payload (``resource/Writable.java:30-153``).  Subclasses implement
    ``write(out)``, ``read(inp)``, ``clear()`` and ``get_num_write_bytes()``."""

    def write(self, out: io.BufferedIOBase) -> None:  # pragma: no cover - interface
        raise NotImplementedError

    def read(self, inp: io.BufferedIOBase) -> None:  # pragma: no cover - interface
        raise NotImplementedError

    def clear(self) -> None:
        pass

    def get_num_write_bytes(self) -> int:
        buf = io.BytesIO()
        self.write(buf)
        return buf.tell()

    # pooled construction --------------------------------------------------------
    @classmethod
    def create(cls) -> "Writable":
        return ResourcePool.get().writable_pool.get_writable(cls)

    def release(self) -> None:
        ResourcePool.get().writable_pool.release_writable(self)

    def free(self) -> None:
        ResourcePool.get().writable_pool.free_writable(self)

    # serialisation used by the object collectives ------------------------------------
    def to_bytes(self) -> bytes:
        buf = io.BytesIO()
        self.write(buf)
        return buf.getvalue()

    @classmethod
    def from_bytes(cls, data: bytes) -> "Writable":
        obj = cls.create()
        obj.read(io.BytesIO(data))
        return obj


_WRITABLE_REGISTRY: Dict[str, Type[Writable]] = {}


def register_writable(cls: Type[Writable]) -> Type[Writable]:
    """Class decorator: makes a Writable reconstructible by name on the receiving rank."""
    _WRITABLE_REGISTRY[f"{cls.__module__}.{cls.__qualname__}"] = cls
    return cls


def writable_class(name: str) -> Type[Writable]:
    return _WRITABLE_REGISTRY[name]


def writable_name(obj: Writable) -> str:
    cls = type(obj)
    return f"{cls.__module__}.{cls.__qualname__}"


class WritablePool:
    """``resource/WritablePool.java:76-159`` — free-lists keyed by class."""

    def __init__(self):
        self._free: Dict[type, List[Writable]] = defaultdict(list)
        self._in_use: Dict[int, Writable] = {}
        self._lock = threading.Lock()

    def get_writable(self, cls: Type[Writable]) -> Writable:
        with self._lock:
            lst = self._free.get(cls)
            obj = lst.pop() if lst else None
        if obj is None:
            obj = cls()
        with self._lock:
            self._in_use[id(obj)] = obj
        return obj

    def release_writable(self, obj: Writable) -> bool:
        with self._lock:
            if self._in_use.pop(id(obj), None) is None:
                return False
        obj.clear()
        with self._lock:
            self._free[type(obj)].append(obj)
        return True

    def free_writable(self, obj: Writable) -> bool:
        with self._lock:
            return self._in_use.pop(id(obj), None) is not None

    def clean(self) -> None:
        with self._lock:
            self._free.clear()

    def stats(self) -> Dict[str, int]:
        with self._lock:
            return {"in_use": len(self._in_use), "cached": sum(len(v) for v in self._free.values())}


class ResourcePool:
    """Per-process singleton (``resource/ResourcePool.java:19-104``)."""

    _inst: Optional["ResourcePool"] = None
    _inst_lock = threading.Lock()

    def __init__(self):
        self.pools: Dict[torch.dtype, ArrayPool] = {dt: ArrayPool(dt) for dt in DATA_TYPE_TAGS}
        self.writable_pool = WritablePool()

    @classmethod
    def get(cls) -> "ResourcePool":
        if cls._inst is None:
            with cls._inst_lock:
                if cls._inst is None:
                    cls._inst = ResourcePool()
        return cls._inst

    def pool(self, dtype: torch.dtype) -> ArrayPool:
        if dtype not in self.pools:
            self.pools[dtype] = ArrayPool(dtype)
        return self.pools[dtype]

    def clean(self) -> None:
        for p in self.pools.values():
            p.clean()
        self.writable_pool.clean()

    def log(self) -> Dict[str, Dict[str, int]]:
        out = {str(dt).replace("torch.", ""): p.stats() for dt, p in self.pools.items()}
        out["writable"] = self.writable_pool.stats()
        return out


class Array:
    """``(array, start, size)`` view (``resource/Array.java:23-75``) over a pooled tensor."""

    dtype: torch.dtype = torch.float64

    def __init__(self, backing: torch.Tensor, start: int, size: int):
        self._backing = backing
        self._start = int(start)
        self._size = int(size)

    @classmethod
    def create(cls, length: int, approximate: bool = False, device="cpu", pinned: bool = False) -> "Array":
        buf = ResourcePool.get().pool(cls.dtype).get_array(length, approximate, device, pinned)
        return cls(buf, 0, length)

    @classmethod
    def wrap(cls, t: torch.Tensor) -> "Array":
        assert t.dtype == cls.dtype and t.dim() == 1
        return cls(t, 0, t.numel())

    def get(self) -> torch.Tensor:
        """The live elements ``backing[start:start+size]`` (a view, no copy)."""
        return self._backing[self._start:self._start + self._size]

    def backing(self) -> torch.Tensor:
        return self._backing

    def start(self) -> int:
        return self._start

    def size(self) -> int:
        return self._size

    def get_num_encode_bytes(self) -> int:
        # 1 type byte + int32 length + elements (resource/DoubleArray.java:44)
        return 5 + self._size * self._backing.element_size()

    def release(self) -> None:
        if self._backing is not None:
            ResourcePool.get().pool(self.dtype).release_array(self._backing)
        self._backing = None
        self._size = 0

    def free(self) -> None:
        if self._backing is not None:
            ResourcePool.get().pool(self.dtype).free_array(self._backing)
        self._backing = None
        self._size = 0


class ByteArray(Array):
    dtype = torch.uint8


class ShortArray(Array):
    dtype = torch.int16


class IntArray(Array):
    dtype = torch.int32


class FloatArray(Array):
    dtype = torch.float32


class LongArray(Array):
    dtype = torch.int64


class DoubleArray(Array):
    dtype = torch.float64


# ---------------------------------------------------------------------------
# Generic payload (de)serialisation for the object collective path
# ---------------------------------------------------------------------------
def encode_payload(obj: Any) -> Tuple[str, Any]:
    """Returns a picklable ``(kind, data)`` description of a partition payload."""
    if isinstance(obj, torch.Tensor):
        return ("tensor", obj.detach().cpu())
    if isinstance(obj, Array):
        return ("array:" + type(obj).__name__, obj.get().detach().cpu().clone())
    if isinstance(obj, Writable):
        return ("writable:" + writable_name(obj), obj.to_bytes())
    return ("pickle", pickle.dumps(obj))


_ARRAY_TYPES = {c.__name__: c for c in (ByteArray, ShortArray, IntArray, FloatArray, LongArray, DoubleArray)}


def decode_payload(desc: Tuple[str, Any], device=None) -> Any:
    kind, data = desc
    if kind == "tensor":
        return data.to(device) if device is not None else data
    if kind.startswith("array:"):
        cls = _ARRAY_TYPES[kind.split(":", 1)[1]]
        return cls.wrap(data.to(device) if device is not None else data)
    if kind.startswith("writable:"):
        cls = writable_class(kind.split(":", 1)[1])
        return cls.from_bytes(data)
    return pickle.loads(data)
