"""Symmetric (peer-mapped) device workspace + the in-kernel NVLink collectives.

Harp's transport layer (``core/harp-collective/.../io``, ``client``, ``server``:
blocking TCP sockets, thread-per-connection server, per-element big-endian
serialisation) is replaced by ONE slab of device memory per rank that every
peer GPU maps into its own address space.  Collectives are kernels
(``csrc/coll.cu``) that load/store/reduce peer slabs directly over NVLink and
synchronise with release/acquire flags that live in the slab.

Slab sources, in order of preference:

1. ``torch.distributed._symmetric_memory`` (CUDA VMM + fabric/fd handle exchange);
2. legacy CUDA IPC on a ``cudaMalloc`` slab owned by ``libharp_b200.so``.

All ranks must create workspaces, allocate from them and call collectives in the
same order (allocation is a deterministic bump pointer, so offsets are identical
everywhere — "symmetric addressing").
"""
from __future__ import annotations

import ctypes
from typing import List

import torch
import torch.distributed as dist

from .. import _native

_DT = {torch.float32: 0, torch.float64: 1, torch.int32: 2, torch.int64: 3}
_OPS = {"sum": 0, "minus": 1, "prod": 2, "max": 3, "min": 4}


def _declare(lib) -> None:
    if getattr(lib, "_hb_coll_declared", False):
        return
    P, S, I, U = ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int, ctypes.c_uint32
    lib.hb_coll_flags_bytes.restype = S
    lib.hb_coll_reduce.argtypes = [P, I, I, S, S, I, I, I, I, S, U, P]
    lib.hb_coll_move.argtypes = [P, I, I, S, S, S, I, I, S, U, P]
    lib.hb_coll_barrier.argtypes = [P, I, I, S, U, P]
    lib.hb_kmeans_regroup_allgather.argtypes = [P, I, I, S, S, I, I, I, S, U, P]
    lib.hb_ipc_export.argtypes = [P, P]
    lib.hb_ipc_open.argtypes = [P, ctypes.POINTER(P)]
    lib.hb_ipc_close.argtypes = [P]
    lib.hb_slab_alloc.argtypes = [S, ctypes.POINTER(P)]
    lib.hb_slab_free.argtypes = [P]
    lib._hb_coll_declared = True


class SymmetricWorkspace:
    """A peer-mapped slab with a bump allocator and in-kernel collectives on its contents."""

    def __init__(self, device: torch.device, bytes_: int = 64 << 20, group=None, backend: str = "auto"):
        assert dist.is_initialized(), "SymmetricWorkspace needs an initialised process group"
        self.lib = _native.lib()
        _declare(self.lib)
        self.device = torch.device(device)
        self.group = group if group is not None else dist.group.WORLD
        self.rank = dist.get_rank(self.group)
        self.world = dist.get_world_size(self.group)
        if self.world > 16:
            raise ValueError("SymmetricWorkspace supports at most 16 ranks (one NVSwitch domain)")
        self.flags_bytes = int(self.lib.hb_coll_flags_bytes())
        self.data_bytes = (int(bytes_) + 255) // 256 * 256
        self.total_bytes = self.data_bytes + self.flags_bytes
        self.flags_offset = self.data_bytes
        self._bump = 0
        self._ticket = 2
        self._ipc_opened: List[int] = []
        self._owned_slab = None
        self.multicast_ptr = 0
        self.backend = None
        errors = []
        for be in (["torch", "ipc"] if backend == "auto" else [backend]):
            try:
                if be == "torch":
                    self._init_torch()
                else:
                    self._init_ipc()
                self.backend = be
                break
            except Exception as e:  # noqa: BLE001
                errors.append(f"{be}: {e!r}")
        # every rank must agree on the backend
        ok = torch.tensor([1 if self.backend is not None else 0], device=self.device)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=self.group)
        if int(ok.item()) == 0:
            raise RuntimeError("no symmetric-memory backend available: " + "; ".join(errors))
        self._ptr_array = (ctypes.c_void_p * 16)(*([ctypes.c_void_p(p) for p in self.peer_ptrs] +
                                                   [ctypes.c_void_p(0)] * (16 - self.world)))
        # zero the flags once, then make sure nobody starts before everyone has zeroed
        self.slab[self.flags_offset:].zero_()
        torch.cuda.synchronize(self.device)
        dist.barrier(group=self.group)

    # ------------------------------------------------------------------ backends
    def _init_torch(self) -> None:
        import torch.distributed._symmetric_memory as symm_mem

        t = symm_mem.empty(self.total_bytes, dtype=torch.uint8, device=self.device)
        hdl = symm_mem.rendezvous(t, self.group)
        self.slab = t
        self.peer_ptrs = [int(p) for p in hdl.buffer_ptrs]
        assert len(self.peer_ptrs) == self.world
        assert self.peer_ptrs[self.rank] == t.data_ptr()
        self._hdl = hdl
        try:
            self.multicast_ptr = int(hdl.multicast_ptr) if hdl.has_multicast_support(
                self.device.type, self.device.index) else 0
        except Exception:  # noqa: BLE001
            self.multicast_ptr = 0

    def _init_ipc(self) -> None:
        # a plain torch allocation (cudaMalloc-backed block); export handle + offset of the block
        t = torch.zeros(self.total_bytes, dtype=torch.uint8, device=self.device)
        st = t.untyped_storage()
        info = st._share_cuda_()
        # (device, handle, storage_size_bytes, storage_offset_bytes, ref_counter_handle, ref_counter_offset, event_handle, event_sync_required)
        handle, offset = bytes(info[1]), int(info[3])
        gathered: List = [None] * self.world
        dist.all_gather_object(gathered, (handle, offset), group=self.group)
        ptrs = []
        for r, (h, off) in enumerate(gathered):
            if r == self.rank:
                ptrs.append(t.data_ptr())
                continue
            out = ctypes.c_void_p(0)
            buf = ctypes.create_string_buffer(h, 64)
            _native.check(self.lib.hb_ipc_open(buf, ctypes.byref(out)), "hb_ipc_open")
            self._ipc_opened.append(out.value)
            ptrs.append(out.value + off)
        self.slab = t
        self.peer_ptrs = ptrs

    # ------------------------------------------------------------------ allocation
    def alloc(self, numel: int, dtype: torch.dtype = torch.float32) -> torch.Tensor:
        """Symmetric allocation: same offset on every rank (call in the same order everywhere)."""
        esz = torch.empty((), dtype=dtype).element_size()
        nbytes = numel * esz
        start = (self._bump + 255) // 256 * 256
        if start + nbytes > self.data_bytes:
            raise MemoryError(f"symmetric workspace exhausted ({start + nbytes} > {self.data_bytes})")
        self._bump = start + nbytes
        return self.slab[start:start + nbytes].view(dtype)

    def reset(self) -> None:
        self._bump = 0

    def offset_of(self, t: torch.Tensor) -> int:
        off = t.data_ptr() - self.slab.data_ptr()
        if off < 0 or off + t.numel() * t.element_size() > self.data_bytes:
            raise ValueError("tensor does not live in this symmetric workspace")
        return off

    def contains(self, t: torch.Tensor) -> bool:
        off = t.data_ptr() - self.slab.data_ptr()
        return 0 <= off and off + t.numel() * t.element_size() <= self.data_bytes

    def _next_ticket(self) -> int:
        t = self._ticket
        self._ticket += 2
        return t & 0x7FFFFFFF

    # ------------------------------------------------------------------ collectives
    def barrier(self) -> None:
        rc = self.lib.hb_coll_barrier(self._ptr_array, self.rank, self.world, self.flags_offset,
                                      self._next_ticket(), _native.stream_ptr())
        _native.check(rc, "hb_coll_barrier")
        _native.count_launch()

    def _reduce(self, t: torch.Tensor, op: str, mode: int, root: int = 0) -> torch.Tensor:
        assert t.is_contiguous() and t.dtype in _DT, f"unsupported dtype {t.dtype}"
        rc = self.lib.hb_coll_reduce(self._ptr_array, self.rank, self.world, self.offset_of(t),
                                     t.numel(), _DT[t.dtype], _OPS[op], mode, root, self.flags_offset,
                                     self._next_ticket(), _native.stream_ptr())
        _native.check(rc, "hb_coll_reduce")
        _native.count_launch()
        return t

    def allreduce_(self, t: torch.Tensor, op: str = "sum") -> torch.Tensor:
        """In-place allreduce of a tensor living at the same offset on every rank."""
        return self._reduce(t, op, 0)

    def allreduce_sum_(self, t: torch.Tensor) -> torch.Tensor:
        return self._reduce(t, "sum", 0)

    def reduce_scatter_(self, t: torch.Tensor, op: str = "sum") -> torch.Tensor:
        """Rank r ends up with the reduced r-th 1/world slice (in place); returns that slice view."""
        self._reduce(t, op, 1)
        vn = 16 // t.element_size()
        per = ((t.numel() + self.world - 1) // self.world + vn - 1) // vn * vn
        lo = min(t.numel(), self.rank * per)
        hi = min(t.numel(), lo + per)
        return t.view(-1)[lo:hi]

    def reduce_(self, t: torch.Tensor, root: int, op: str = "sum") -> torch.Tensor:
        return self._reduce(t, op, 2, root)

    def allgather_(self, t: torch.Tensor) -> torch.Tensor:
        """``t`` has world equal shards along dim 0; shard ``rank`` is valid on entry, all on exit."""
        assert t.is_contiguous()
        nbytes = t.numel() * t.element_size()
        assert nbytes % self.world == 0
        rc = self.lib.hb_coll_move(self._ptr_array, self.rank, self.world, self.offset_of(t), 0,
                                   nbytes // self.world, 0, 0, self.flags_offset, self._next_ticket(),
                                   _native.stream_ptr())
        _native.check(rc, "hb_coll_move(allgather)")
        _native.count_launch()
        return t

    def broadcast_(self, t: torch.Tensor, root: int) -> torch.Tensor:
        rc = self.lib.hb_coll_move(self._ptr_array, self.rank, self.world, self.offset_of(t), 0,
                                   t.numel() * t.element_size(), 1, root, self.flags_offset,
                                   self._next_ticket(), _native.stream_ptr())
        _native.check(rc, "hb_coll_move(broadcast)")
        _native.count_launch()
        return t

    def rotate_(self, dst: torch.Tensor, src: torch.Tensor, src_rank: int) -> torch.Tensor:
        """``dst`` (local) ← ``src`` as it exists on ``src_rank`` (ring rotation / arbitrary map)."""
        nbytes = src.numel() * src.element_size()
        assert dst.numel() * dst.element_size() >= nbytes
        rc = self.lib.hb_coll_move(self._ptr_array, self.rank, self.world, self.offset_of(dst),
                                   self.offset_of(src), nbytes, 2, int(src_rank), self.flags_offset,
                                   self._next_ticket(), _native.stream_ptr())
        _native.check(rc, "hb_coll_move(rotate)")
        _native.count_launch()
        return dst

    def kmeans_regroup_allgather_(self, acc: torch.Tensor, C: torch.Tensor, k: int, d: int,
                                  keep_empty: bool = False) -> torch.Tensor:
        """Fused regroup → divide → allgather of K-means partial sums (one kernel per rank)."""
        rc = self.lib.hb_kmeans_regroup_allgather(
            self._ptr_array, self.rank, self.world, self.offset_of(acc), self.offset_of(C), k, d,
            1 if keep_empty else 0, self.flags_offset, self._next_ticket(), _native.stream_ptr())
        _native.check(rc, "hb_kmeans_regroup_allgather")
        _native.count_launch()
        return C

    def close(self) -> None:
        for p in self._ipc_opened:
            self.lib.hb_ipc_close(ctypes.c_void_p(p))
        self._ipc_opened = []
