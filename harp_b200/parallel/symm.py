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

All ranks must create workspaces, allocate / free from them and call collectives in the
same order: the allocator (first-fit free list with coalescing, 256-byte granules) is
deterministic, so offsets are identical everywhere — "symmetric addressing".

Rules the caller must keep (also in INTEGRATION.md):

* ONE stream per workspace: all collectives of a workspace share one flag array and one ticket counter, so they
  must be issued on the same CUDA stream (the first one used is remembered and later calls assert it).
* Waits for a peer inside the kernels are bounded by ``HB_WAIT_TIMEOUT_S`` (default 1800 s, the reference's
  ``DATA_MAX_WAIT_TIME``; 0 = for ever).  An expired wait does not trap: the kernel records it and skips its data
  phase; ``check()`` raises.  Construct engines / tables behind a ``dist.barrier`` so that a slow rank's set-up is
  not charged to the first collective.
"""
from __future__ import annotations

import ctypes
from typing import List, Optional

import torch
import torch.distributed as dist

from .. import _native

_DT = {torch.float32: 0, torch.float64: 1, torch.int32: 2, torch.int64: 3, torch.int16: 4, torch.int8: 5}
_OPS = {"sum": 0, "minus": 1, "prod": 2, "max": 3, "min": 4}


def _declare(lib) -> None:
    if getattr(lib, "_hb_coll_declared", False):
        return
    P, S, I, U = ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int, ctypes.c_uint32
    lib.hb_coll_flags_bytes.restype = S
    lib.hb_coll_reduce.argtypes = [P, I, I, S, S, I, I, I, I, S, U, P]
    lib.hb_coll_move.argtypes = [P, I, I, S, S, S, I, I, S, U, P]
    lib.hb_coll_barrier.argtypes = [P, I, I, S, U, P]
    lib.hb_coll_rows.argtypes = [P, I, I, S, S, ctypes.c_longlong, ctypes.c_longlong, I, I, S, U, P]
    lib.hb_coll_set_timeout_ns.argtypes = [ctypes.c_ulonglong]
    lib.hb_coll_nvls.argtypes = [P, P, I, I, S, S, I, I, S, U, P]
    ULLP, IP = ctypes.POINTER(ctypes.c_ulonglong), ctypes.POINTER(ctypes.c_int)
    lib.hb_vmm_granularity.argtypes = [I, ctypes.POINTER(S)]
    lib.hb_vmm_create.argtypes = [S, ULLP, IP]
    lib.hb_vmm_import.argtypes = [I, ULLP]
    lib.hb_vmm_map.argtypes = [ctypes.c_ulonglong, S, S, ctypes.POINTER(P)]
    lib.hb_vmm_unmap.argtypes = [P, S]
    lib.hb_vmm_release.argtypes = [ctypes.c_ulonglong]
    lib.hb_mc_create.argtypes = [I, S, ULLP, IP]
    lib.hb_mc_add_device.argtypes = [ctypes.c_ulonglong]
    lib.hb_mc_bind.argtypes = [ctypes.c_ulonglong, ctypes.c_ulonglong, S]
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
        self._free = [(0, self.data_bytes)]     # sorted (offset, size) holes
        self._live = {}                         # offset -> size of live allocations
        self._ticket = 2
        self._stream = None
        import os
        self.lib.hb_coll_set_timeout_ns(int(float(os.environ.get("HB_WAIT_TIMEOUT_S", "1800")) * 1e9))
        self._ipc_opened: List[int] = []
        self._owned_slab = None
        self.multicast_ptr = 0
        self.backend = None
        errors = []
        import os
        order = ["vmm", "torch", "ipc"] if os.environ.get("HB_SYMM_VMM", "1") != "0" else ["torch", "ipc"]
        for be in (order if backend == "auto" else [backend]):
            try:
                if be == "vmm":
                    self._init_vmm()
                elif be == "torch":
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

    def _init_vmm(self) -> None:
        """Own slab: cuMemCreate physical memory on every rank, POSIX-fd handles passed between the processes over a
        unix socket (SCM_RIGHTS), every peer's memory mapped locally, and the whole slab bound to ONE NVSwitch
        multicast object whose mapping (``multicast_ptr``) the NVLS kernels use (``csrc/nvls.cu``)."""
        import array
        import os
        import socket

        lib, world, rank = self.lib, self.world, self.rank
        gran = ctypes.c_size_t(0)
        _native.check(lib.hb_vmm_granularity(world, ctypes.byref(gran)), "hb_vmm_granularity")
        g = int(gran.value)
        size = (self.total_bytes + g - 1) // g * g
        handle, fd = ctypes.c_ulonglong(0), ctypes.c_int(-1)
        _native.check(lib.hb_vmm_create(size, ctypes.byref(handle), ctypes.byref(fd)), "hb_vmm_create")
        mc_handle, mc_fd = ctypes.c_ulonglong(0), ctypes.c_int(-1)
        if rank == 0 and world > 1:
            _native.check(lib.hb_mc_create(world, size, ctypes.byref(mc_handle), ctypes.byref(mc_fd)), "hb_mc_create")
        # ---- fd exchange: every rank listens on an abstract unix socket, sends its fd(s) to every peer
        SymmetricWorkspace._vmm_seq = getattr(SymmetricWorkspace, "_vmm_seq", 0) + 1       # same count on every rank
        tag = f"{os.environ.get('MASTER_PORT', '0')}_{os.getppid()}_{SymmetricWorkspace._vmm_seq}"
        names = [f"\0hb_vmm_{tag}_{r}" for r in range(world)]
        srv = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
        srv.bind(names[rank]); srv.listen(world)
        dist.barrier(group=self.group)
        for r in range(world):
            if r == rank:
                continue
            c = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
            c.connect(names[r])
            fds = [fd.value] + ([mc_fd.value] if rank == 0 else [])
            socket.send_fds(c, [rank.to_bytes(4, "little")], fds)
            c.close()
        peer_fd, got_mc = {}, None
        for _ in range(world - 1):
            conn, _a = srv.accept()
            msg, fds, _f, _ad = socket.recv_fds(conn, 4, 2)
            src = int.from_bytes(msg, "little")
            peer_fd[src] = fds[0]
            if src == 0 and len(fds) > 1:
                got_mc = fds[1]
            conn.close()
        srv.close()
        # ---- map everybody's memory; bind ours to the multicast object; map the multicast object
        ptrs, self._vmm = [], {"size": size, "handles": [], "maps": []}
        for r in range(world):
            h = handle if r == rank else ctypes.c_ulonglong(0)
            if r != rank:
                _native.check(lib.hb_vmm_import(peer_fd[r], ctypes.byref(h)), "hb_vmm_import")
                os.close(peer_fd[r])
            out = ctypes.c_void_p(0)
            _native.check(lib.hb_vmm_map(h.value, size, g, ctypes.byref(out)), "hb_vmm_map")
            ptrs.append(int(out.value))
            self._vmm["handles"].append(h.value); self._vmm["maps"].append(int(out.value))
        os.close(fd.value)
        self.multicast_ptr = 0
        if world > 1:
            if rank != 0:
                _native.check(lib.hb_vmm_import(got_mc, ctypes.byref(mc_handle)), "hb_vmm_import(multicast)")
                os.close(got_mc)
            else:
                os.close(mc_fd.value)
            _native.check(lib.hb_mc_add_device(mc_handle.value), "hb_mc_add_device")
            dist.barrier(group=self.group)                      # every device added before anybody binds
            _native.check(lib.hb_mc_bind(mc_handle.value, handle.value, size), "hb_mc_bind")
            dist.barrier(group=self.group)
            out = ctypes.c_void_p(0)
            _native.check(lib.hb_vmm_map(mc_handle.value, size, g, ctypes.byref(out)), "hb_vmm_map(multicast)")
            self.multicast_ptr = int(out.value)
            self._vmm["mc_handle"] = mc_handle.value

        class _Raw:                # torch tensor over the mapping (kept alive by this workspace)
            def __init__(self, ptr, nbytes):
                self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 3}
        self.slab = torch.as_tensor(_Raw(ptrs[rank], self.total_bytes), device=self.device)
        self.slab.zero_()
        self.peer_ptrs = ptrs

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
        """Symmetric allocation: same offset on every rank (call alloc / free in the same order everywhere).
        First fit over the free list, 256-byte granules."""
        esz = torch.empty((), dtype=dtype).element_size()
        nbytes = numel * esz
        need = max(256, (nbytes + 255) // 256 * 256)
        for i, (off, size) in enumerate(self._free):
            if size >= need:
                if size == need:
                    self._free.pop(i)
                else:
                    self._free[i] = (off + need, size - need)
                self._live[off] = need
                return self.slab[off:off + nbytes].view(dtype)
        raise MemoryError(f"symmetric workspace exhausted: {need} bytes wanted, largest hole "
                          f"{max([sz for _, sz in self._free], default=0)} of {self.data_bytes}")

    def free(self, t: torch.Tensor) -> None:
        """Returns an allocation (the tensor ``alloc`` returned, or any view starting at its first byte) to the
        free list; neighbouring holes are merged.  Every rank must free in the same order it would allocate."""
        off = t.data_ptr() - self.slab.data_ptr()
        size = self._live.pop(off, None)
        if size is None:
            raise ValueError("not a live allocation of this symmetric workspace")
        import bisect
        i = bisect.bisect_left(self._free, (off, 0))
        self._free.insert(i, (off, size))
        if i + 1 < len(self._free) and self._free[i][0] + self._free[i][1] == self._free[i + 1][0]:
            self._free[i] = (self._free[i][0], self._free[i][1] + self._free[i + 1][1])
            self._free.pop(i + 1)
        if i > 0 and self._free[i - 1][0] + self._free[i - 1][1] == self._free[i][0]:
            self._free[i - 1] = (self._free[i - 1][0], self._free[i - 1][1] + self._free[i][1])
            self._free.pop(i)

    def reset(self) -> None:
        self._free = [(0, self.data_bytes)]
        self._live = {}

    def bytes_in_use(self) -> int:
        return sum(self._live.values())

    def offset_of(self, t: torch.Tensor) -> int:
        off = t.data_ptr() - self.slab.data_ptr()
        if off < 0 or off + t.numel() * t.element_size() > self.data_bytes:
            raise ValueError("tensor does not live in this symmetric workspace")
        return off

    def contains(self, t: torch.Tensor) -> bool:
        off = t.data_ptr() - self.slab.data_ptr()
        return 0 <= off and off + t.numel() * t.element_size() <= self.data_bytes

    def _next_ticket(self) -> int:
        # natural 32-bit wrap: the kernels compare with a signed difference, which is wrap-safe only that way
        t = self._ticket
        self._ticket += 2
        cur = torch.cuda.current_stream(self.device).cuda_stream
        if self._stream is None:
            self._stream = cur
        elif self._stream != cur:
            raise RuntimeError("SymmetricWorkspace collectives must all be issued on one CUDA stream (flags and "
                               "tickets are shared); this call is on a different stream than the first one")
        return t & 0xFFFFFFFF

    def rebind_stream(self) -> None:
        """Forget the remembered stream (after a synchronize, when the owner moves its collectives elsewhere)."""
        self._stream = None

    def check(self) -> None:
        """Raises if an in-kernel wait for a peer expired since the last check (call after a synchronize)."""
        w = self.slab[self.flags_offset + self.flags_bytes - 256: self.flags_offset + self.flags_bytes - 248].view(torch.int32).cpu()
        if int(w[0]) != 0:
            tick, peer = int(w[0]) & 0x7FFFFFFF, int(w[1])
            self.slab[self.flags_offset + self.flags_bytes - 256: self.flags_offset + self.flags_bytes - 248].zero_()
            raise _native.NativeError(f"NVLink collective (ticket {tick}) gave up waiting for rank {peer} after "
                                      f"HB_WAIT_TIMEOUT_S; the ranks' collective call order diverged or a rank died")

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

    def allreduce_(self, t: torch.Tensor, op: str = "sum", deterministic: bool = False) -> torch.Tensor:
        """In-place allreduce of a tensor living at the same offset on every rank.  float32 sums take the NVLS path
        (in-switch reduction) when the slab is multicast-bound and at least four ranks take part (measured: 5 MB at 2
        ranks 42.7 us NVLS vs 32.6 us peer loads; at 8 ranks NVLS is ahead at every size, 671 vs 537 GB/s at 64 MB);
        ``deterministic=True`` keeps the rank-ordered peer-load kernel, whose result is bit-identical on every rank and
        from run to run."""
        if (op == "sum" and t.dtype == torch.float32 and self.has_nvls and not deterministic and self.world >= 4
                and (t.numel() * 4) % 16 == 0 and self.offset_of(t) % 16 == 0):
            return self._nvls(t, 0)
        return self._reduce(t, op, 0)

    def allreduce_sum_(self, t: torch.Tensor) -> torch.Tensor:
        return self._reduce(t, "sum", 0)

    # ---- NVLS data path (needs backend "vmm": the slab is bound to a multicast object) ----
    @property
    def has_nvls(self) -> bool:
        return bool(self.multicast_ptr) and self.world > 1

    def _nvls(self, t: torch.Tensor, mode: int, root: int = 0) -> torch.Tensor:
        nbytes = t.numel() * t.element_size()
        rc = self.lib.hb_coll_nvls(self._ptr_array, ctypes.c_void_p(self.multicast_ptr), self.rank, self.world,
                                   self.offset_of(t), nbytes, mode, int(root), self.flags_offset, self._next_ticket(),
                                   _native.stream_ptr())
        _native.check(rc, "hb_coll_nvls")
        _native.count_launch()
        return t

    def nvls_allreduce_(self, t: torch.Tensor) -> torch.Tensor:
        """In-place float32 sum allreduce with ``multimem.ld_reduce`` + ``multimem.st`` (size a multiple of 16 bytes)."""
        assert t.dtype == torch.float32 and t.is_contiguous()
        return self._nvls(t, 0)

    def nvls_broadcast_(self, t: torch.Tensor, root: int) -> torch.Tensor:
        return self._nvls(t, 1, root)

    def nvls_allgather_(self, t: torch.Tensor) -> torch.Tensor:
        return self._nvls(t, 2)

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

    def _rows(self, local: torch.Tensor, glob: Optional[torch.Tensor], mode: int) -> None:
        assert local.is_contiguous() and local.dim() >= 1 and local.dtype in (torch.float32, torch.float64)
        rows = local.shape[0]
        row_elems = local.numel() // max(rows, 1)
        rc = self.lib.hb_coll_rows(self._ptr_array, self.rank, self.world, self.offset_of(local),
                                   self.offset_of(glob) if glob is not None else 0, rows, row_elems, _DT[local.dtype], mode,
                                   self.flags_offset, self._next_ticket(), _native.stream_ptr())
        _native.check(rc, "hb_coll_rows")
        _native.count_launch()

    def regroup_rows_(self, t: torch.Tensor) -> torch.Tensor:
        """Dense regroup with a SUM combiner: afterwards rank r holds the sum over ranks of the rows it owns
        (``row % world == r``), in place; the other rows keep this rank's stale partials."""
        self._rows(t, None, 0)
        return t

    def allgather_rows_(self, t: torch.Tensor) -> torch.Tensor:
        """Every owner's rows are stored into every peer's copy of ``t`` (same offset)."""
        self._rows(t, None, 1)
        return t

    def push_rows_(self, local: torch.Tensor, glob: torch.Tensor) -> None:
        """``glob`` (this rank's rows of the global table, id p at index p // world) += every rank's ``local[p]``."""
        self._rows(local, glob, 2)

    def pull_rows_(self, local: torch.Tensor, glob: torch.Tensor, use_bcast: bool = False) -> None:
        """``local[p]`` += the owner's global row of p, for every p (gather; with ``use_bcast`` the owners push)."""
        self._rows(local, glob, 4 if use_bcast else 3)

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
