"""Host side of the in-kernel model rotation (``csrc/rotate_ring.cuh``).

``RingSchedule`` turns the reference's rotation orders into the per-step tables the epoch kernels need.  It is
the same walk as ``RotateTask.updateRotationMap`` (``ml/java/src/main/java/edu/iu/dymoro/RotateTask.java:158-210``,
via ``dymoro.RotationMapper``), evaluated identically on every rank: after each rotate call the schedule knows
which shard sits on which worker and which worker each rank must pull from.  ``orders=None`` is the plain ring
``next = (self + 1) % N`` (``LocalGlobalSyncCollective.java:718-730``), what LDA uses
(``LDAMPCollectiveMapper.java:255-259``); SGD and CCD pass randomised orders
(``RotationUtil.java:58-90``).

``RingState`` owns the peer-visible memory of one engine: the 64-byte control block of counters/flags and the
ping-pong buffers of the S model slices, as ONE allocation in the symmetric slab (or a plain device buffer on
one GPU), and marshals the arguments of ``hb_sgd_epoch`` / ``hb_lda_epoch``.

Waits inside the kernels are bounded by ``HB_WAIT_TIMEOUT_S`` (default 1800 s, the reference's
``DATA_MAX_WAIT_TIME``, ``core/harp-collective/.../io/Constant.java:35``; 0 disables the bound).  When a wait
expires the kernel records a code in the control block and drains instead of trapping; ``RingState.check()``
turns it into a Python exception.
"""
from __future__ import annotations

import ctypes
import os
from typing import List, Optional, Sequence

import torch

from .. import _native
from ..dymoro import RotationMapper

_ERR = {1: "the consumer rank never pulled this rank's previous slice",
        2: "the source rank never finished the slice this rank wanted to pull",
        3: "the local communication warps never delivered the slice",
        4: "a column-block ticket never arrived",
        5: "aborted after another waiter gave up"}


def wait_timeout_ns() -> int:
    return int(float(os.environ.get("HB_WAIT_TIMEOUT_S", "1800")) * 1e9)


class RingSchedule:
    """Where every shard of ONE model slice sits, rotate call after rotate call."""

    def __init__(self, num_workers: int, orders: Optional[Sequence[int]] = None):
        self.n = int(num_workers)
        self.mapper = RotationMapper(orders, self.n) if orders is not None else None
        self.shard_at = list(range(self.n))          # shard_at[worker] = shard currently held by that worker
        self.rotations = 0

    def rotate(self) -> List[int]:
        """Advances by one rotate call; returns ``src`` with ``src[w]`` = the worker that ``w`` receives from."""
        n = self.n
        rmap = self.mapper.next_map() if self.mapper is not None else None
        if rmap is None:
            rmap = {w: (w + 1) % n for w in range(n)}
        src = [0] * n
        new = [0] * n
        for w in range(n):
            dst = rmap[w]
            src[dst] = w
            new[dst] = self.shard_at[w]
        assert sorted(src) == list(range(n)), "rotation map is not a permutation"
        self.shard_at = new
        self.rotations += 1
        return src


class RingState:
    """Control block + ping-pong slice buffers of a rotation engine (see module docstring)."""

    def __init__(self, device: torch.device, slice_numel: int, dtype: torch.dtype, num_slices: int, *, symm=None,
                 rank: int = 0, world: int = 1, loopback: bool = False):
        self.lib = _native.lib()
        self.device = device
        self.S = int(num_slices)
        self.rank, self.world = int(rank), int(world)
        self.ring_world = self.world if self.world > 1 else (1 if loopback else 0)
        esz = torch.empty((), dtype=dtype).element_size()
        self.slice_bytes = (slice_numel * esz + 255) // 256 * 256
        nbuf = 2 if self.ring_world > 0 else 1
        self.nbuf = nbuf
        total = 256 + self.S * nbuf * self.slice_bytes
        self._symm = symm if (symm is not None and self.world > 1) else None
        if symm is not None and self.world > 1:
            self.state = symm.alloc(total, torch.uint8)
            base = symm.offset_of(self.state)
            self.peer_ptrs = [int(p) + base for p in symm.peer_ptrs]
        else:
            assert self.world == 1, "a multi-rank ring needs a symmetric workspace"
            self.state = torch.zeros(total, dtype=torch.uint8, device=device)
            self.peer_ptrs = [self.state.data_ptr()]
        self.state.zero_()
        self.buf: List[List[torch.Tensor]] = []
        self.off_buf: List[int] = []
        for k in range(self.S):
            pair = []
            for p in range(2):
                off = 256 + (k * nbuf + (p if nbuf == 2 else 0)) * self.slice_bytes
                self.off_buf.append(off)
                pair.append(self.state[off:off + slice_numel * esz].view(dtype))
            self.buf.append(pair)
        self.cur = [0] * self.S            # parity of the buffer that holds the slice's current contents
        self.cnt = 0                       # uses per slice completed by the epoch kernels
        self._peer_arr = (ctypes.c_void_p * 16)(*([ctypes.c_void_p(p) for p in self.peer_ptrs] +
                                                  [ctypes.c_void_p(0)] * (16 - len(self.peer_ptrs))))
        self._off_arr = (ctypes.c_ulonglong * 4)(*(self.off_buf + [0] * (4 - len(self.off_buf))))
        self._bytes_arr = (ctypes.c_ulonglong * 2)(*([self.slice_bytes] * self.S + [0] * (2 - self.S)))

    def release(self) -> None:
        """Returns the ring's slab allocation to the workspace (every rank, same order; nothing may be in flight)."""
        if self._symm is not None and self.state is not None:
            self._symm.free(self.state)
        self.state = None

    def current(self, k: int) -> torch.Tensor:
        return self.buf[k][self.cur[k]]

    def other(self, k: int) -> torch.Tensor:
        return self.buf[k][1 - self.cur[k]]

    def ring_args(self, src_of_step: Optional[Sequence[int]]):
        """Positional tail shared by the epoch entry points: (peer_state, rank, world, off_ctl, off_buf4, bytes2,
        src_rank, cnt_base, par_base2, timeout_ns)."""
        par = (ctypes.c_uint32 * 2)(*(self.cur + [0] * (2 - self.S)))
        src = None
        if src_of_step is not None and self.ring_world > 0:
            src = (ctypes.c_int * len(src_of_step))(*[int(x) % self.ring_world for x in src_of_step])
        return (self._peer_arr, self.rank if self.ring_world > 0 else 0, self.ring_world, 0, self._off_arr,
                self._bytes_arr, src, self.cnt & 0xFFFFFFFF, par, wait_timeout_ns())

    def advance(self, uses: int) -> None:
        """Book-keeping after an epoch launch that trained ``uses`` times on every slice."""
        if self.ring_world > 0:
            self.cnt += uses
            if uses & 1:
                self.cur = [c ^ 1 for c in self.cur]

    def check(self) -> None:
        """Raises if a bounded wait inside an epoch kernel expired (call after a synchronize)."""
        words = self.state[:64].view(torch.int32).cpu()
        code, detail = int(words[10]), int(words[11])
        if code:
            raise _native.NativeError(f"model rotation stalled at step {detail}: {_ERR.get(code, code)} "
                                      f"(HB_WAIT_TIMEOUT_S={os.environ.get('HB_WAIT_TIMEOUT_S', '1800')})")
