"""Gram matrix / column moments ops (``csrc/gram.cu``) with CPU restatements.

These are the "partial result" producers of the DAAL-style distributed algorithms: every rank
computes ``XᵀX`` (and column sums) of its rows; the partials are then allreduced so that every
rank holds the result (the reference gathers them to one master instead,
``HD/data_comm/HarpDAALComm.java:203-246``)."""
from __future__ import annotations

import ctypes
from typing import NamedTuple, Optional

import torch

from .. import _native


def _declare(lib) -> None:
    if getattr(lib, "_hb_gram_declared", False):
        return
    P, I, LL = ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong
    lib.hb_gram_tc_supported.argtypes = [I]
    lib.hb_gram.argtypes = [P, LL, I, P, I, P]
    lib.hb_moments.argtypes = [P, LL, I, P, P, P, P, I, P]
    lib.hb_xty.argtypes = [P, LL, I, P, I, P, P]
    lib.hb_center_rows.argtypes = [P, LL, I, P, P, P]
    lib._hb_gram_declared = True


def gram(X: torch.Tensor, out: Optional[torch.Tensor] = None, accumulate: bool = False,
         force_simt: bool = False) -> torch.Tensor:
    """``XᵀX`` of a row-major matrix.  CUDA: tcgen05 TF32 kernel (fp32 SIMT kernel when ``d % 32 != 0``
    or ``force_simt``).  CPU: exact matmul in the input precision."""
    n, d = X.shape
    if X.device.type != "cuda":
        g = X.t() @ X
        if out is None:
            return g
        return out.add_(g) if accumulate else out.copy_(g)
    lib = _native.lib()
    _declare(lib)
    assert X.dtype == torch.float32
    X = X.contiguous()
    if out is None:
        out = torch.empty(d, d, dtype=torch.float32, device=X.device)
        accumulate = False
    flags = (1 if force_simt else 0) | (2 if accumulate else 0)
    _native.check(lib.hb_gram(_native.ptr(X), n, d, _native.ptr(out), flags, _native.stream_ptr()), "hb_gram")
    _native.count_launch(2)
    return out


def xty(X: torch.Tensor, Y: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``[Σx ; Yᵀ X]`` as ``[(m + 1), d]`` float64 in one streaming pass over ``X`` (K14, ``hb_xty``): row 0 holds the
    column sums, row ``1 + j`` the cross products with dependent variable ``j``.  The normal-equation partials of
    linear / ridge regression without ever building ``[1 | X | Y]``."""
    n, d = X.shape
    m = 0 if Y is None else Y.shape[1]
    if X.device.type != "cuda" or d % 4 != 0 or m > 8 or X.dtype != torch.float32:
        Xd = X.double()
        rows = [Xd.sum(0)] + ([Y.double().t() @ Xd] if m else [])
        return torch.cat([rows[0][None, :]] + rows[1:], 0)
    lib = _native.lib()
    _declare(lib)
    X = X.contiguous()
    Yc = Y.to(torch.float32).contiguous() if m else None
    out = torch.zeros(m + 1, d, dtype=torch.float64, device=X.device)
    _native.check(lib.hb_xty(_native.ptr(X), n, d, _native.ptr(Yc), m, _native.ptr(out), _native.stream_ptr()), "hb_xty")
    _native.count_launch()
    return out


def centered_gram(X: torch.Tensor, shift: torch.Tensor, chunk_rows: int = 1 << 20) -> torch.Tensor:
    """``Σ (x - shift)(x - shift)ᵀ`` (float32 result, tensor-core Gram of centred rows).  X is centred chunk by chunk
    into one scratch buffer of ``chunk_rows`` rows (``hb_center_rows``) and the chunks' Grams are accumulated, so the
    extra memory is bounded; used when the mean of the data is large against its spread, where TF32 products of the raw
    values keep almost nothing of the spread."""
    n, d = X.shape
    if X.device.type != "cuda" or d % 4 != 0:
        Xc = X - shift.to(X.dtype)
        return Xc.t() @ Xc
    lib = _native.lib()
    _declare(lib)
    X = X.contiguous()
    sh = shift.to(torch.float32).contiguous()
    G = torch.zeros(d, d, dtype=torch.float32, device=X.device)
    buf = torch.empty(min(n, chunk_rows), d, dtype=torch.float32, device=X.device)
    for r0 in range(0, n, chunk_rows):
        rows = min(chunk_rows, n - r0)
        _native.check(lib.hb_center_rows(_native.ptr(X[r0:]), rows, d, _native.ptr(sh), _native.ptr(buf),
                                         _native.stream_ptr()), "hb_center_rows")
        _native.count_launch()
        gram(buf[:rows], out=G, accumulate=True)
    return G


class Moments(NamedTuple):
    sum: torch.Tensor      # [d] float64
    sumsq: torch.Tensor    # [d] float64
    min: torch.Tensor      # [d] float32 (input dtype on CPU)
    max: torch.Tensor
    n: int


def moments(X: torch.Tensor) -> Moments:
    n, d = X.shape
    if X.device.type != "cuda":
        Xd = X.double()
        if n == 0:
            inf = torch.full((d,), float("inf"), dtype=X.dtype)
            return Moments(torch.zeros(d, dtype=torch.float64), torch.zeros(d, dtype=torch.float64), inf, -inf, 0)
        return Moments(Xd.sum(0), (Xd * Xd).sum(0), X.min(0).values, X.max(0).values, n)
    lib = _native.lib()
    _declare(lib)
    X = X.contiguous()
    s = torch.empty(d, dtype=torch.float64, device=X.device)
    ss = torch.empty(d, dtype=torch.float64, device=X.device)
    mn = torch.empty(d, dtype=torch.float32, device=X.device)
    mx = torch.empty(d, dtype=torch.float32, device=X.device)
    _native.check(lib.hb_moments(_native.ptr(X), n, d, _native.ptr(s), _native.ptr(ss), _native.ptr(mn),
                                 _native.ptr(mx), 0, _native.stream_ptr()), "hb_moments")
    _native.count_launch(2)
    return Moments(s, ss, mn, mx, n)


class FusedGram:
    """``XᵀX`` of this rank's rows **and** the allreduce of the ranks' partial results in one launch
    (``csrc/gram.cu``: ``gram_tc_kernel<1>`` + ``gram_fused_tail``) — the DAAL-style "partial result →
    master → broadcast" step (``HD/data_comm/HarpDAALComm.java:203-246``) done over NVLink peer memory by the
    kernel that produced the partial result.  ``state`` is this rank's block of a symmetric slab (same offset
    on every rank), ``peer_ptrs`` the device addresses of that block on every rank.  Works on one GPU too
    (Gram + mirror in one launch)."""

    def __init__(self, state: torch.Tensor, peer_ptrs, rank: int, d: int):
        lib = _native.lib()
        _declare(lib)
        lib.hb_gram_fused_state_bytes.restype = ctypes.c_size_t
        lib.hb_gram_fused_state_bytes.argtypes = [ctypes.c_int]
        lib.hb_gram_allreduce_fused.argtypes = [ctypes.c_void_p, ctypes.c_longlong, ctypes.c_int, ctypes.c_void_p,
                                                ctypes.c_int, ctypes.c_int, ctypes.c_uint, ctypes.c_void_p]
        assert lib.hb_gram_tc_supported(d), f"fused Gram needs d % 32 == 0 (got {d})"
        need = self.state_bytes(d)
        assert state.dtype == torch.uint8 and state.numel() >= need and state.data_ptr() % 256 == 0
        assert int(peer_ptrs[rank]) == state.data_ptr()
        self.lib, self.d, self.rank, self.world = lib, int(d), int(rank), len(peer_ptrs)
        self.state = state
        self._peers = (ctypes.c_void_p * self.world)(*[ctypes.c_void_p(int(p)) for p in peer_ptrs])
        self.G = state[256:256 + d * d * 4].view(torch.float32).view(d, d)
        self.epoch = 0
        self._stats = None
        lib.hb_gram_allreduce_fused_stats.argtypes = [ctypes.c_void_p, ctypes.c_longlong, ctypes.c_int, ctypes.c_void_p,
                                                      ctypes.c_int, ctypes.c_int, ctypes.c_uint, ctypes.c_void_p,
                                                      ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
        state[:256].zero_()          # flags start at 0; the caller barriers across ranks before the first call

    @staticmethod
    def state_bytes(d: int) -> int:
        return 256 + d * d * 4

    def __call__(self, X: torch.Tensor, want_moments: bool = False):
        """Returns the global Gram matrix (a view of the state, overwritten by the next call).  With ``want_moments``
        also the column ``Moments`` of THIS rank's rows, produced by the same launch (the statistics warps of the
        Gram kernel re-read through L2 what its TMA warp streams, so X leaves HBM once for both)."""
        assert X.dtype == torch.float32 and X.dim() == 2 and X.shape[1] == self.d
        X = X.contiguous()
        P = ctypes.c_void_p
        if want_moments:
            if self._stats is None:
                dev = X.device
                self._stats = (torch.empty(self.d, dtype=torch.float64, device=dev), torch.empty(self.d, dtype=torch.float64, device=dev),
                               torch.empty(self.d, dtype=torch.float32, device=dev), torch.empty(self.d, dtype=torch.float32, device=dev))
            sp = [_native.ptr(t) for t in self._stats]
        else:
            sp = [P(0)] * 4
        rc = self.lib.hb_gram_allreduce_fused_stats(_native.ptr(X), X.shape[0], self.d, ctypes.cast(self._peers, P),
                                                    self.rank, self.world, self.epoch, *sp, _native.stream_ptr())
        _native.check(rc, "hb_gram_allreduce_fused_stats")
        _native.count_launch(2 if want_moments else 1)
        self.epoch += 1
        if want_moments:
            s, ss, mn, mx = self._stats
            return self.G, Moments(s, ss, mn, mx, X.shape[0])
        return self.G


def fused_gram_for(comm, d: int, device: torch.device) -> Optional["FusedGram"]:
    """A cached ``FusedGram`` bound to ``comm``'s symmetric workspace (None when the fused path does not apply:
    CPU, ``d % 32 != 0``, or several ranks without a symmetric workspace)."""
    if device.type != "cuda" or d % 32 != 0 or d < 32 or d > 16384:
        return None
    world = 1 if comm is None else comm.world
    symm = None if comm is None else getattr(comm, "symm", None)
    if world > 1 and symm is None:
        return None
    holder = comm if comm is not None else fused_gram_for
    cache = getattr(holder, "_fused_gram_cache", None)
    if cache is None:
        cache = {}
        setattr(holder, "_fused_gram_cache", cache)
    key = (d, device.index)
    if key not in cache:
        nbytes = FusedGram.state_bytes(d)
        if world > 1:
            try:
                st = symm.alloc(nbytes, torch.uint8)
            except MemoryError:
                cache[key] = None
                return None
            off = symm.offset_of(st)
            fg = FusedGram(st, [p + off for p in symm.peer_ptrs], comm.rank, d)
            torch.cuda.synchronize(device)
            import torch.distributed as dist
            dist.barrier(group=comm.group)
        else:
            st = torch.empty(nbytes, dtype=torch.uint8, device=device)
            fg = FusedGram(st, [st.data_ptr()], 0, d)
        cache[key] = fg
    return cache[key]
