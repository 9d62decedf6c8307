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
