"""Dense linear-algebra ops of the DAAL-style algorithms (``csrc/linalg.cu``) with CPU restatements.

``tsqr_r``        R factor of a tall-skinny matrix in one streaming pass (K21; daal_qr / daal_svd / daal_pca step 1-2,
                  ``ml/daal/.../daal_svd/SVDDaalCollectiveMapper.java:136-219``).
``potrs_batched`` batched Cholesky factor / solve, one warp per system (K16).
``als_solve``     implicit-ALS step 4 for a set of rows (K22; ``ml/daal/.../daal_als/ALSDaalCollectiveMapper.java:354-492``).
"""
from __future__ import annotations

import ctypes
from typing import Optional, Tuple

import torch

from .. import _native


def _declare(lib) -> None:
    if getattr(lib, "_hb_linalg_declared", False):
        return
    P, I, LL, F = ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong, ctypes.c_float
    lib.hb_tsqr_pad.argtypes = [I]
    lib.hb_tsqr_workspace_bytes.restype = ctypes.c_size_t
    lib.hb_tsqr_workspace_bytes.argtypes = [I]
    lib.hb_tsqr.argtypes = [P, LL, I, LL, P, P, P]
    lib.hb_chol_pad.argtypes = [I]
    lib.hb_potrs_batched.argtypes = [P, P, P, P, I, LL, P, P]
    lib.hb_als_solve.argtypes = [P, P, P, P, P, F, F, P, I, I, P, P, P, P]
    lib._hb_linalg_declared = True


def tsqr_supported(X: torch.Tensor) -> bool:
    return (X.device.type == "cuda" and X.dtype == torch.float32 and X.dim() == 2 and 1 <= X.shape[1] <= 64
            and X.layout == torch.strided)


def tsqr_r(X: torch.Tensor) -> torch.Tensor:
    """Upper-triangular ``R`` (``[d, d]``, same dtype as ``X``) with ``RᵀR = XᵀX``; the sign of each row is whatever
    the reflections produced (callers normalise).  CUDA fp32 with d ≤ 64: the Householder streaming kernel; otherwise
    ``torch.linalg.qr``."""
    n, d = X.shape
    if not tsqr_supported(X):
        return torch.linalg.qr(X, mode="r").R
    lib = _native.lib()
    _declare(lib)
    if X.stride(1) != 1:
        X = X.contiguous()
    D = lib.hb_tsqr_pad(d)
    work = torch.empty(lib.hb_tsqr_workspace_bytes(d) // 4, dtype=torch.float32, device=X.device)
    R = torch.empty(D, D, dtype=torch.float32, device=X.device)
    _native.check(lib.hb_tsqr(_native.ptr(X), n, d, X.stride(0), _native.ptr(work), _native.ptr(R),
                              _native.stream_ptr()), "hb_tsqr")
    _native.count_launch(2)
    return R[:d, :d]


def potrs_batched(A: torch.Tensor, B: Optional[torch.Tensor] = None, want_factor: bool = False
                  ) -> Tuple[Optional[torch.Tensor], Optional[torch.Tensor], int]:
    """Cholesky of every ``A[i]`` (``[batch, f, f]`` SPD) and, when ``B`` (``[batch, f]``) is given, the solutions of
    ``A[i] x = B[i]``.  Returns ``(X or None, L or None, number of matrices with a non-positive pivot)``."""
    batch, f, _ = A.shape
    if A.device.type != "cuda" or A.dtype != torch.float32 or f > 64:
        L, info = torch.linalg.cholesky_ex(A)
        X = torch.cholesky_solve(B.unsqueeze(-1), L).squeeze(-1) if B is not None else None
        return X, (L if want_factor else None), int((info != 0).sum())
    lib = _native.lib()
    _declare(lib)
    A = A.contiguous()
    X = torch.empty(batch, f, dtype=torch.float32, device=A.device) if B is not None else None
    L = torch.empty_like(A) if want_factor else None
    info = torch.zeros(1, dtype=torch.int32, device=A.device)
    _native.check(lib.hb_potrs_batched(_native.ptr(A), _native.ptr(B.contiguous() if B is not None else None),
                                       _native.ptr(X), _native.ptr(L), f, batch, _native.ptr(info),
                                       _native.stream_ptr()), "hb_potrs_batched")
    _native.count_launch(1)
    return X, L, int(info.item())


def als_pad(f: int) -> int:
    return 16 if f <= 16 else 32 if f <= 32 else 64 if f <= 64 else 0


class ALSSide:
    """Device-resident CSR of one side's ratings (rows owned by this rank) plus the scratch of ``hb_als_solve``."""

    def __init__(self, ptr: torch.Tensor, idx: torch.Tensor, val: torch.Tensor, rows: Optional[torch.Tensor] = None):
        self.ptr = ptr.to(torch.int64).contiguous()
        self.idx = idx.to(torch.int32).contiguous()
        self.val = val.to(torch.float32).contiguous()
        self.rows = rows.to(torch.int32).contiguous() if rows is not None else None
        self.n_rows = int(self.rows.numel()) if rows is not None else int(self.ptr.numel()) - 1
        self.counter = torch.zeros(1, dtype=torch.int32, device=ptr.device)
        self.info = torch.zeros(1, dtype=torch.int32, device=ptr.device)


def als_solve(side: ALSSide, other: torch.Tensor, out: torch.Tensor, f: int, lam: float, alpha: float,
              YtY: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``out[u] = (YᵀY + Yᵤᵀ(Cᵤ − I)Yᵤ + λI)⁻¹ Yᵤᵀ Cᵤ 1`` for the rows of ``side``; ``other`` and ``out`` are fp32
    ``[*, F]`` with ``F = als_pad(f)`` and zero padding columns.  ``YtY`` may be passed when it was reduced over ranks."""
    lib = _native.lib()
    _declare(lib)
    F = als_pad(f)
    assert F and other.shape[1] == F and out.shape[1] == F and other.dtype == torch.float32 and out.dtype == torch.float32
    if YtY is None:
        YtY = other.t() @ other
    YtY = YtY.to(torch.float32).contiguous()
    _native.check(lib.hb_als_solve(_native.ptr(side.ptr), _native.ptr(side.idx), _native.ptr(side.val),
                                   _native.ptr(other), _native.ptr(YtY), float(lam), float(alpha),
                                   _native.ptr(side.rows), side.n_rows, f, _native.ptr(out), _native.ptr(side.counter),
                                   _native.ptr(side.info), _native.stream_ptr()), "hb_als_solve")
    _native.count_launch(1)
    return out
