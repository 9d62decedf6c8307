"""WDA-MDS device op (``csrc/mds.cu``): rows of ``B(X) X`` and the stress terms in one pass over the local rows of the
target-distance matrix, with a torch restatement for CPU tensors (``BCCalcTask`` / ``StressCalcTask``)."""
from __future__ import annotations

import ctypes
from typing import Tuple

import torch

from .. import _native


def bc_stress(D: torch.Tensor, ids: torch.Tensor, X: torch.Tensor, shift: float) -> Tuple[torch.Tensor, torch.Tensor]:
    """``(BX [m, dim], stress_terms [2])`` for the local rows: ``dt = max(D - shift, 0)``, weights 1 off the diagonal,
    ``r = dt / d(X)`` (0 where the layout distance vanishes), ``BX_i = (Σ_j r_ij) X_i − Σ_j r_ij X_j``,
    ``stress_terms = (Σ w (d − dt)², Σ w dt²)``."""
    m, n = D.shape
    dim = X.shape[1]
    if D.device.type != "cuda" or dim > 8:
        Xd = X.to(D.dtype)
        W = torch.ones_like(D)
        W[torch.arange(m, device=D.device), ids] = 0.0
        Dt = (D - shift).clamp(min=0.0)
        dist = (Xd[ids][:, None, :] - Xd[None, :, :]).pow(2).sum(2).sqrt()
        ratio = torch.where(dist > 1e-12, W * Dt / dist.clamp(min=1e-12), torch.zeros_like(dist))
        BX = ratio.sum(1)[:, None] * Xd[ids] - ratio @ Xd
        return BX, torch.stack([(W * (dist - Dt) ** 2).sum(), (W * Dt ** 2).sum()]).double()
    lib = _native.lib()
    if not getattr(lib, "_hb_mds_declared", False):
        P, I = ctypes.c_void_p, ctypes.c_int
        lib.hb_mds_bc_stress.argtypes = [P, P, P, I, I, I, ctypes.c_float, P, P, P]
        lib._hb_mds_declared = True
    Df = D if D.dtype == torch.float32 else D.float()
    Xf = X.to(torch.float32).contiguous()
    BX = torch.empty(m, dim, dtype=torch.float32, device=D.device)
    st = torch.zeros(2, dtype=torch.float64, device=D.device)
    _native.check(lib.hb_mds_bc_stress(_native.ptr(Df.contiguous()), _native.ptr(ids.to(torch.int64).contiguous()), _native.ptr(Xf),
                                       m, n, dim, float(shift), _native.ptr(BX), _native.ptr(st), _native.stream_ptr()),
                  "hb_mds_bc_stress")
    _native.count_launch()
    return BX, st
