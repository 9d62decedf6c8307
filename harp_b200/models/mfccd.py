"""MF-CCD++ engine: W sharded by rows, H by columns (owner = id % N), residuals local, two kernel passes per feature.

The reference rotates the rank-one rows of W and H through the workers while every worker sweeps its own rows /
columns (``ml/java/src/main/java/edu/iu/ccd/CCDMPCollectiveMapper.java:546-547``, ``:635-636``;
``CCDMPTask.java:67-125``).  Here a feature's vector is needed by every rank at once, so the owners' entries are
all-gathered over NVLink (``SymmetricWorkspace.allgather_rows_``: owners store their rows into every peer) and each
rank runs ONE kernel pass per side and feature over the ratings of the ids it owns (``csrc/mf.cu::ccd_sweep_kernel``).
Every rank keeps its ratings twice — by row for the W sweeps, by column for the H sweeps — each with its own residual
array; a sweep over one order leaves the other stale by one rank-one term, which the next sweep over that order applies
on the fly, so no pass exists only to keep the two orders in step.

Order inside a feature: H sweep, then W sweep (as ``apps/mf.py`` has always done); ``inner`` repeats the pair.
"""
from __future__ import annotations

import math
from typing import Optional

import torch
import torch.distributed as dist

from ..ops import mf


def _csr(key_local: torch.Tensor, other: torch.Tensor, vals: torch.Tensor, n_keys: int):
    order = torch.argsort(key_local, stable=True)
    cnt = torch.bincount(key_local, minlength=n_keys)
    ptr = torch.zeros(n_keys + 1, dtype=torch.int32, device=key_local.device)
    ptr[1:] = torch.cumsum(cnt, 0).to(torch.int32)
    return ptr, other[order].to(torch.int32).contiguous(), vals[order].contiguous()


class MFCCDEngine:
    """``rows_r/cols_r/vals_r``: the ratings whose ROW this rank owns (global ids, row % N == rank);
    ``rows_c/cols_c/vals_c``: the ratings whose COLUMN it owns.  On one rank both are the whole set."""

    def __init__(self, rows_r, cols_r, vals_r, rows_c, cols_c, vals_c, n_rows: int, n_cols: int, r: int, lam: float, *,
                 symm=None, group=None, seed: int = 0, inner: int = 1):
        self.dev = rows_r.device
        self.N = dist.get_world_size(group) if dist.is_initialized() else 1
        self.me = dist.get_rank(group) if dist.is_initialized() else 0
        self.group = group
        N, me, dev = self.N, self.me, self.dev
        self.r, self.lam, self.inner = int(r), float(lam), int(inner)
        self.n_rows, self.n_cols = int(n_rows), int(n_cols)
        self.dt = torch.float32 if dev.type == "cuda" else torch.float64
        self.own_r = torch.arange(me, n_rows, N, device=dev)
        self.own_c = torch.arange(me, n_cols, N, device=dev)
        self.ptr_r, self.idx_r, vr = _csr(rows_r.long() // N, cols_r.long(), vals_r.to(self.dt), self.own_r.numel())
        self.ptr_c, self.idx_c, vc = _csr(cols_c.long() // N, rows_c.long(), vals_c.to(self.dt), self.own_c.numel())
        self.res_r, self.res_c = vr.clone(), vc.clone()                    # H = 0 at the start: residual = rating
        self.nnz = int(vr.numel())
        # one CTA per id where the lists are long (a column of a tall matrix), one warp otherwise
        self.wide_r = self.nnz > 512 * max(int(self.own_r.numel()), 1)
        self.wide_c = int(vc.numel()) > 512 * max(int(self.own_c.numel()), 1)
        g = torch.Generator(device=dev).manual_seed(seed + 11)
        scale = 1.0 / math.sqrt(self.r)
        W0 = torch.rand(self.r, n_rows, device=dev, generator=g, dtype=self.dt) * scale      # same stream on every rank
        self.W_loc = W0[:, self.own_r].contiguous()                                          # feature-major shards
        self.H_loc = torch.zeros(self.r, self.own_c.numel(), device=dev, dtype=self.dt)
        del W0
        self.symm = symm if (N > 1 and dev.type == "cuda" and symm is not None) else None
        alloc = (lambda n: self.symm.alloc(n, torch.float32)) if self.symm is not None else \
                (lambda n: torch.zeros(n, device=dev, dtype=self.dt))
        self.u = alloc(n_rows)             # the feature's full vectors, gathered from the owners
        self.v = alloc(n_cols)
        self.du = alloc(n_rows)            # ... and their change in the sweep just done
        self.dv = alloc(n_cols)
        self.du_loc = torch.zeros(self.own_r.numel(), device=dev, dtype=self.dt)
        self.dv_loc = torch.zeros(self.own_c.numel(), device=dev, dtype=self.dt)
        self.pend_c: Optional[torch.Tensor] = None     # owned-column values of the feature whose W sweep res_c has not seen
        self.pend_r: Optional[torch.Tensor] = None
        self.iteration = 0
        if dist.is_initialized() and N > 1:
            dist.barrier(group=group)

    # -- the owners' entries of a vector reach every rank -------------------------------------------------------
    def _gather(self, full: torch.Tensor, own: torch.Tensor, local: torch.Tensor) -> None:
        if self.N == 1:
            full.copy_(local)
            return
        if self.symm is not None:
            full[self.me::self.N] = local              # owners' entries, then every owner stores its rows into every peer
            self.symm.allgather_rows_(full.view(-1, 1))
            return
        full.zero_()
        full[self.me::self.N] = local
        dist.all_reduce(full, group=self.group)

    def iterate(self) -> float:
        """One outer iteration (all r features); returns the training RMSE from the residuals."""
        for f in range(self.r):
            self._gather(self.u, self.own_r, self.W_loc[f])
            for _ in range(self.inner):
                # H sweep over the owned columns; res_c may still lack the previous W sweep
                v_loc = self.H_loc[f]
                mf.ccd_sweep(self.ptr_c, self.idx_c, self.res_c, self.u, v_loc, self.lam,
                             pd_other=self.du if self.pend_c is not None else None, pd_self=self.pend_c,
                             dself=self.dv_loc, wide=self.wide_c)
                self.pend_c = None
                self._gather(self.v, self.own_c, v_loc)
                self._gather(self.dv, self.own_c, self.dv_loc)
                # W sweep over the owned rows; res_r lacks the H sweep just done: dv[j] * u_old[i]
                u_loc = self.W_loc[f]
                u_old = u_loc.clone()
                mf.ccd_sweep(self.ptr_r, self.idx_r, self.res_r, self.v, u_loc, self.lam,
                             pd_other=self.dv, pd_self=u_old, dself=self.du_loc, wide=self.wide_r)
                self._gather(self.u, self.own_r, u_loc)
                self._gather(self.du, self.own_r, self.du_loc)
                self.pend_c = v_loc.clone()            # res_c lacks du[i] * v_new[j]
        self.iteration += 1
        return self.rmse()

    def close(self) -> None:
        """Gives the gathered-vector buffers back to the symmetric workspace (every rank, same order)."""
        if self.symm is not None:
            torch.cuda.synchronize(self.dev)
            for t in (self.dv, self.du, self.v, self.u):
                self.symm.free(t)
            self.symm = None

    def rmse(self) -> float:
        st = torch.stack([(self.res_r.double() ** 2).sum(), torch.tensor(float(self.res_r.numel()), dtype=torch.float64, device=self.dev)])
        if self.N > 1:
            dist.all_reduce(st, group=self.group)
        return math.sqrt(float(st[0]) / max(float(st[1]), 1.0))

    def full_factors(self):
        """(W [r, n_rows], H [r, n_cols]) gathered from the shards (model files, test RMSE)."""
        W = torch.zeros(self.r, self.n_rows, device=self.dev, dtype=self.dt)
        H = torch.zeros(self.r, self.n_cols, device=self.dev, dtype=self.dt)
        W[:, self.own_r] = self.W_loc
        H[:, self.own_c] = self.H_loc
        if self.N > 1:
            dist.all_reduce(W, group=self.group)
            dist.all_reduce(H, group=self.group)
        return W, H
