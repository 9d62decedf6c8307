"""MF-SGD engine with model rotation.

Reference skeleton (``ml/java/src/main/java/edu/iu/sgd/SGDCollectiveMapper.java:154-320``): W rows stay
on the worker that owns them, H columns live in ``numModelSlices = 2`` slices per worker that travel
a ring; while slice k is being trained on, slice 1-k is in flight to the next worker
(``dymoro/Rotator.java:53-67``).  One epoch = N rotations x 2 slices, so every H column meets every
worker's ratings exactly once.

Here: column j has home worker ``j % N``, slice ``(j // N) % S`` and slice-local index ``j // SN``
(``S = num_slices``: 2 like the reference when panels are large; 1 when they are small — over NVLink a slice
moves in microseconds, and one fatter panel per rotation step keeps the block schedule efficient).
Each slice has a ping-pong pair of buffers in the symmetric workspace; the rotation is a peer *pull*
over NVLink (``SymmetricWorkspace.rotate_``) issued on a side stream, overlapping the SGD kernel of
the other slice.  The per-panel block schedule and the update itself are ``ops.mf.sgd_pass``.
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional, Tuple

import torch
import torch.distributed as dist

from ..ops import mf


class MFSGDEngine:
    def __init__(self, rows: torch.Tensor, cols: torch.Tensor, vals: torch.Tensor, num_rows_local: int,
                 num_cols: int, rank: int, eps: float, lam: float, *, symm=None, blocks: Optional[int] = None,
                 seed: int = 0, group=None, num_slices: Optional[int] = None):
        self.dev = rows.device
        self.rank_k = int(rank)
        self.eps, self.lam = float(eps), float(lam)
        self.group = group
        self.N = dist.get_world_size(group) if dist.is_initialized() else 1
        self.me = dist.get_rank(group) if dist.is_initialized() else 0
        self.symm = symm if self.N > 1 else None
        self.num_cols = int(num_cols)
        self.num_rows_local = int(num_rows_local)
        N = self.N
        if num_slices is None:
            # two slices (compute one while the other travels) only pay off while a half panel is still big
            nnz_max = float(rows.numel())
            if dist.is_initialized() and N > 1:
                t = torch.tensor([nnz_max], dtype=torch.float64, device=self.dev)
                dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
                nnz_max = float(t.item())
            num_slices = 2 if (N == 1 or nnz_max / (2.0 * N) >= 4.0e6) else 1
        self.S = S = int(num_slices)
        assert S in (1, 2)
        self.cap = max(1, (self.num_cols + S * N - 1) // (S * N))          # columns per (shard, slice)
        if self.dev.type == "cuda":
            mx = mf.max_blocks(self.rank_k)
            sms = torch.cuda.get_device_properties(self.dev).multi_processor_count
            self.B = int(blocks) if blocks else mf.auto_blocks(rows.numel() / (float(S) * N), min(mx, 2 * sms))
        else:
            self.B = int(blocks) if blocks else 4
        g = torch.Generator(device=self.dev).manual_seed(seed + 7919 * self.me)
        scale = 1.0 / math.sqrt(self.rank_k)
        dt = torch.float32 if self.dev.type == "cuda" else torch.float64
        self.dt = dt
        self.W = (torch.rand(self.num_rows_local, self.rank_k, device=self.dev, generator=g, dtype=dt) * scale).contiguous()
        # H slices: ping-pong buffers (symmetric memory when rotating)
        self.Hbuf: List[List[torch.Tensor]] = []
        for k in range(S):
            pair = []
            for _ in range(2 if N > 1 else 1):
                if self.symm is not None:
                    t = self.symm.alloc(self.cap * self.rank_k, dt).view(self.cap, self.rank_k)
                else:
                    t = torch.empty(self.cap, self.rank_k, dtype=dt, device=self.dev)
                pair.append(t)
            self.Hbuf.append(pair)
        self.cur = [0] * S
        self.rot = [0] * S
        gh = torch.Generator(device=self.dev).manual_seed(seed + 104729)      # same H init on every rank, then
        for k in range(S):                                                    # each keeps its own shard
            full = torch.rand(N, self.cap, self.rank_k, device=self.dev, generator=gh, dtype=dt) * scale
            self.Hbuf[k][0].copy_(full[self.me])
        # panels[shard][slice]
        cols = cols.to(torch.int64)
        shard = cols % N
        slc = (cols // N) % S
        local = cols // (S * N)
        self.panels: List[List[mf.SgdPanel]] = []
        self.nnz = int(rows.numel())
        for s in range(N):
            per = []
            for k in range(S):
                m = (shard == s) & (slc == k)
                per.append(mf.build_panel(rows[m], local[m], vals[m], self.B))
            self.panels.append(per)
        self.trained = torch.zeros(1, dtype=torch.int64, device=self.dev)
        if self.dev.type == "cuda":
            self.comm_stream = torch.cuda.Stream(device=self.dev)
            self.ev_compute = [torch.cuda.Event(), torch.cuda.Event()]
            self.ev_comm = [torch.cuda.Event(), torch.cuda.Event()]
            for e in self.ev_comm:
                e.record()
        self.comm_time_hidden = True

    # ------------------------------------------------------------------------------------
    def _visiting_shard(self, k: int) -> int:
        return (self.me - self.rot[k]) % self.N

    def _rotate(self, k: int) -> None:
        """slice k moves to the next worker: every worker pulls its predecessor's copy."""
        if self.N == 1:
            return
        src = self.Hbuf[k][self.cur[k]]
        dst = self.Hbuf[k][1 - self.cur[k]]
        pred = (self.me - 1) % self.N
        if self.symm is not None and self.dev.type == "cuda":
            self.ev_compute[k].record()
            with torch.cuda.stream(self.comm_stream):
                self.comm_stream.wait_event(self.ev_compute[k])
                self.symm.rotate_(dst, src, pred)
                self.ev_comm[k].record()
        else:
            # gloo / NCCL send-recv ring (baseline path and CPU tests)
            send = src.contiguous()
            ops = [dist.P2POp(dist.isend, send, (self.me + 1) % self.N, group=self.group),
                   dist.P2POp(dist.irecv, dst, pred, group=self.group)]
            for r in dist.batch_isend_irecv(ops):
                r.wait()
        self.cur[k] ^= 1
        self.rot[k] += 1

    def _wait_slice(self, k: int) -> None:
        if self.N > 1 and self.symm is not None and self.dev.type == "cuda":
            torch.cuda.current_stream().wait_event(self.ev_comm[k])

    def epoch(self, deadline_ms: Optional[float] = None) -> torch.Tensor:
        """N rotations x S slices (``SGDCollectiveMapper.java:263-277``). Returns the trained-rating counter."""
        self.trained.zero_()
        for _ in range(self.N):
            for k in range(self.S):
                self._wait_slice(k)
                panel = self.panels[self._visiting_shard(k)][k]
                mf.sgd_pass(panel, self.W, self.Hbuf[k][self.cur[k]], self.eps, self.lam, deadline_ms, self.trained)
                self._rotate(k)
        return self.trained

    def sq_err(self, rows: torch.Tensor, cols: torch.Tensor, vals: torch.Tensor) -> Tuple[float, int]:
        """Global (Σ squared error, count) of a rating list given with worker-local rows / global columns:
        one extra full rotation (``SGDCollectiveMapper.java:670-735``)."""
        N, S = self.N, self.S
        cols = cols.to(torch.int64)
        shard, slc, local = cols % N, (cols // N) % S, cols // (S * N)
        acc = torch.zeros(1, dtype=torch.float64, device=self.dev)
        for _ in range(N):
            for k in range(S):
                self._wait_slice(k)
                s = self._visiting_shard(k)
                m = (shard == s) & (slc == k)
                mf.sq_err(self.W, self.Hbuf[k][self.cur[k]], rows[m], local[m], vals[m].to(self.dt), acc)
                self._rotate(k)
        cnt = torch.tensor([float(rows.numel())], dtype=torch.float64, device=self.dev)
        if N > 1:
            dist.all_reduce(acc, group=self.group)
            dist.all_reduce(cnt, group=self.group)
        return float(acc.item()), int(cnt.item())

    def rmse(self, rows, cols, vals) -> float:
        se, n = self.sq_err(rows, cols, vals)
        return math.sqrt(se / max(n, 1))

    def synchronize(self) -> None:
        if self.dev.type == "cuda":
            for k in range(self.S):
                self._wait_slice(k)
            torch.cuda.synchronize(self.dev)

    def home_H(self) -> Dict[int, torch.Tensor]:
        """The H columns currently held by this worker as ``{global column id: vector}`` blocks
        (``[ids, matrix]`` per slice) — after whole epochs every shard is home again."""
        out = {}
        for k in range(self.S):
            self._wait_slice(k)
            s = self._visiting_shard(k)
            idx = torch.arange(self.cap, device=self.dev)
            gid = idx * (self.S * self.N) + k * self.N + s
            keep = gid < self.num_cols
            out[k] = (gid[keep], self.Hbuf[k][self.cur[k]][keep])
        return out
