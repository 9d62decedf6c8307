"""MF-SGD engine with model rotation.

Reference skeleton (``ml/java/src/main/java/edu/iu/sgd/SGDCollectiveMapper.java:154-320``): W rows stay
on the worker that owns them, H columns live in ``numModelSlices = 2`` slices per worker that travel
a ring; while slice k is being trained on, slice 1-k is in flight to the next worker
(``dymoro/Rotator.java:53-67``).  One epoch = N rotations x 2 slices, so every H column meets every
worker's ratings exactly once.

Here: column j has home worker ``j % N``, slice ``(j // N) % S`` and slice-local index ``j // SN``; always
``S = 2`` slices in flight like the reference.  On GPUs a whole epoch is ONE cooperative launch
(``csrc/mf.cu::sgd_epoch_kernel``): 2N panel passes back to back, and in every CTA a communication warp pulls
the next use of the other slice from the source rank's symmetric slab over NVLink while the training warps
work (``csrc/rotate_ring.cuh``).  Rotation orders may be randomised like the reference's
(``dymoro/RotationUtil.java:58-90``): the per-step source ranks come from ``parallel.ring.RingSchedule``.
The older path — one launch per panel, rotation as a separate pull kernel on a side stream — is kept for CPU
tensors (gloo) and as the ``fused=False`` A/B arm; the RMSE pass uses it too.
"""
from __future__ import annotations

import math
import os
from typing import Dict, List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist

from ..ops import mf
from ..parallel.ring import RingSchedule, RingState


class MFSGDEngine:
    def __init__(self, rows: torch.Tensor, cols: torch.Tensor, vals: torch.Tensor, num_rows_local: int,
                 num_cols: int, rank: int, eps: float, lam: float, *, symm=None, blocks: Optional[int] = None,
                 seed: int = 0, group=None, num_slices: Optional[int] = None, fused: Optional[bool] = None,
                 cluster: Optional[int] = None, orders: Optional[Sequence[int]] = None, loopback: bool = False,
                 max_run: Optional[int] = None, virtual_world: int = 0, col_blocks: Optional[int] = None):
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
        # Tuning aid (single GPU only): lay the problem out as ONE rank of a `virtual_world`-rank job would see it
        # (N_v x S panels, N_v x S steps per epoch, loop-back ring) so that the per-rank kernel time of a large
        # job can be measured on one GPU.  The slices' contents are not meaningful in this mode.
        self.Nv = int(virtual_world) if (virtual_world and N == 1) else 0
        if self.Nv:
            N = self.Nv
            loopback = True
        cuda = self.dev.type == "cuda"
        self.S = S = int(num_slices) if num_slices else 2        # two slices in flight, like the reference
        assert S in (1, 2)
        if fused is None:
            # one rank has nothing to rotate: two per-panel launches per epoch measured faster there (15.4 vs 19.5 ms
            # on BASELINE config 3) than the epoch kernel, whose cells end with a fence that drains their stores
            env = os.environ.get("HB_SGD_FUSED")
            fused = cuda and ((self.N > 1 and self.symm is not None) or loopback or bool(self.Nv))
            if env is not None:
                fused = cuda and env != "0" and (self.N == 1 or self.symm is not None)
        self.fused = bool(fused)
        self.cap = max(1, (self.num_cols + S * N - 1) // (S * N))          # columns per (shard, slice)
        nnz_max = float(rows.numel())
        if dist.is_initialized() and N > 1:
            t = torch.tensor([nnz_max], dtype=torch.float64, device=self.dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
            nnz_max = float(t.item())
        per_panel = nnz_max / (float(S) * N)
        self.cs = 1
        self.C = None
        if cuda and self.fused:
            # worker = cluster of cs CTAs; B = as many workers as one cooperative launch keeps resident
            self.cs = int(cluster) if cluster else mf.pick_cluster(per_panel)
            mx = mf.epoch_max_workers(self.rank_k, self.cs)
            while mx < 1 and self.cs > 1:
                self.cs //= 2
                mx = mf.epoch_max_workers(self.rank_k, self.cs)
            self.B = int(blocks) if blocks else mx
            self.B = max(1, min(self.B, mx))
            # column-blocks per panel: a few more than workers, so that a block released by one worker has a whole
            # cell of slack before the next one asks for it and the cross-worker hand-over latency leaves the
            # critical path (B workers passing B tokens round a ring pay it at every step)
            self.C = mf.pick_col_blocks(self.B, self.B - col_blocks if (col_blocks is not None and col_blocks < 0) else col_blocks)
            if max_run is None:
                max_run = mf.pick_max_run(per_panel * self.B / self.C, self.B, self.cs, mf.epoch_run_slots(self.rank_k))
        elif cuda:
            mx = mf.max_blocks(self.rank_k)
            sms = torch.cuda.get_device_properties(self.dev).multi_processor_count
            self.B = int(blocks) if blocks else mf.auto_blocks(per_panel, min(mx, 2 * sms))
        else:
            self.B = int(blocks) if blocks else 4
        self.max_run = int(max_run) if max_run else 32
        g = torch.Generator(device=self.dev).manual_seed(seed + 7919 * self.me)
        scale = 1.0 / math.sqrt(self.rank_k)
        dt = torch.float32 if cuda else torch.float64
        self.dt = dt
        self.W = (torch.rand(self.num_rows_local, self.rank_k, device=self.dev, generator=g, dtype=dt) * scale).contiguous()
        # H slices: control block + ping-pong buffers, one allocation (symmetric memory when there are peers)
        self.ring = None
        if cuda:
            self.ring = RingState(self.dev, self.cap * self.rank_k, dt, S, symm=self.symm, rank=self.me, world=self.N,
                                  loopback=loopback)
            self.Hbuf = [[b.view(self.cap, self.rank_k) for b in pair] for pair in self.ring.buf]
        else:
            self.Hbuf = [[torch.empty(self.cap, self.rank_k, dtype=dt, device=self.dev) for _ in range(2 if N > 1 else 1)]
                         for _ in range(S)]
        self._cur = [0] * S
        self.sched = [RingSchedule(self.N, orders) for _ in range(S)]   # every slice walks the same order array
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
                per.append(mf.build_panel(rows[m], local[m], vals[m], self.B, max_run=self.max_run, C=self.C))
            self.panels.append(per)
        self.trained = torch.zeros(1, dtype=torch.int64, device=self.dev)
        self.epochs_fused = 0
        if cuda:
            self.desc = mf.panel_descriptors([self.panels[s][k] for s in range(N) for k in range(S)])
            self.comm_stream = torch.cuda.Stream(device=self.dev)
            self.ev_compute = [torch.cuda.Event(), torch.cuda.Event()]
            self.ev_comm = [torch.cuda.Event(), torch.cuda.Event()]
            for e in self.ev_comm:
                e.record()
            torch.cuda.synchronize(self.dev)
        self.comm_time_hidden = True
        if dist.is_initialized() and N > 1:
            dist.barrier(group=group)       # nobody starts a device collective before every rank's state exists

    # cur parity lives in the ring state when there is one
    @property
    def cur(self) -> List[int]:
        return self.ring.cur if self.ring is not None else self._cur

    @property
    def rot(self) -> List[int]:
        return [s.rotations for s in self.sched]

    # ------------------------------------------------------------------------------------
    def _visiting_shard(self, k: int) -> int:
        return self.sched[k].shard_at[self.me]

    def _rotate(self, k: int) -> None:
        """slice k moves on: every worker pulls the copy of the worker the rotation map names as its source."""
        if self.N == 1:
            self.sched[k].rotate()
            return
        src = self.Hbuf[k][self.cur[k]]
        dst = self.Hbuf[k][1 - self.cur[k]]
        srcs = self.sched[k].rotate()
        pred = srcs[self.me]
        succ = srcs.index(self.me)
        if self.symm is not None and self.dev.type == "cuda":
            # on the stream of every other collective of this workspace (its flags / tickets are per workspace);
            # the per-panel path is the A/B arm, the whole-epoch kernel is what overlaps the transfer
            self.symm.rotate_(dst, src, pred)
        else:
            # gloo / NCCL send-recv ring (baseline path and CPU tests)
            send = src.contiguous()
            ops = [dist.P2POp(dist.isend, send, succ, group=self.group),
                   dist.P2POp(dist.irecv, dst, pred, group=self.group)]
            for r in dist.batch_isend_irecv(ops):
                r.wait()
        self.cur[k] ^= 1

    def _wait_slice(self, k: int) -> None:
        if self.N > 1 and self.symm is not None and self.dev.type == "cuda":
            torch.cuda.current_stream().wait_event(self.ev_comm[k])

    def _epoch_fused(self, deadline_ms: Optional[float]) -> None:
        """The whole epoch as one launch; the per-step panel / source-rank tables come from the schedules."""
        N, S = self.N, self.S
        panel_of_step, src_of_step = [], []
        if self.Nv:
            N = self.Nv
            panel_of_step = [j * S + k for j in range(N) for k in range(S)]
            src_of_step = [0] * (N * S)
        else:
            for _ in range(N):
                for k in range(S):
                    panel_of_step.append(self.sched[k].shard_at[self.me] * S + k)
                    src_of_step.append(self.sched[k].rotate()[self.me])
        for k in range(S):
            self._wait_slice(k)
        mf.sgd_epoch(self.W, self.desc, panel_of_step, S, self.B, self.cs | ((self.C or self.B) << 8), self.eps, self.lam, self.epochs_fused,
                     deadline_ms, self.trained, self.ring, src_of_step)
        self.epochs_fused += 1
        self.ring.advance(N)
        tb = (self.epochs_fused * self.B) & 0x7FFFFFFF
        for per in self.panels:
            for p in per:
                p.ticket_base = tb

    def epoch(self, deadline_ms: Optional[float] = None) -> torch.Tensor:
        """N rotations x S slices (``SGDCollectiveMapper.java:263-277``). Returns the trained-rating counter."""
        self.trained.zero_()
        if self.fused and self.dev.type == "cuda":
            self._epoch_fused(deadline_ms)
            return self.trained
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

    def close(self) -> None:
        """Gives the engine's block of the symmetric workspace back (call on every rank, in the same order as other
        allocations / frees); the engine is unusable afterwards."""
        self.synchronize()
        if self.ring is not None:
            self.ring.release()

    def rmse(self, rows, cols, vals) -> float:
        se, n = self.sq_err(rows, cols, vals)
        return math.sqrt(se / max(n, 1))

    def synchronize(self) -> None:
        if self.dev.type == "cuda":
            for k in range(self.S):
                self._wait_slice(k)
            torch.cuda.synchronize(self.dev)
            if self.ring is not None and self.fused:
                self.ring.check()

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
