"""K-means engine: the device-resident iteration used by every K-means launcher and by bench.py.

One iteration = assignment + partial-sum accumulation (``ops.kmeans``) →
cross-worker combination of ``[sums | counts]`` → centroid update.  The
cross-worker step is selectable:

``comm="nccl"``     plain ``dist.all_reduce`` — the baseline build (SURVEY §5.9 column 2)
``comm="p2p"``      the in-kernel NVLink peer-memory reduce-scatter + all-gather
                    (``csrc/coll.cu``; Harp's regroup→allgather shape,
                    ``ml/java/.../kmeans/regroupallgather/KMeansCollectiveMapper.java:168-185``)
``comm="fused"``    ONE launch per iteration: the tcgen05 distance/arg-min/accumulate kernel also does the
                    regroup → update → allgather over NVLink and prepares the next iteration's operands
                    (``ops.kmeans.FusedKMeans``); works on one GPU too (3 launches become 1)
``comm="auto"``     fused on CUDA when the shape is supported (and, for >1 rank, a symmetric workspace is
                    given), else p2p / nccl

Reference semantics kept: an empty cluster collapses to the zero vector
(``KMeansCollectiveMapper.java:170-183``) unless ``keep_empty=True``.
"""
from __future__ import annotations

import math
import os
from typing import Optional

import torch
import torch.distributed as dist

from ..ops import kmeans as kops


class KMeansEngine:
    def __init__(self, X: torch.Tensor, k: int, *, comm: str = "auto", group=None,
                 keep_empty: bool = False, symm=None, precision: Optional[str] = None, local: bool = False):
        assert X.dim() == 2
        self.X = X.contiguous() if X.device.type == "cuda" else X
        self.n, self.d = X.shape
        self.k = int(k)
        self.dev = X.device
        self.group = group
        self.keep_empty = keep_empty
        # local=True: a single-GPU engine even inside a multi-rank job (no collective, e.g. to score a sample)
        self.world = dist.get_world_size(group) if (dist.is_available() and dist.is_initialized() and not local) else 1
        self.rank = dist.get_rank(group) if self.world > 1 else 0
        self.symm = symm
        self.variant = os.environ.get("HB_KMEANS_VARIANT", "auto")
        k, d = self.k, self.d
        if comm == "auto":
            fused_ok = (self.dev.type == "cuda" and X.dtype == torch.float32 and kops.tc_supported(d, k)
                        and self.variant in ("auto", "pair") and (self.world == 1 or symm is not None))
            comm = "fused" if fused_ok else ("p2p" if (symm is not None and self.world > 1) else "nccl")
        self.comm = comm
        self.fused = None
        # Products of the distance GEMM (fused kernel only): "tf32" = fp32 inputs truncated to 10 mantissa bits by the
        # tensor core; "fp16x3" / "bf16x3" = every value split into two 16-bit halves and three MMAs per product
        # (fp32-faithful with fp16 halves, see csrc/kmeans.cu).  HB_KMEANS_PRECISION overrides the default.
        self.precision = precision or os.environ.get("HB_KMEANS_PRECISION", "tf32")
        if self.comm != "fused":
            self.precision = "tf32" if self.dev.type == "cuda" else "exact"
        self.Xs = None
        self._symm_allocs = []          # what close() hands back to the workspace
        self._symm_ws = symm
        if self.comm == "fused":
            nbytes = kops.FusedKMeans.state_bytes(k, d)
            if self.world > 1:
                assert symm is not None, "comm='fused' on several ranks needs a SymmetricWorkspace"
                st = symm.alloc(nbytes, torch.uint8)
                self._symm_allocs.append(st)
                off = symm.offset_of(st)
                peers = [p + off for p in symm.peer_ptrs]
            else:
                st = torch.empty(nbytes, dtype=torch.uint8, device=self.dev)
                peers = [st.data_ptr()]
            scale = 1.0
            if self.precision == "fp16x3":
                # power of two that brings the largest |x| of the whole data set to [2^13, 2^14): halves stay inside
                # fp16's range, and centroids (means of points) can never exceed it
                mx = self.X.abs().max().to(torch.float64) if self.n else torch.zeros((), dtype=torch.float64, device=self.dev)
                if self.world > 1:
                    dist.all_reduce(mx, op=dist.ReduceOp.MAX, group=group)
                mxv = float(mx.item())
                scale = 2.0 ** (13 - math.floor(math.log2(mxv))) if mxv > 0 else 1.0
            self.fused = kops.FusedKMeans(st, peers, self.rank, k, d, keep_empty, precision=self.precision, scale=scale)
            self.fused.init(None)
            if self.fused.prec:
                self.Xs = self.fused.split_rows(self.X)
            if self.world > 1:
                torch.cuda.synchronize(self.dev)
                dist.barrier(group=group)          # nobody signals a peer whose flags are not zeroed yet
            self.obj = self.fused.obj_total
            self.workspace = None
            self.total_points = float(self.n)
            if self.world > 1:
                t = torch.tensor([float(self.n)], dtype=torch.float64, device=self.dev)
                dist.all_reduce(t, group=group)
                self.total_points = float(t.item())
            return
        wdt = torch.float32 if self.dev.type == "cuda" else X.dtype      # the CPU path keeps the input precision
        # one slab [sums | counts] so the cross-worker combine is a single collective
        if self.comm == "p2p" and self.world > 1:
            assert symm is not None, "comm='p2p' needs a SymmetricWorkspace"
            self.acc = symm.alloc(k * d + k, torch.float32)
            self._symm_allocs.append(self.acc)
        else:
            self.acc = torch.zeros(k * d + k, dtype=wdt, device=self.dev)
        self._sums = self.acc[: k * d].view(k, d)
        self._counts = self.acc[k * d:]
        if self.comm == "p2p" and self.world > 1:
            self.obj = symm.alloc(2, torch.float64)[:1]
            self.obj.zero_()
            self._C = symm.alloc(k * d, torch.float32).view(k, d)
            self._C.zero_()
            self._symm_allocs += [self.obj, self._C]
        else:
            self.obj = torch.zeros(1, dtype=torch.float64, device=self.dev)
            self._C = torch.zeros(k, d, dtype=wdt, device=self.dev)
        self.workspace = None
        if self.dev.type == "cuda":
            need = kops.workspace_bytes(self.n, k, d)
            self.workspace = torch.empty((need + 255) // 256 * 256, dtype=torch.uint8, device=self.dev)
        self.total_points = float(self.n)
        if self.world > 1:
            t = torch.tensor([float(self.n)], dtype=torch.float64, device=self.dev)
            dist.all_reduce(t, group=group)
            self.total_points = float(t.item())

    @property
    def C(self) -> torch.Tensor:
        """Current centroids ``[k, d]`` (device)."""
        return self.fused.centroids if self.fused is not None else self._C

    @property
    def sums(self) -> torch.Tensor:
        if self.fused is not None:
            return self.fused.last_sums[: self.k * self.d].view(self.k, self.d)
        return self._sums

    @property
    def counts(self) -> torch.Tensor:
        if self.fused is not None:
            return self.fused.last_sums[self.k * self.d:]
        return self._counts

    def set_centroids(self, C: torch.Tensor, non_blocking: bool = True) -> None:
        if self.fused is not None:
            self.fused.set_centroids(C, non_blocking=non_blocking)
        else:
            self._C.copy_(C.view(self.k, self.d), non_blocking=non_blocking)

    def close(self) -> None:
        """Gives the engine's blocks of the symmetric workspace back (every rank, same order); unusable afterwards."""
        if self._symm_allocs:
            torch.cuda.synchronize(self.dev)
            for t in reversed(self._symm_allocs):
                self._symm_ws.free(t)
            self._symm_allocs = []

    def step(self, want_obj: bool = True) -> Optional[torch.Tensor]:
        """One Lloyd iteration; returns the (device) Σ min-distance² of the *input* centroids."""
        if self.fused is not None:
            self.fused.step(self.X, Xs=self.Xs)
            return self.obj if want_obj else None
        kops.kmeans_step(self.X, self.C, self.sums, self.counts, None,
                         self.obj if want_obj else None, self.workspace, variant=self.variant)
        if self.world > 1 and self.comm == "p2p" and self.d % 4 != 0:
            self.symm.allreduce_sum_(self.acc)
            if want_obj:
                self.symm.allreduce_sum_(self.obj)
        elif self.world > 1 and self.comm == "p2p":
            # regroup -> divide -> allgather fused in one NVLink kernel per rank
            self.symm.kmeans_regroup_allgather_(self.acc, self.C, self.k, self.d, self.keep_empty)
            if want_obj:
                self.symm.allreduce_sum_(self.obj)
            return self.obj if want_obj else None
        elif self.world > 1:
            dist.all_reduce(self.acc, group=self.group)
            if want_obj:
                dist.all_reduce(self.obj, group=self.group)
        kops.kmeans_finalize(self.sums, self.counts, self.C, keep_empty=self.keep_empty)
        return self.obj if want_obj else None

    def mse(self) -> float:
        return float(self.obj.item()) / max(self.total_points, 1.0)
