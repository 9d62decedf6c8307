"""K-means engine: the device-resident iteration used by every K-means launcher and by bench.py.

One iteration = assignment + partial-sum accumulation (``ops.kmeans``) →
cross-worker combination of ``[sums | counts]`` → centroid update.  The
cross-worker step is selectable:

``comm="nccl"``     plain ``dist.all_reduce`` — the baseline build (SURVEY §5.9 column 2)
``comm="p2p"``      the in-kernel NVLink peer-memory reduce-scatter + all-gather
                    (``csrc/coll.cu``; Harp's regroup→allgather shape,
                    ``ml/java/.../kmeans/regroupallgather/KMeansCollectiveMapper.java:168-185``)
``comm="auto"``     p2p when a symmetric workspace is available, else nccl

Reference semantics kept: an empty cluster collapses to the zero vector
(``KMeansCollectiveMapper.java:170-183``) unless ``keep_empty=True``.
"""
from __future__ import annotations

import os
from typing import Optional

import torch
import torch.distributed as dist

from ..ops import kmeans as kops


class KMeansEngine:
    def __init__(self, X: torch.Tensor, k: int, *, comm: str = "auto", group=None,
                 keep_empty: bool = False, symm=None):
        assert X.dim() == 2
        self.X = X
        self.n, self.d = X.shape
        self.k = int(k)
        self.dev = X.device
        self.group = group
        self.keep_empty = keep_empty
        self.world = dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if self.world > 1 else 0
        self.symm = symm
        self.variant = os.environ.get("HB_KMEANS_VARIANT", "auto")
        if comm == "auto":
            comm = "p2p" if (symm is not None and self.world > 1) else "nccl"
        self.comm = comm
        k, d = self.k, self.d
        # one slab [sums | counts] so the cross-worker combine is a single collective
        if self.comm == "p2p" and self.world > 1:
            assert symm is not None, "comm='p2p' needs a SymmetricWorkspace"
            self.acc = symm.alloc(k * d + k, torch.float32)
        else:
            self.acc = torch.zeros(k * d + k, dtype=torch.float32, device=self.dev)
        self.sums = self.acc[: k * d].view(k, d)
        self.counts = self.acc[k * d:]
        if self.comm == "p2p" and self.world > 1:
            self.obj = symm.alloc(2, torch.float64)[:1]
            self.obj.zero_()
            self.C = symm.alloc(k * d, torch.float32).view(k, d)
            self.C.zero_()
        else:
            self.obj = torch.zeros(1, dtype=torch.float64, device=self.dev)
            self.C = torch.zeros(k, d, dtype=torch.float32, device=self.dev)
        self.workspace = None
        if self.dev.type == "cuda":
            need = kops.workspace_bytes(self.n, k, d)
            self.workspace = torch.empty((need + 255) // 256 * 256, dtype=torch.uint8, device=self.dev)
        self.total_points = float(self.n)
        if self.world > 1:
            t = torch.tensor([float(self.n)], dtype=torch.float64, device=self.dev)
            dist.all_reduce(t, group=group)
            self.total_points = float(t.item())

    def set_centroids(self, C: torch.Tensor, non_blocking: bool = True) -> None:
        self.C.copy_(C.view(self.k, self.d), non_blocking=non_blocking)

    def step(self, want_obj: bool = True) -> Optional[torch.Tensor]:
        """One Lloyd iteration; returns the (device) Σ min-distance² of the *input* centroids."""
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
