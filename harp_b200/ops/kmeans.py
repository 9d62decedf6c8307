"""K-means assignment / accumulation op.

CUDA tensors run the hand-written sm_100a kernels in ``csrc/kmeans.cu``
(tcgen05 TF32 distance GEMM + TMEM arg-min epilogue + L2 vector reductions).
CPU tensors run a plain PyTorch restatement of the reference's hot loop
(``ml/java/.../kmeans/regroupallgather/CenCalcTask.java:67-100``) so that the
multi-process semantics can be tested without a GPU.
"""
from __future__ import annotations

import os
from dataclasses import dataclass
from typing import Optional

import torch

from .. import _native


@dataclass
class KMeansStepResult:
    sums: torch.Tensor      # [k, d] fp32: Σ of the points assigned to each centroid
    counts: torch.Tensor    # [k] fp32
    assign: Optional[torch.Tensor]  # [n] int32 or None
    obj: Optional[torch.Tensor]     # [1] float64: Σ min squared distance, or None


def workspace_bytes(n: int, k: int, d: int, simt: bool = False) -> int:
    lib = _native.lib()
    if simt or not lib.hb_kmeans_tc_supported(d, k):
        return int(lib.hb_kmeans_simt_workspace_bytes(n, k, d))
    return int(lib.hb_kmeans_workspace_bytes(k, d))


def tc_supported(d: int, k: int) -> bool:
    """Whether the tcgen05 kernels (and so the fused one-launch iteration) take this shape."""
    return bool(_native.lib().hb_kmeans_tc_supported(int(d), int(k)))


def _cpu_step(X, C, sums, counts, assign, obj, zero):
    # exact (x-c)^2 in the tensor's own precision, lowest index wins ties
    d2 = torch.cdist(X.double(), C.double()).pow(2) if X.shape[0] * C.shape[0] <= (1 << 24) else None
    if d2 is None:
        best = torch.empty(X.shape[0], dtype=torch.int64)
        bestd = torch.empty(X.shape[0], dtype=torch.float64)
        for s in range(0, X.shape[0], 65536):
            blk = torch.cdist(X[s:s + 65536].double(), C.double()).pow(2)
            bestd[s:s + 65536], best[s:s + 65536] = blk.min(dim=1)
    else:
        bestd, best = d2.min(dim=1)
    if zero:
        sums.zero_()
        counts.zero_()
        if obj is not None:
            obj.zero_()
    sums.index_add_(0, best, X.to(sums.dtype))
    counts.index_add_(0, best, torch.ones_like(best, dtype=counts.dtype))
    if assign is not None:
        assign.copy_(best.to(assign.dtype))
    if obj is not None:
        obj += bestd.sum()


def kmeans_step(
    X: torch.Tensor,
    C: torch.Tensor,
    sums: Optional[torch.Tensor] = None,
    counts: Optional[torch.Tensor] = None,
    assign: Optional[torch.Tensor] = None,
    obj: Optional[torch.Tensor] = None,
    workspace: Optional[torch.Tensor] = None,
    *,
    want_assign: bool = False,
    want_obj: bool = False,
    accumulate: bool = False,
    force_simt: bool = False,
    variant: str = "auto",
    faithful: bool = False,
) -> KMeansStepResult:
    """One pass of Lloyd's assignment over ``X[n,d]`` against ``C[k,d]``.

    ``faithful=True`` (what the application mappers pass) keeps fp32 arithmetic wherever that is affordable: problems
    up to 2^32 multiply-adds per pass take the exact SIMT kernel (milliseconds), larger ones the TF32 tensor-core kernel.
    The reference computes these distances in fp64, and on data far from the origin a TF32 product (10-bit mantissa)
    can flip assignments between centroids whose distances differ by less than ~0.1 %, after which two variants of the
    same algorithm no longer produce the same centroids.  The engines choose their precision explicitly."""
    assert X.dim() == 2 and C.dim() == 2 and X.shape[1] == C.shape[1]
    n, d = X.shape
    k = C.shape[0]
    dev = X.device
    if sums is None:
        acc_dt = torch.float32 if dev.type == "cuda" else X.dtype     # CPU path keeps the input precision
        sums = torch.empty(k, d, dtype=acc_dt, device=dev)
        counts = torch.empty(k, dtype=acc_dt, device=dev)
        if accumulate:
            sums.zero_(); counts.zero_()
    if want_assign and assign is None:
        assign = torch.empty(n, dtype=torch.int32, device=dev)
    if want_obj and obj is None:
        obj = torch.zeros(1, dtype=torch.float64, device=dev)

    if dev.type != "cuda":
        _cpu_step(X, C, sums, counts, assign, obj, zero=not accumulate)
        return KMeansStepResult(sums, counts, assign, obj)

    lib = _native.lib()
    assert X.dtype == torch.float32 and C.dtype == torch.float32, "CUDA K-means path is fp32/TF32"
    X = X.contiguous()
    C = C.contiguous()
    simt = force_simt or not lib.hb_kmeans_tc_supported(d, k) or (faithful and n * k * d <= (1 << 32))
    need = workspace_bytes(n, k, d, simt)
    if workspace is None or workspace.numel() * workspace.element_size() < need:
        workspace = torch.empty((need + 255) // 256 * 256, dtype=torch.uint8, device=dev)
    # variant: "auto"/"pair" = CTA-pair tcgen05 kernel (‖c‖² folded into the GEMM), "pair_noext" = CTA-pair with
    # ‖c‖² added in the epilogue, "single" = 1-CTA tcgen05 kernel
    flags = (1 if simt else 0) | (2 if accumulate else 0) | (4 if variant == "single" else 0) | (8 if variant == "pair_noext" else 0) | (16 if variant == "pair_x128" else 0)
    flags |= (int(os.environ.get("HB_KMEANS_DBG", "0")) & 0xff) << 8   # perf-triage switches (pair kernel)
    rc = lib.hb_kmeans_step(
        _native.ptr(X), n, d, _native.ptr(C), k, _native.ptr(sums), _native.ptr(counts),
        _native.ptr(assign), _native.ptr(obj), _native.ptr(workspace),
        workspace.numel() * workspace.element_size(), flags, _native.stream_ptr(),
    )
    _native.check(rc, "hb_kmeans_step")
    _native.count_launch(3 if simt else 2)
    return KMeansStepResult(sums, counts, assign, obj)


def kmeans_finalize(sums: torch.Tensor, counts: torch.Tensor, C_out: torch.Tensor, keep_empty: bool = False) -> torch.Tensor:
    """``C_out[j] = sums[j] / counts[j]``; empty clusters → zero vector (or untouched)."""
    k, d = sums.shape
    if sums.device.type != "cuda":
        nz = counts > 0
        if keep_empty:
            C_out[nz] = (sums[nz] / counts[nz, None]).to(C_out.dtype)
        else:
            C_out.zero_()
            C_out[nz] = (sums[nz] / counts[nz, None]).to(C_out.dtype)
        return C_out
    lib = _native.lib()
    rc = lib.hb_kmeans_finalize(_native.ptr(sums), _native.ptr(counts), k, d, _native.ptr(C_out),
                                1 if keep_empty else 0, _native.stream_ptr())
    _native.check(rc, "hb_kmeans_finalize")
    _native.count_launch(1)
    return C_out


class FusedKMeans:
    """One-launch-per-iteration K-means (``csrc/kmeans.cu``: ``assign_tc_pair_kernel<0,1,1>`` + ``fused_tail``).

    The launch that runs the tcgen05 distance GEMM, arg-min and partial-sum accumulation also does Harp's
    regroup → centroid update → allgather over NVLink peer memory
    (``ml/java/.../kmeans/regroupallgather/KMeansCollectiveMapper.java:168-185``) and prepares the GEMM operands
    of the next iteration.  ``state`` is this rank's block of a symmetric slab (same offset on every rank);
    ``peer_ptrs`` are the device addresses of that block on every rank (for one GPU: just its own).
    """

    PREC = {"tf32": 0, "fp16x3": 1, "bf16x3": 2}

    def __init__(self, state: torch.Tensor, peer_ptrs, rank: int, k: int, d: int, keep_empty: bool = False,
                 precision: str = "tf32", scale: float = 1.0):
        import ctypes

        self.prec = self.PREC[precision]
        self.precision = precision
        self.scale = float(scale) if self.prec == 1 else 1.0

        lib = _native.lib()
        lib.hb_kmeans_fused_state_bytes.restype = ctypes.c_size_t
        lib.hb_kmeans_fused_state_bytes.argtypes = [ctypes.c_int, ctypes.c_int]
        lib.hb_kmeans_fused_offsets.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_void_p]
        lib.hb_kmeans_fused_prep.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_void_p]
        lib.hb_kmeans_fused_init.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p]
        lib.hb_kmeans_fused_step.argtypes = [ctypes.c_void_p, ctypes.c_longlong, ctypes.c_int, ctypes.c_int, ctypes.c_void_p,
                                             ctypes.c_int, ctypes.c_int, ctypes.c_uint, ctypes.c_int, ctypes.c_void_p,
                                             ctypes.c_void_p]
        lib.hb_kmeans_fused_prep2.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                              ctypes.c_float, ctypes.c_void_p]
        lib.hb_kmeans_fused_step2.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_longlong, ctypes.c_int, ctypes.c_int,
                                              ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_uint, ctypes.c_int,
                                              ctypes.c_void_p, ctypes.c_int, ctypes.c_float, ctypes.c_void_p]
        lib.hb_kmeans_split_row_bytes.restype = ctypes.c_size_t
        lib.hb_kmeans_split_row_bytes.argtypes = [ctypes.c_int]
        lib.hb_kmeans_split_rows.argtypes = [ctypes.c_void_p, ctypes.c_longlong, ctypes.c_int, ctypes.c_float, ctypes.c_int,
                                             ctypes.c_void_p, ctypes.c_void_p]
        self.lib = lib
        self.k, self.d = int(k), int(d)
        self.rank, self.world = int(rank), len(peer_ptrs)
        self.keep_empty = bool(keep_empty)
        need = self.state_bytes(k, d)
        assert state.dtype == torch.uint8 and state.numel() >= need and state.data_ptr() % 256 == 0
        assert int(peer_ptrs[rank]) == state.data_ptr()
        self.state = state
        self._peers = (ctypes.c_void_p * self.world)(*[ctypes.c_void_p(int(p)) for p in peer_ptrs])
        off = (ctypes.c_ulonglong * 8)()
        _native.check(lib.hb_kmeans_fused_offsets(k, d, ctypes.cast(off, ctypes.c_void_p)), "hb_kmeans_fused_offsets")
        o = [int(v) for v in off]
        kd = k * d
        self.C = [state[o[i]: o[i] + kd * 4].view(torch.float32).view(k, d) for i in range(2)]
        self.acc = [state[o[2 + i]: o[2 + i] + (kd + k) * 4].view(torch.float32) for i in range(2)]
        self.obj_terms = state[o[4]: o[4] + 16].view(torch.float64)
        self.obj_total = state[o[5]: o[5] + 8].view(torch.float64)
        self.epoch = 0

    @staticmethod
    def state_bytes(k: int, d: int) -> int:
        import ctypes

        lib = _native.lib()
        lib.hb_kmeans_fused_state_bytes.restype = ctypes.c_size_t
        lib.hb_kmeans_fused_state_bytes.argtypes = [ctypes.c_int, ctypes.c_int]
        return int(lib.hb_kmeans_fused_state_bytes(k, d))

    @property
    def centroids(self) -> torch.Tensor:
        """The centroid buffer the next step will read (= the result of the previous one)."""
        return self.C[self.epoch & 1]

    def init(self, C0: Optional[torch.Tensor] = None) -> None:
        """Zero the state and load ``C0``.  The caller must barrier across ranks before the first ``step``."""
        if C0 is not None:
            C0 = C0.to(torch.float32).contiguous()
        _native.check(self.lib.hb_kmeans_fused_init(_native.ptr(self.state), self.k, self.d, _native.ptr(C0),
                                                    _native.stream_ptr()), "hb_kmeans_fused_init")
        _native.count_launch(2)
        self.epoch = 0
        if self.prec:
            for par in (1, 0):
                _native.check(self.lib.hb_kmeans_fused_prep2(_native.ptr(self.state), self.k, self.d, par, self.prec,
                                                             self.scale, _native.stream_ptr()), "hb_kmeans_fused_prep2")
            _native.count_launch(2)

    def set_centroids(self, C: torch.Tensor, non_blocking: bool = True) -> None:
        """Overwrite the current centroids (every rank must do the same) and rebuild the GEMM operands."""
        self.centroids.copy_(C.view(self.k, self.d), non_blocking=non_blocking)
        _native.check(self.lib.hb_kmeans_fused_prep2(_native.ptr(self.state), self.k, self.d, self.epoch & 1, self.prec,
                                                     self.scale, _native.stream_ptr()), "hb_kmeans_fused_prep2")
        _native.count_launch(1)

    def split_rows(self, X: torch.Tensor) -> torch.Tensor:
        """The points as 16-bit hi/lo rows for the split-precision GEMM (``hb_kmeans_split_rows``); done once per
        data set, the rows never change."""
        assert self.prec and X.dtype == torch.float32 and X.is_contiguous() and X.shape[1] == self.d
        rb = int(self.lib.hb_kmeans_split_row_bytes(self.d))
        Xs = torch.empty(X.shape[0], rb // 2, dtype=torch.int16, device=X.device)
        _native.check(self.lib.hb_kmeans_split_rows(_native.ptr(X), X.shape[0], self.d, self.scale, self.prec,
                                                    _native.ptr(Xs), _native.stream_ptr()), "hb_kmeans_split_rows")
        _native.count_launch(1)
        return Xs

    def step(self, X: torch.Tensor, assign: Optional[torch.Tensor] = None, Xs: Optional[torch.Tensor] = None) -> torch.Tensor:
        """One full Lloyd iteration in one launch; returns the new centroids (a view of the state)."""
        assert X.dtype == torch.float32 and X.is_contiguous() and X.shape[1] == self.d
        assert (Xs is not None) == bool(self.prec) or X.shape[0] == 0, "split-precision steps need the split rows"
        rc = self.lib.hb_kmeans_fused_step2(_native.ptr(X), _native.ptr(Xs), X.shape[0], self.d, self.k,
                                            ctypes_cast_void(self._peers), self.rank, self.world, self.epoch,
                                            1 if self.keep_empty else 0, _native.ptr(assign), self.prec, self.scale,
                                            _native.stream_ptr())
        _native.check(rc, "hb_kmeans_fused_step2")
        _native.count_launch(1)
        self.epoch += 1
        return self.centroids

    @property
    def last_sums(self) -> torch.Tensor:
        """[sums | counts] of this rank from the last step (valid until the step after next)."""
        return self.acc[(self.epoch - 1) & 1]


def ctypes_cast_void(arr):
    import ctypes

    return ctypes.cast(arr, ctypes.c_void_p)
