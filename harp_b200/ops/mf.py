"""Matrix-factorisation ops: SGD panel pass, squared error, CCD++ rank-one update.

CUDA tensors run ``csrc/mf.cu``; CPU tensors run restatements of the Java loops
(``SGDMPTask.doRun``, ``RMSETask``, ``CCDMPTask``) so that apps and tests work without a GPU.
"""
from __future__ import annotations

import ctypes
from dataclasses import dataclass
from typing import Optional

import torch

from .. import _native


def _declare(lib) -> None:
    if getattr(lib, "_hb_mf_declared", False):
        return
    P, I, LL, F, U, ULL = (ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong, ctypes.c_float, ctypes.c_uint32,
                           ctypes.c_ulonglong)
    lib.hb_sgd_max_blocks.argtypes = [I]
    lib.hb_sgd_pass.argtypes = [P, P, P, P, P, P, P, I, I, F, F, P, U, ULL, P, P]
    lib.hb_device_time_ns.argtypes = [P, P]
    lib.hb_sgd_pass_dynamic.argtypes = [P, P, P, P, P, P, P, I, I, I, F, F, P, ctypes.c_ulonglong, P, P]
    lib.hb_mf_sq_err.argtypes = [P, P, P, P, P, LL, I, P, P]
    lib.hb_ccd_update.argtypes = [P, P, P, P, P, P, I, F, P]
    lib.hb_ccd_sweep.argtypes = [P, P, P, P, P, P, P, P, I, F, I, P]
    lib.hb_sgd_epoch_max_workers.argtypes = [I, I]
    lib.hb_sgd_epoch.argtypes = [P, P, I, P, I, I, I, I, I, F, F, U, ULL, P,
                                 P, I, I, ULL, P, P, P, U, P, ULL, P]
    lib._hb_mf_declared = True


@dataclass
class SgdPanel:
    """Ratings of (this worker's rows) x (one H slice), bucketed B x C, column-sorted inside blocks."""
    B: int
    blk_seg_ptr: torch.Tensor   # int32 [B*C+1]
    seg_col: torch.Tensor       # int32 [nseg]
    seg_ptr: torch.Tensor       # int32 [nseg+1]
    r_row: torch.Tensor         # int32 [nnz]
    r_val: torch.Tensor         # float32 [nnz]
    col_ticket: torch.Tensor    # uint32-as-int32 [C]: ring tickets, or (dynamic schedule) locks that rest at 0
    ticket_base: int = 0
    C: int = 0                  # column-blocks; 0 / == B -> Latin-square ticket ring, > B -> dynamic schedule
    extra: Optional[torch.Tensor] = None   # caller's per-rating payload, permuted like the ratings

    @property
    def ncolblocks(self) -> int:
        return self.C if self.C else self.B

    @property
    def nnz(self) -> int:
        return int(self.r_row.numel())


def max_blocks(rank: int) -> int:
    lib = _native.lib()
    _declare(lib)
    return int(lib.hb_sgd_max_blocks(rank))


def epoch_max_workers(rank: int, cs: int) -> int:
    """Workers (clusters of ``cs`` CTAs) the whole-epoch kernel keeps resident on this GPU."""
    lib = _native.lib()
    _declare(lib)
    return int(lib.hb_sgd_epoch_max_workers(rank, cs))


def pick_cluster(nnz_per_panel: float, kind: str = "sgd") -> int:
    """CTAs per worker for the whole-epoch kernels.  A pass costs B hand-overs (each ends with a gpu-scope fence that
    has to drain the cell's stores: ~5 us measured) plus nnz / (B x warps per worker) updates, so few, fat workers
    win at every size tried: clusters of 8 (15 workers on a B200: two 8-SM clusters fit per GPC) beat 4, 2 and 1 at
    781 K, 3 M and 12 M ratings per panel.  For small panels even clusters of 16 (non-portable size; only 7 fit, 112
    of 148 SMs) win, because 7 hand-overs per pass instead of 15 outweigh the idle SMs: SGD at 781 K ratings per panel
    (8 ranks) 3.33 -> 2.97 ms per epoch, LDA at 781 K / 3.1 M tokens per panel 13.3 -> 12.5 / 21.6 -> 20.6 ms; at
    3.1 M ratings (SGD) and 12.5 M (both) clusters of 8 are ahead again (``scripts/tune_*_virtual.py``,
    ``profiles/EXPERIMENTS.md``).  ``HB_EPOCH_CS`` overrides."""
    import os
    if os.environ.get("HB_EPOCH_CS"):
        return int(os.environ["HB_EPOCH_CS"])
    small = 1.5e6 if kind == "sgd" else 5e6
    return 16 if nnz_per_panel < small else 8


def pick_col_blocks(workers: int, col_blocks: Optional[int] = None) -> int:
    """Column-blocks per panel for the whole-epoch kernels (>= workers).  ``HB_EPOCH_C`` overrides: an absolute count,
    or ``+k`` for workers + k."""
    import os
    env = os.environ.get("HB_EPOCH_C")
    if col_blocks is None and env:
        col_blocks = workers + int(env[1:]) if env.startswith("+") else int(env)
    if col_blocks is None:
        col_blocks = workers
    return max(workers, min(int(col_blocks), 255))


def epoch_run_slots(rank: int) -> int:
    """Column runs one CTA of the whole-epoch kernel trains at the same time: 16 warps x 4 runs with the quad-run
    cell (rank <= 128), 31 warps x 1 run otherwise (``csrc/mf.cu``; ``HB_SGD_QUAD=0`` switches the quad cell off)."""
    import os
    return 64 if (rank <= 128 and os.environ.get("HB_SGD_QUAD", "1") != "0") else 31


def pick_max_run(nnz_per_panel: float, workers: int, cs: int, slots_per_cta: int = 31) -> int:
    """Longest column run trained sequentially: short enough that a cell has ~2 runs per run slot of its worker."""
    cell = nnz_per_panel / float(workers * workers)
    x = cell / (2.0 * cs * slots_per_cta)
    if slots_per_cta == 64:          # quad-run cell: eight ratings per batch, so whole multiples of 8 (measured)
        return 8 if x < 12 else (16 if x < 24 else 32)
    return int(max(4, min(32, x)))


def panel_descriptors(panels):
    """Host array of the 6-pointer descriptors (``SgdPanelDesc``) that ``hb_sgd_epoch`` copies into the kernel's
    parameter block."""
    flat = [t.data_ptr() for p in panels for t in (p.blk_seg_ptr, p.seg_col, p.seg_ptr, p.r_row, p.r_val, p.col_ticket)]
    return (ctypes.c_uint64 * len(flat))(*flat)


def sgd_epoch(W: torch.Tensor, desc, panel_of_step, S: int, workers: int,
              cs: int, eps: float, lam: float, ticket_epoch: int, deadline_ms: Optional[float],
              trained: torch.Tensor, ring, src_of_step) -> None:
    """One launch = ``steps`` panel passes with the rotation done in-kernel (``hb_sgd_epoch``)."""
    lib = _native.lib()
    _declare(lib)
    deadline = int(deadline_ms * 1e6) if deadline_ms is not None else 0
    steps = len(panel_of_step)
    pos = (ctypes.c_int * steps)(*panel_of_step)
    rc = lib.hb_sgd_epoch(_native.ptr(W), desc, len(desc) // 6, pos, steps, S, workers, cs,
                          W.shape[1], float(eps), float(lam), ticket_epoch & 0xFFFFFFFF, deadline,
                          _native.ptr(trained), *ring.ring_args(src_of_step), _native.stream_ptr())
    _native.check(rc, "hb_sgd_epoch")
    _native.count_launch()


def auto_blocks(nnz_per_panel: float, limit: int, per_block: int = 500) -> int:
    """Blocks per side for a panel: a pass costs ~B hand-overs plus nnz/(B * warps) updates, so small panels
    want fewer, fatter blocks; capped by what one cooperative launch keeps resident."""
    import math
    return int(max(8, min(limit, math.isqrt(max(1, int(nnz_per_panel / per_block))))))


def _serpentine_blocks(weight: torch.Tensor, B: int, greedy_rounds: int = 64) -> torch.Tensor:
    """Block id per item, block totals nearly equal.  The ``greedy_rounds * B`` heaviest items are placed one by one
    on the currently lightest block (LPT, a small host-side heap); the long tail of light items is dealt in
    alternating direction (0..B-1, B-1..0, ...) over the blocks ordered by the load reached so far."""
    import heapq

    n = weight.numel()
    dev = weight.device
    order = torch.argsort(weight, descending=True, stable=True)
    blk_sorted = torch.empty(n, dtype=torch.int64, device=dev)
    head = min(n, greedy_rounds * B)
    w_head = weight[order[:head]].tolist()
    heap = [(0, b) for b in range(B)]
    assign = [0] * head
    for j, w in enumerate(w_head):
        load, b = heapq.heappop(heap)
        assign[j] = b
        heapq.heappush(heap, (load + w, b))
    blk_sorted[:head] = torch.tensor(assign, dtype=torch.int64, device=dev)
    if head < n:
        base = torch.tensor([b for _, b in sorted(heap)], dtype=torch.int64, device=dev)    # lightest block first
        pos = torch.arange(n - head, device=dev)
        off, rnd = pos % B, pos // B
        blk_sorted[head:] = base[torch.where(rnd % 2 == 0, off, B - 1 - off)]
    blk = torch.empty(n, dtype=torch.int64, device=dev)
    blk[order] = blk_sorted
    return blk


def build_panel(rows: torch.Tensor, cols: torch.Tensor, vals: torch.Tensor, B: int, max_run: int = 32,
                hot_spread: bool = True, balance: bool = True, C: Optional[int] = None, lpt: bool = True,
                payload: Optional[torch.Tensor] = None) -> SgdPanel:
    """Buckets a COO rating list (worker-local row ids, slice-local column ids).  Rows and columns are dealt
    to the B row-blocks / column-blocks by weight (``balance=False``: row % B, col % B); inside a block
    ratings are ordered by column (then by original order).
    Column runs longer than ``max_run`` are cut into sub-runs (flagged in bit 31 of ``seg_col``) so that a
    popular column does not serialise a whole block on one warp."""
    dev = rows.device
    rows = rows.to(torch.int64)
    cols = cols.to(torch.int64)
    n = rows.numel()
    C = B if not C else int(C)
    assert C >= B
    if n == 0:
        z = torch.zeros(B * C + 1, dtype=torch.int32, device=dev)
        e = torch.zeros(0, dtype=torch.int32, device=dev)
        return SgdPanel(B, z, e, torch.zeros(1, dtype=torch.int32, device=dev), e,
                        torch.zeros(0, dtype=torch.float32, device=dev), torch.zeros(C, dtype=torch.int32, device=dev), 0, C)
    ncols = int(cols.max().item()) + 1
    # Hot columns (a blockbuster item, a stop word): if all their ratings sat in ONE column-block, that
    # heavy block would be handed round the ring through all B CTAs one after the other.  Their ratings
    # are dealt round-robin over all column-blocks instead; every run of such a column is flagged shared
    # (bit 31), i.e. its updates are merged with atomics, since several CTAs now work on it at once.
    cnt = torch.bincount(cols, minlength=ncols)
    hot_thr = max(4 * max_run if max_run else 256, n // (4 * C) if hot_spread else n + 1)
    hot_col = cnt > hot_thr if hot_spread else torch.zeros(ncols, dtype=torch.bool, device=dev)
    # The ticket ring couples all CTAs: the pass runs at the pace of the heaviest row-block / column-block.
    # Rows and columns are therefore dealt to blocks by weight (heaviest first, serpentine order), not by id.
    if balance and B > 1:
        row_blk = _serpentine_blocks(torch.bincount(rows), B)
        col_blk = _serpentine_blocks(torch.where(hot_col, torch.zeros_like(cnt), cnt), C)
        rb_of, cb = row_blk[rows], col_blk[cols]
    else:
        rb_of, cb = rows % B, cols % C
    if bool(hot_col.any()):
        order_c = torch.argsort(cols, stable=True)
        first = torch.zeros(ncols + 1, dtype=torch.int64, device=dev)
        first[1:] = torch.cumsum(cnt, 0)
        rank_in_col = torch.empty(n, dtype=torch.int64, device=dev)
        rank_in_col[order_c] = torch.arange(n, device=dev) - first[cols[order_c]]
        is_hot = hot_col[cols]
        cb = torch.where(is_hot, (cols + rank_in_col) % C, cb)
    blk = rb_of * C + cb
    key = blk * ncols + cols
    order = torch.argsort(key, stable=True)
    key_s = key[order]
    uniq, counts = torch.unique_consecutive(key_s, return_counts=True)
    if max_run and int(counts.max().item()) > max_run:
        pieces = (counts + max_run - 1) // max_run                       # sub-runs per run
        rep = torch.repeat_interleave(torch.arange(uniq.numel(), device=dev), pieces)
        first = torch.zeros(uniq.numel(), dtype=torch.int64, device=dev)
        first[1:] = torch.cumsum(pieces, 0)[:-1]
        j = torch.arange(rep.numel(), device=dev) - first[rep]           # index of the piece inside its run
        base, extra = counts[rep] // pieces[rep], counts[rep] % pieces[rep]
        sub_counts = base + (j < extra).to(torch.int64)                  # near-equal pieces
        split_flag = (pieces[rep] > 1)
        uniq, counts = uniq[rep], sub_counts
    else:
        split_flag = torch.zeros(uniq.numel(), dtype=torch.bool, device=dev)
    split_flag = split_flag | hot_col[uniq % ncols]
    seg_ptr = torch.zeros(uniq.numel() + 1, dtype=torch.int64, device=dev)
    seg_ptr[1:] = torch.cumsum(counts, 0)
    seg_blk = uniq // ncols
    if lpt and uniq.numel() > 1:
        # Warps of a CTA claim a block's runs in stored order and the block is over when the last warp finishes:
        # longest runs first (LPT) leaves only short runs for the end, so the idle tail shrinks from ~half the
        # longest run to ~half the shortest one.  Ratings keep their order inside a run.
        perm = torch.argsort(seg_blk * (int(counts.max().item()) + 1) - counts, stable=True)
        new_ptr = torch.zeros_like(seg_ptr)
        new_ptr[1:] = torch.cumsum(counts[perm], 0)
        gather = torch.repeat_interleave(seg_ptr[:-1][perm] - new_ptr[:-1], counts[perm]) + torch.arange(n, device=dev)
        order = order[gather]
        uniq, counts, split_flag, seg_blk, seg_ptr = uniq[perm], counts[perm], split_flag[perm], seg_blk[perm], new_ptr
    seg_col = (uniq % ncols).to(torch.int32)
    seg_col = torch.where(split_flag, seg_col | torch.tensor(-2147483648, dtype=torch.int32, device=dev), seg_col)
    blk_seg_ptr = torch.searchsorted(seg_blk, torch.arange(B * C + 1, device=dev, dtype=torch.int64)).to(torch.int32)
    return SgdPanel(B, blk_seg_ptr.contiguous(), seg_col.contiguous(), seg_ptr.to(torch.int32).contiguous(),
                    rows[order].to(torch.int32).contiguous(), vals[order].to(torch.float32).contiguous(),
                    torch.zeros(C, dtype=torch.int32, device=dev), 0, C,
                    payload[order].contiguous() if payload is not None else None)


def row_collision_bound(panel: SgdPanel) -> float:
    """Upper bound on the fraction of a panel's ratings whose W row another warp of the same worker could be
    updating at the same moment: ratings whose row occurs in more than one column run of their cell.  (Different
    cells of a row-block never run concurrently, and different row-blocks never share a row: that is the
    reference's Scheduler exclusivity, ``dymoro/Scheduler.java:95-237``; only runs INSIDE one cell overlap.)"""
    n = panel.nnz
    if n == 0:
        return 0.0
    seg_len = (panel.seg_ptr[1:] - panel.seg_ptr[:-1]).long()
    nseg = seg_len.numel()
    seg_of = torch.repeat_interleave(torch.arange(nseg, device=panel.r_row.device), seg_len)
    cell_of_seg = torch.searchsorted(panel.blk_seg_ptr.long(), torch.arange(nseg, device=seg_of.device), right=True) - 1
    cell = cell_of_seg[seg_of]
    nrow = int(panel.r_row.max().item()) + 1
    key = cell * nrow + panel.r_row.long()
    # a (cell, row) pair is exposed when it shows up in two different runs
    pair = torch.unique(torch.stack([key, seg_of], 1), dim=0)
    uk, inv = torch.unique(key, return_inverse=True)
    exposed_keys = torch.zeros(uk.numel(), dtype=torch.bool, device=key.device)
    k2, c2 = torch.unique(pair[:, 0], return_counts=True)
    exposed_keys[torch.searchsorted(uk, k2)] = c2 > 1
    return float(exposed_keys[inv].float().mean().item())


def _cpu_sgd_pass(panel: SgdPanel, W, H, eps, lam) -> int:
    # sequential restatement in the kernel's own order: Latin-square steps, blocks, column runs
    B, C = panel.B, panel.ncolblocks
    bsp, sc, sp = panel.blk_seg_ptr.tolist(), (panel.seg_col & 0x7FFFFFFF).tolist(), panel.seg_ptr.tolist()
    rr, rv = panel.r_row.tolist(), panel.r_val.tolist()
    n = 0
    for t in range(C):
        for i in range(B):
            c = (i * C // B + t) % C if C != B else (i + t) % B
            for s in range(bsp[i * C + c], bsp[i * C + c + 1]):
                h = H[sc[s]]
                for k in range(sp[s], sp[s + 1]):
                    w = W[rr[k]]
                    e = float(w @ h) - rv[k]
                    w_new = w - eps * (e * h + lam * w)
                    h.sub_(eps * (e * w + lam * h))
                    W[rr[k]] = w_new
                    n += 1
    return n


def _host_sgd_pass(panel: SgdPanel, W, H, eps, lam) -> int:
    """The same sequential pass in C++ (``csrc/host/mf_host.cpp::hh_sgd_pass``); the Python loop above stays as its
    oracle and as the fallback when the host library is not built or the tensors are not plain fp32/fp64."""
    ok = (_native.have_host_lib() and W.dtype == H.dtype and W.dtype in (torch.float32, torch.float64)
          and W.is_contiguous() and H.is_contiguous() and panel.r_val.dtype == torch.float32)
    if not ok:
        return _cpu_sgd_pass(panel, W, H, eps, lam)
    lib = _native.host()
    P = ctypes.c_void_p
    lib.hh_sgd_pass.restype = ctypes.c_longlong
    lib.hh_sgd_pass.argtypes = [P, P, ctypes.c_int, P, P, P, P, P, ctypes.c_int, ctypes.c_int, ctypes.c_double,
                                ctypes.c_double, ctypes.c_int]
    keep = [t.contiguous() for t in (panel.blk_seg_ptr, panel.seg_col, panel.seg_ptr, panel.r_row, panel.r_val)]
    return int(lib.hh_sgd_pass(_native.ptr(W), _native.ptr(H), W.shape[1], *[_native.ptr(t) for t in keep], panel.B,
                               panel.ncolblocks, float(eps), float(lam), 1 if W.dtype == torch.float64 else 0))


def sgd_pass(panel: SgdPanel, W: torch.Tensor, H: torch.Tensor, eps: float, lam: float,
             deadline_ms: Optional[float] = None, trained: Optional[torch.Tensor] = None) -> torch.Tensor:
    """One pass over the panel; returns a 1-element int64 tensor with the number of ratings trained."""
    if W.device.type != "cuda":
        n = _host_sgd_pass(panel, W, H, eps, lam)
        out = trained if trained is not None else torch.zeros(1, dtype=torch.int64)
        out += n
        return out
    lib = _native.lib()
    _declare(lib)
    assert W.dtype == torch.float32 and H.dtype == torch.float32 and W.is_contiguous() and H.is_contiguous()
    rank = W.shape[1]
    if trained is None:
        trained = torch.zeros(1, dtype=torch.int64, device=W.device)
    deadline = int(deadline_ms * 1e6) if deadline_ms is not None else 0     # budget from kernel entry
    if panel.ncolblocks != panel.B:        # dynamic schedule: any free, not yet visited column-block
        rc = lib.hb_sgd_pass_dynamic(_native.ptr(W), _native.ptr(H), _native.ptr(panel.blk_seg_ptr), _native.ptr(panel.seg_col),
                                     _native.ptr(panel.seg_ptr), _native.ptr(panel.r_row), _native.ptr(panel.r_val), panel.B,
                                     panel.ncolblocks, rank, float(eps), float(lam), _native.ptr(panel.col_ticket), deadline,
                                     _native.ptr(trained), _native.stream_ptr())
        _native.check(rc, "hb_sgd_pass_dynamic")
        _native.count_launch()
        return trained
    rc = lib.hb_sgd_pass(_native.ptr(W), _native.ptr(H), _native.ptr(panel.blk_seg_ptr), _native.ptr(panel.seg_col),
                         _native.ptr(panel.seg_ptr), _native.ptr(panel.r_row), _native.ptr(panel.r_val), panel.B, rank,
                         float(eps), float(lam), _native.ptr(panel.col_ticket), panel.ticket_base & 0xFFFFFFFF,
                         deadline, _native.ptr(trained), _native.stream_ptr())
    _native.check(rc, "hb_sgd_pass")
    panel.ticket_base = (panel.ticket_base + panel.B) & 0x7FFFFFFF
    _native.count_launch()
    return trained


def sq_err(W: torch.Tensor, H: torch.Tensor, rows: torch.Tensor, cols: torch.Tensor, vals: torch.Tensor,
           out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``Σ (v - w·h)²`` (float64, accumulated into ``out``)."""
    if out is None:
        out = torch.zeros(1, dtype=torch.float64, device=W.device)
    if rows.numel() == 0:
        return out
    if W.device.type != "cuda":
        pred = (W[rows.long()] * H[cols.long()]).sum(1)
        out += ((vals - pred).double() ** 2).sum()
        return out
    lib = _native.lib()
    _declare(lib)
    r32 = rows.to(torch.int32).contiguous()      # keep the converted tensors alive across the launch
    c32 = cols.to(torch.int32).contiguous()
    v32 = vals.float().contiguous()
    rc = lib.hb_mf_sq_err(_native.ptr(W), _native.ptr(H), _native.ptr(r32), _native.ptr(c32), _native.ptr(v32),
                          rows.numel(), W.shape[1], _native.ptr(out), _native.stream_ptr())
    _native.check(rc, "hb_mf_sq_err")
    _native.count_launch()
    return out


def ccd_update(ptr: torch.Tensor, idx: torch.Tensor, resid: torch.Tensor, perm: Optional[torch.Tensor],
               Wf: torch.Tensor, Hf: torch.Tensor, lam: float) -> None:
    """CCD++ rank-one update of ``Wf`` (one feature) given ``Hf``; ``resid`` (per rating, addressed through
    ``perm`` when the CSR order differs from the residual order) is updated in place."""
    nrows = Wf.numel()
    if Wf.device.type != "cuda":
        for i in range(nrows):
            k0, k1 = int(ptr[i]), int(ptr[i + 1])
            if k0 == k1:
                continue
            j = idx[k0:k1].long()
            rk = perm[k0:k1].long() if perm is not None else torch.arange(k0, k1)
            hj = Hf[j]
            wi = Wf[i]
            wnew = ((resid[rk] + wi * hj) * hj).sum() / (lam * (k1 - k0) + (hj * hj).sum())     # CCDMPTask.java:80
            resid[rk] += (wi - wnew) * hj
            Wf[i] = wnew
        return
    lib = _native.lib()
    _declare(lib)
    rc = lib.hb_ccd_update(_native.ptr(ptr), _native.ptr(idx), _native.ptr(resid), _native.ptr(perm), _native.ptr(Wf),
                           _native.ptr(Hf), nrows, float(lam), _native.stream_ptr())
    _native.check(rc, "hb_ccd_update")
    _native.count_launch()


def ccd_sweep(ptr: torch.Tensor, idx: torch.Tensor, res: torch.Tensor, other: torch.Tensor, self_: torch.Tensor,
              lam: float, pd_other: Optional[torch.Tensor] = None, pd_self: Optional[torch.Tensor] = None,
              dself: Optional[torch.Tensor] = None, wide: bool = False) -> None:
    """One CCD++ sweep of ONE feature over the owned ids of one side, in a single pass over their ratings
    (``hb_ccd_sweep``; ``CCDMPTask.java:67-125``).  ``res`` is this side's residual order, stale by the rank-one term
    ``pd_other[idx] * pd_self[i]`` when the previous sweep ran on the other order; it is brought up to date and then
    updated for the new value, ``self_`` is overwritten with the new values and ``dself`` receives their change."""
    n = self_.numel()
    if self_.device.type != "cuda":
        for i in range(n):
            k0, k1 = int(ptr[i]), int(ptr[i + 1])
            if k0 == k1:
                if dself is not None:
                    dself[i] = 0
                continue
            j = idx[k0:k1].long()
            o = other[j]
            r = res[k0:k1]
            if pd_other is not None:
                r = r - pd_other[j] * pd_self[i]
            s_old = self_[i].clone()
            s_new = ((r + s_old * o) * o).sum() / (lam * (k1 - k0) + (o * o).sum())
            res[k0:k1] = r - (s_new - s_old) * o
            self_[i] = s_new
            if dself is not None:
                dself[i] = s_new - s_old
        return
    lib = _native.lib()
    _declare(lib)
    rc = lib.hb_ccd_sweep(_native.ptr(ptr), _native.ptr(idx), _native.ptr(res), _native.ptr(pd_other), _native.ptr(pd_self),
                          _native.ptr(other), _native.ptr(self_), _native.ptr(dself), n, float(lam), 1 if wide else 0,
                          _native.stream_ptr())
    _native.check(rc, "hb_ccd_sweep")
    _native.count_launch()
