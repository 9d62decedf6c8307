"""Colour-coding DP step of subgraph counting (``csrc/subgraph.cu``) with a torch restatement for CPU tensors.

Tables are compressed the way the reference's are (``colorcount_HJ.java``: one entry per colour set of the
sub-template's size): a sub-template of size s over k colours has ``C(k, s)`` columns, ordered by ascending bit mask.
"""
from __future__ import annotations

import ctypes
import itertools
from functools import lru_cache
from typing import Dict, List, Tuple

import torch

from .. import _native


@lru_cache(maxsize=None)
def colour_sets(k: int, s: int) -> Tuple[int, ...]:
    """Bit masks of the ``C(k, s)`` colour sets of size s, ascending (= the column order of a size-s table)."""
    return tuple(sorted(sum(1 << c for c in comb) for comb in itertools.combinations(range(k), s)))


@lru_cache(maxsize=None)
def _split_lists(k: int, s: int, sa: int):
    ia: Dict[int, int] = {m: i for i, m in enumerate(colour_sets(k, sa))}
    ip: Dict[int, int] = {m: i for i, m in enumerate(colour_sets(k, s - sa))}
    A: List[List[int]] = []
    P: List[List[int]] = []
    for C in colour_sets(k, s):
        bits = [c for c in range(k) if C >> c & 1]
        a_row, p_row = [], []
        for comb in itertools.combinations(bits, sa):
            m = sum(1 << c for c in comb)
            a_row.append(ia[m]); p_row.append(ip[C ^ m])
        A.append(a_row); P.append(p_row)
    return A, P


_SPLIT_CACHE: Dict[tuple, Tuple[torch.Tensor, torch.Tensor]] = {}


def split_tables(k: int, s: int, sa: int, device: torch.device) -> Tuple[torch.Tensor, torch.Tensor]:
    """``(split_a, split_p)`` int16 ``[C(k, s), C(s, sa)]``: for colour set c and split j, the column of the active
    table (size sa) and of the passive table (size s - sa) whose colour sets partition c."""
    key = (k, s, sa, str(device))
    if key not in _SPLIT_CACHE:
        A, P = _split_lists(k, s, sa)
        _SPLIT_CACHE[key] = (torch.tensor(A, dtype=torch.int16, device=device), torch.tensor(P, dtype=torch.int16, device=device))
    return _SPLIT_CACHE[key]


def combine(nbr_ptr: torch.Tensor, nbr_idx: torch.Tensor, act: torch.Tensor, pas: torch.Tensor, split_a: torch.Tensor,
            split_p: torch.Tensor) -> torch.Tensor:
    """``out[v, c] = Σ_j act[v, split_a[c, j]] · Σ_{u ∈ N(v)} pas[u, split_p[c, j]]``; ``nbr_idx`` holds, for every
    neighbour, its ROW in ``pas`` (local rows first, then whatever remote rows the caller appended)."""
    n, na = act.shape
    np_ = pas.shape[1]
    nC, nsplit = split_a.shape
    if act.device.type != "cuda" or act.dtype != torch.float32:
        deg = (nbr_ptr[1:] - nbr_ptr[:-1]).long()
        src = torch.repeat_interleave(torch.arange(n, device=act.device), deg)
        pasN = torch.zeros(n, np_, dtype=act.dtype, device=act.device).index_add_(0, src, pas[nbr_idx.long()])
        out = torch.empty(n, nC, dtype=act.dtype, device=act.device)
        step = max(1, (1 << 24) // max(nC * nsplit, 1))
        sa_l, sp_l = split_a.long(), split_p.long()
        for r0 in range(0, n, step):
            out[r0:r0 + step] = (act[r0:r0 + step][:, sa_l] * pasN[r0:r0 + step][:, sp_l]).sum(-1)
        return out
    lib = _native.lib()
    if not getattr(lib, "_hb_subgraph_declared", False):
        P, I = ctypes.c_void_p, ctypes.c_int
        lib.hb_subgraph_combine.argtypes = [P, P, P, P, P, P, I, I, I, I, P, I, P]
        lib._hb_subgraph_declared = True
    out = torch.empty(n, nC, dtype=torch.float32, device=act.device)
    _native.check(lib.hb_subgraph_combine(_native.ptr(nbr_ptr.to(torch.int64).contiguous()), _native.ptr(nbr_idx.to(torch.int32).contiguous()),
                                          _native.ptr(act.contiguous()), _native.ptr(pas.contiguous()),
                                          _native.ptr(split_a.contiguous()), _native.ptr(split_p.contiguous()),
                                          na, np_, nC, nsplit, _native.ptr(out), n, _native.stream_ptr()),
                  "hb_subgraph_combine")
    _native.count_launch()
    return out
