"""Sparse graph ops of the contrib applications (``csrc/graph.cu``) with torch restatements for CPU tensors."""
from __future__ import annotations

import ctypes

import torch

from .. import _native


def pagerank_push(ptr: torch.Tensor, idx: torch.Tensor, src_ids: torch.Tensor, pr: torch.Tensor, out: torch.Tensor,
                  dangling: torch.Tensor) -> None:
    """``out[t] += pr[s] / outdeg(s)`` for every link ``s → t`` of the local source pages; ``dangling[0] +=`` the rank
    of local pages without out-links (``contrib/.../pagerank/PageRankMapper.java``)."""
    m = src_ids.numel()
    if m == 0:
        return
    if pr.device.type != "cuda":
        deg = (ptr[1:] - ptr[:-1])
        p = pr[src_ids]
        has = deg > 0
        share = torch.where(has, p / deg.clamp(min=1).to(p.dtype), torch.zeros_like(p))
        out.index_add_(0, idx.long(), torch.repeat_interleave(share, deg))
        dangling += p[~has].sum()
        return
    lib = _native.lib()
    if not getattr(lib, "_hb_graph_declared", False):
        P, I = ctypes.c_void_p, ctypes.c_int
        lib.hb_pagerank_push.argtypes = [P, P, P, P, P, P, I, P]
        lib._hb_graph_declared = True
    assert pr.dtype == torch.float64 and out.dtype == torch.float64 and out.is_contiguous() and dangling.is_contiguous()
    _native.check(lib.hb_pagerank_push(_native.ptr(ptr.to(torch.int64).contiguous()), _native.ptr(idx.to(torch.int32).contiguous()),
                                       _native.ptr(src_ids.to(torch.int64).contiguous()), _native.ptr(pr.contiguous()),
                                       _native.ptr(out), _native.ptr(dangling), m, _native.stream_ptr()), "hb_pagerank_push")
    _native.count_launch()
