"""LDA collapsed-Gibbs ops (``csrc/lda.cu``) with a CPU restatement of the same sampler
(``ml/java/src/main/java/edu/iu/lda/LDAMPTask.java:85-331`` — same conditional distribution, dense
instead of SparseLDA's bucket decomposition)."""
from __future__ import annotations

import ctypes
from dataclasses import dataclass
from typing import Optional

import torch

from .. import _native
from . import mf


def _declare(lib) -> None:
    if getattr(lib, "_hb_lda_declared", False):
        return
    P, I, LL, F, U, ULL = (ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong, ctypes.c_float, ctypes.c_uint32,
                           ctypes.c_ulonglong)
    lib.hb_lda_max_blocks.argtypes = [I]
    lib.hb_lda_pass.argtypes = [P, P, P, P, P, P, P, P, I, I, F, F, ULL, P, U, ULL, P, P]
    lib.hb_lda_pass_dk16.argtypes = [P, P, P, P, P, P, P, P, I, I, F, F, ULL, P, U, ULL, P, P]
    lib.hb_lda_topic_sums.argtypes = [P, LL, I, P, P]
    lib.hb_lda_count.argtypes = [P, P, LL, I, P, P]
    lib.hb_lda_epoch_supported.argtypes = [I, I]
    lib.hb_lda_epoch_max_workers.argtypes = [I, I, I]
    lib.hb_lda_epoch.argtypes = [P, I, P, P, I, P, I, I, I, I, I, F, F, ULL, U, ULL, P,
                                 P, I, I, ULL, P, P, P, U, P, ULL, P]
    lib.hb_lda_tl_supported.argtypes = [I]
    lib.hb_lda_tl_max_workers.argtypes = [I, I]
    lib.hb_lda_tl_epoch.argtypes = [P, P, P, P, I, P, I, I, I, I, I, F, F, ULL, U, ULL, P,
                                    P, I, I, ULL, P, P, P, U, P, ULL, P]
    lib.hb_lda_tl_epoch2.argtypes = [P, P, P, P, I, P, I, I, I, I, I, F, F, ULL, U, ULL, P,
                                     P, I, I, ULL, P, P, P, U, P, ULL, ULL, P]
    lib.hb_lgamma_sum.argtypes = [P, LL, ctypes.c_double, P, P, P]
    lib._hb_lda_declared = True


@dataclass
class TokenPanel:
    """Tokens of (this worker's documents) x (one word slice): B x B blocks, word-sorted inside a block."""
    B: int
    blk_seg_ptr: torch.Tensor
    seg_word: torch.Tensor
    seg_ptr: torch.Tensor
    tok_doc: torch.Tensor     # int32 [ntok]
    tok_z: torch.Tensor       # int32 [ntok] (updated in place by the sampler)
    tok_word: torch.Tensor    # int32 [ntok] slice-local word id (for table initialisation / checks)
    col_ticket: torch.Tensor
    ticket_base: int = 0
    tok_pos: Optional[torch.Tensor] = None   # token -> index into the rank's document-ordered topic list (token-list sampler)

    @property
    def ntok(self) -> int:
        return int(self.tok_doc.numel())


def max_blocks(K: int) -> int:
    lib = _native.lib()
    _declare(lib)
    return int(lib.hb_lda_max_blocks(K))


def epoch_supported(K: int, dk16: bool) -> bool:
    lib = _native.lib()
    _declare(lib)
    return bool(lib.hb_lda_epoch_supported(int(K), 1 if dk16 else 0))


def epoch_max_workers(K: int, dk16: bool, cs: int) -> int:
    lib = _native.lib()
    _declare(lib)
    return int(lib.hb_lda_epoch_max_workers(int(K), 1 if dk16 else 0, int(cs)))


def panel_descriptors(panels):
    """Host array of the 6-pointer descriptors of ``lda_epoch_kernel`` (``LdaPanelDesc``)."""
    flat = [t.data_ptr() for p in panels for t in (p.blk_seg_ptr, p.seg_word, p.seg_ptr, p.tok_doc, p.tok_z, p.col_ticket)]
    return (ctypes.c_uint64 * len(flat))(*flat)


def lda_epoch(n_dk: torch.Tensor, commons: torch.Tensor, desc, panel_of_step,
              S: int, workers: int, cs: int, alpha: float, beta: float, seed: int, ticket_epoch: int,
              deadline_ms: Optional[float], trained: torch.Tensor, ring, src_of_step) -> None:
    """One launch = a whole iteration's ``steps`` panel passes with the rotation done in-kernel (``hb_lda_epoch``)."""
    lib = _native.lib()
    _declare(lib)
    deadline = int(deadline_ms * 1e6) if deadline_ms is not None else 0
    steps = len(panel_of_step)
    pos = (ctypes.c_int * steps)(*panel_of_step)
    rc = lib.hb_lda_epoch(_native.ptr(n_dk), 1 if n_dk.dtype == torch.int16 else 0, _native.ptr(commons),
                          desc, len(desc) // 6, pos, steps, S, workers, cs, n_dk.shape[1],
                          float(alpha), float(beta), int(seed) & 0xFFFFFFFFFFFFFFFF, ticket_epoch & 0xFFFFFFFF, deadline,
                          _native.ptr(trained), *ring.ring_args(src_of_step), _native.stream_ptr())
    _native.check(rc, "hb_lda_epoch")
    _native.count_launch()


def tl_max_workers(K: int, cs: int) -> int:
    lib = _native.lib()
    _declare(lib)
    return int(lib.hb_lda_tl_max_workers(int(K), int(cs)))


def tl_panel_descriptors(panels):
    """Host array of the 7-pointer descriptors of the token-list kernel (``csrc/lda_tl.cu::PanelDesc``)."""
    flat = [t.data_ptr() for p in panels for t in (p.blk_seg_ptr, p.seg_word, p.seg_ptr, p.tok_doc, p.tok_z, p.col_ticket,
                                                    p.tok_pos)]
    return (ctypes.c_uint64 * len(flat))(*flat)


def lda_tl_epoch(doc_z: torch.Tensor, d_ptr: torch.Tensor, commons: torch.Tensor, K: int, desc, panel_of_step, S: int,
                 workers: int, cs: int, alpha: float, beta: float, seed: int, ticket_epoch: int,
                 deadline_ms: Optional[float], trained: torch.Tensor, ring, src_of_step, wb_off: int = 0) -> None:
    """One iteration with the SparseLDA-style token-list sampler (``hb_lda_tl_epoch2``), any K <= 16384.  ``wb_off``:
    element offset of the per-32-topic block counts inside every slice (0: slices hold the word rows only)."""
    lib = _native.lib()
    _declare(lib)
    deadline = int(deadline_ms * 1e6) if deadline_ms is not None else 0
    steps = len(panel_of_step)
    pos = (ctypes.c_int * steps)(*panel_of_step)
    rc = lib.hb_lda_tl_epoch2(_native.ptr(doc_z), _native.ptr(d_ptr), _native.ptr(commons), desc, len(desc) // 7, pos, steps,
                              S, workers, cs, int(K), float(alpha), float(beta), int(seed) & 0xFFFFFFFFFFFFFFFF,
                              ticket_epoch & 0xFFFFFFFF, deadline, _native.ptr(trained), *ring.ring_args(src_of_step),
                              int(wb_off), _native.stream_ptr())
    _native.check(rc, "hb_lda_tl_epoch2")
    _native.count_launch()


def lgamma_sum(counts: torch.Tensor, add: float):
    """``(Σ_{c>0} lgamma(c + add), #non-zeros)`` of an int32 count table in one fused pass (K11, ``hb_lgamma_sum``)."""
    if counts.device.type != "cuda" or counts.dtype != torch.int32:
        c = counts[counts > 0].double()
        return float(torch.lgamma(c + add).sum()), int(c.numel())
    lib = _native.lib()
    _declare(lib)
    out = torch.zeros(1, dtype=torch.float64, device=counts.device)
    nnz = torch.zeros(1, dtype=torch.int64, device=counts.device)
    c = counts.contiguous()
    _native.check(lib.hb_lgamma_sum(_native.ptr(c), c.numel(), float(add), _native.ptr(out), _native.ptr(nnz),
                                    _native.stream_ptr()), "hb_lgamma_sum")
    _native.count_launch()
    return float(out.item()), int(nnz.item())


def build_token_panel(docs: torch.Tensor, words: torch.Tensor, z: torch.Tensor, B: int, max_run: int = 64,
                      pos: Optional[torch.Tensor] = None) -> TokenPanel:
    assert int(z.max().item()) < (1 << 24) if z.numel() else True      # topics travel through a float32 column
    sp = mf.build_panel(docs, words, z.to(torch.float32), B, max_run=max_run, payload=pos)
    # recover each token's word from the run structure
    seg_len = (sp.seg_ptr[1:] - sp.seg_ptr[:-1]).long()
    tok_word = torch.repeat_interleave((sp.seg_col & 0x7FFFFFFF).to(torch.int32), seg_len)
    return TokenPanel(B, sp.blk_seg_ptr, sp.seg_col, sp.seg_ptr, sp.r_row, sp.r_val.to(torch.int32).contiguous(),
                      tok_word.contiguous(), sp.col_ticket, 0,
                      sp.extra.to(torch.int32).contiguous() if sp.extra is not None else None)


def count_into(idx0: torch.Tensor, z: torch.Tensor, K: int, counts: torch.Tensor) -> None:
    """``counts[idx0[t], z[t]] += 1``."""
    if idx0.numel() == 0:
        return
    if counts.device.type != "cuda":
        counts.view(-1).index_add_(0, idx0.long() * K + z.long(), torch.ones(idx0.numel(), dtype=counts.dtype))
        return
    lib = _native.lib()
    _declare(lib)
    i32, z32 = idx0.to(torch.int32).contiguous(), z.to(torch.int32).contiguous()
    _native.check(lib.hb_lda_count(_native.ptr(i32), _native.ptr(z32), idx0.numel(), K, _native.ptr(counts),
                                   _native.stream_ptr()), "hb_lda_count")
    _native.count_launch()


def topic_sums(n_wk: torch.Tensor, out: torch.Tensor) -> torch.Tensor:
    """``out[k] += Σ_w n_wk[w, k]`` (int64)."""
    if n_wk.device.type != "cuda":
        out += n_wk.long().sum(0)
        return out
    lib = _native.lib()
    _declare(lib)
    _native.check(lib.hb_lda_topic_sums(_native.ptr(n_wk), n_wk.shape[0], n_wk.shape[1], _native.ptr(out),
                                        _native.stream_ptr()), "hb_lda_topic_sums")
    _native.count_launch()
    return out


def _cpu_pass(panel: TokenPanel, n_dk, n_wk, commons, alpha, beta, seed) -> int:
    g = torch.Generator().manual_seed(int(seed) & 0x7FFFFFFF)
    B = panel.B
    bsp, sw, sp = panel.blk_seg_ptr.tolist(), (panel.seg_word & 0x7FFFFFFF).tolist(), panel.seg_ptr.tolist()
    td = panel.tok_doc.tolist()
    n = 0
    cm = commons.double()
    for t in range(B):
        for i in range(B):
            c = (i + t) % B
            for s in range(bsp[i * B + c], bsp[i * B + c + 1]):
                w = sw[s]
                for tk in range(sp[s], sp[s + 1]):
                    d, zo = td[tk], int(panel.tok_z[tk])
                    n_dk[d, zo] -= 1; n_wk[w, zo] -= 1
                    p = (n_dk[d].double() + alpha) * (n_wk[w].double() + beta) * cm
                    zn = int(torch.multinomial(p / p.sum(), 1, generator=g))
                    n_dk[d, zn] += 1; n_wk[w, zn] += 1
                    panel.tok_z[tk] = zn
                    n += 1
    return n


def _host_pass(panel: TokenPanel, n_dk, n_wk, commons, alpha, beta, seed) -> int:
    """The sequential sampler in C++ (``csrc/host/mf_host.cpp::hh_lda_pass``, counter-based RNG like the device
    kernel); falls back to the Python loop when the host library is missing or the tables are not int32/int16."""
    ok = (_native.have_host_lib() and n_wk.dtype == torch.int32 and n_dk.dtype in (torch.int32, torch.int16)
          and n_dk.is_contiguous() and n_wk.is_contiguous() and panel.tok_z.dtype == torch.int32
          and commons.dtype == torch.float32)
    if not ok:
        return _cpu_pass(panel, n_dk, n_wk, commons, alpha, beta, seed)
    lib = _native.host()
    P = ctypes.c_void_p
    lib.hh_lda_pass.restype = ctypes.c_longlong
    lib.hh_lda_pass.argtypes = [P, ctypes.c_int, P, P, P, P, P, P, P, ctypes.c_int, ctypes.c_int, ctypes.c_double,
                                ctypes.c_double, ctypes.c_ulonglong]
    keep = [t.contiguous() for t in (commons, panel.blk_seg_ptr, panel.seg_word, panel.seg_ptr, panel.tok_doc)]
    return int(lib.hh_lda_pass(_native.ptr(n_dk), 1 if n_dk.dtype == torch.int16 else 0, _native.ptr(n_wk),
                               *[_native.ptr(t) for t in keep], _native.ptr(panel.tok_z), panel.B, n_dk.shape[1],
                               float(alpha), float(beta), int(seed) & 0xFFFFFFFFFFFFFFFF))


def lda_pass(panel: TokenPanel, n_dk: torch.Tensor, n_wk: torch.Tensor, commons: torch.Tensor, alpha: float,
             beta: float, seed: int, deadline_ms: Optional[float] = None,
             trained: Optional[torch.Tensor] = None) -> torch.Tensor:
    if n_dk.device.type != "cuda":
        n = _host_pass(panel, n_dk, n_wk, commons, alpha, beta, seed)
        out = trained if trained is not None else torch.zeros(1, dtype=torch.int64)
        out += n
        return out
    lib = _native.lib()
    _declare(lib)
    # n_dk may be int16 (documents shorter than 32768 tokens): half the row traffic, the kernel's bound
    assert n_dk.dtype in (torch.int32, torch.int16) and n_wk.dtype == torch.int32 and commons.dtype == torch.float32
    assert n_dk.is_contiguous() and n_wk.is_contiguous()
    K = n_dk.shape[1]
    if trained is None:
        trained = torch.zeros(1, dtype=torch.int64, device=n_dk.device)
    deadline = int(deadline_ms * 1e6) if deadline_ms is not None else 0
    fn = lib.hb_lda_pass_dk16 if n_dk.dtype == torch.int16 else lib.hb_lda_pass
    rc = fn(_native.ptr(n_dk), _native.ptr(n_wk), _native.ptr(commons), _native.ptr(panel.blk_seg_ptr),
            _native.ptr(panel.seg_word), _native.ptr(panel.seg_ptr), _native.ptr(panel.tok_doc),
            _native.ptr(panel.tok_z), panel.B, K, float(alpha), float(beta), int(seed) & 0xFFFFFFFFFFFFFFFF,
            _native.ptr(panel.col_ticket), panel.ticket_base & 0xFFFFFFFF, deadline, _native.ptr(trained),
            _native.stream_ptr())
    _native.check(rc, "hb_lda_pass")
    panel.ticket_base = (panel.ticket_base + panel.B) & 0x7FFFFFFF
    _native.count_launch()
    return trained
