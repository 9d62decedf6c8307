"""LDA-CGS engine with model rotation (``ml/java/src/main/java/edu/iu/lda/LDAMPCollectiveMapper.java:183-378``).

Documents (and their dense doc-topic rows) stay on their worker; the word-topic table is sharded by
word (home worker ``w % N``, slice ``(w // N) % S``, slice-local row ``w // SN``) and its S slices
travel the ring, double-buffered in symmetric memory and pulled over NVLink on a side stream while
the other slice is being sampled (``S = num_slices``: 2 like the reference for large panels, 1 for small
ones, where one fatter panel per rotation step keeps the block schedule efficient).  Global topic sums are refreshed once per iteration
(``:442-475``); log-likelihood follows ``CalcLikelihoodTask`` (``:699-752``).
"""
from __future__ import annotations

import math
import os
from typing import List, Optional

import torch
import torch.distributed as dist

from ..ops import lda as L
from ..ops import mf as _mf
from ..parallel.ring import RingSchedule, RingState


class LDAEngine:
    def __init__(self, docs: torch.Tensor, words: torch.Tensor, num_docs_local: int, vocab: int, K: int,
                 alpha: float, beta: float, *, symm=None, blocks: Optional[int] = None, seed: int = 0, group=None,
                 num_slices: Optional[int] = None, dk_dtype: str = "auto", fused: Optional[bool] = None,
                 cluster: Optional[int] = None, loopback: bool = False, max_run: Optional[int] = None,
                 virtual_world: int = 0, sampler: Optional[str] = None):
        self.dev = docs.device
        self.K, self.V = int(K), int(vocab)
        self.alpha, self.beta = float(alpha), float(beta)
        self.group = group
        self.N = dist.get_world_size(group) if dist.is_initialized() else 1
        self.me = dist.get_rank(group) if dist.is_initialized() else 0
        self.symm = symm if self.N > 1 else None
        N = self.N
        import os
        # tuning aid, one GPU only: the per-rank layout of a `virtual_world`-rank job (see MFSGDEngine)
        self.Nv = int(virtual_world) if (virtual_world and N == 1) else 0
        if self.Nv:
            N = self.Nv
            loopback = True
        cuda = self.dev.type == "cuda"
        self.S = S = int(num_slices) if num_slices else 2          # two slices in flight, like the reference
        assert S in (1, 2)
        self.cap = max(1, (self.V + S * N - 1) // (S * N))
        ntok_max = float(docs.numel())
        if dist.is_initialized() and N > 1:
            t = torch.tensor([ntok_max], dtype=torch.float64, device=self.dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
            ntok_max = float(t.item())
        per_panel = ntok_max / (float(S) * N)
        # The sampling kernel is bound by reading one document-topic row per token; when no document is longer
        # than an int16 can count, the table is kept in 16 bits (dk_dtype="int32" forces the wide table).
        dk_dtype = os.environ.get("HB_LDA_DK", dk_dtype)          # triage knob
        if dk_dtype == "auto":
            longest = int(torch.bincount(docs.long(), minlength=1).max().item()) if docs.numel() else 0
            dk_dtype = "int16" if (cuda and longest < 32768) else "int32"
        self.dk16 = dk_dtype == "int16"
        # Sampler: "dense" forms p_k for all K topics per token (csrc/lda.cu, K <= 2048); "tokenlist" is the
        # SparseLDA-style sampler of csrc/lda_tl.cu (document bucket = sum over the document's token list, word and
        # smoothing buckets through block sums; any K <= 16384, no document-topic table).  Default: tokenlist on GPUs.
        sampler = sampler or os.environ.get("HB_LDA_SAMPLER")
        can_fuse = cuda and (self.N == 1 or self.symm is not None)
        if sampler is None:
            # measured on BASELINE config 4 (K = 1000, 100-token documents): dense 1.42 G tokens/s, token list 0.61 G on
            # one B200 — the dense kernel prefetches the next document row while it samples, the token-list kernel
            # pays two dependent gathers per token; so dense wherever it exists, token list beyond
            dense_ok = L.epoch_supported(self.K, self.dk16) if cuda else True
            sampler = "dense" if (dense_ok or not can_fuse or fused is False) else "tokenlist"
        self.sampler = sampler
        self.tl = sampler == "tokenlist"
        if self.tl:
            assert can_fuse, "the token-list sampler runs in the whole-iteration kernel (CUDA; symmetric workspace when N > 1)"
            fused = True
        if fused is None:
            fused = (can_fuse and os.environ.get("HB_LDA_FUSED", "1") != "0" and L.epoch_supported(self.K, self.dk16))
        self.fused = bool(fused)
        self.cs = 1
        if cuda and self.fused:
            self.cs = int(cluster) if cluster else _mf.pick_cluster(per_panel, "lda")
            maxw = (lambda c: L.tl_max_workers(self.K, c)) if self.tl else (lambda c: L.epoch_max_workers(self.K, self.dk16, c))
            mx = maxw(self.cs)
            while mx < 1 and self.cs > 1:
                self.cs //= 2
                mx = maxw(self.cs)
            self.B = max(1, min(int(blocks) if blocks else mx, mx))
            if max_run is None:
                x = per_panel / (self.B * self.B) / (2.0 * self.cs * 24)        # ~2 runs per sampling warp of a worker
                max_run = 16 if x < 24 else (32 if x < 48 else 64)
        elif cuda:
            sms = torch.cuda.get_device_properties(self.dev).multi_processor_count
            self.B = int(blocks) if blocks else _mf.auto_blocks(per_panel, min(L.max_blocks(self.K), sms), per_block=32)
        else:
            self.B = int(blocks) if blocks else 2
        self.max_run = int(max_run) if max_run else 64
        self.seed = int(seed)
        self.iteration = 0
        g = torch.Generator(device=self.dev).manual_seed(seed + 31 * self.me)
        self.ntok = int(docs.numel())
        z = torch.randint(0, self.K, (self.ntok,), device=self.dev, generator=g)
        words = words.to(torch.int64)
        pos = None
        if self.tl:
            # the rank's tokens in document order: doc_z[d_ptr[d] .. d_ptr[d+1]) are the current topics of document d
            order = torch.argsort(docs.to(torch.int64), stable=True)
            pos = torch.empty(self.ntok, dtype=torch.int64, device=self.dev)
            pos[order] = torch.arange(self.ntok, device=self.dev)
            self.doc_z = z[order].to(torch.int32).contiguous()
            self.d_ptr = torch.zeros(num_docs_local + 1, dtype=torch.int32, device=self.dev)
            self.d_ptr[1:] = torch.cumsum(torch.bincount(docs.to(torch.int64), minlength=num_docs_local), 0).to(torch.int32)
            self.doc_of_pos = docs.to(torch.int64)[order]
        shard, slc, local = words % N, (words // N) % S, words // (S * N)
        self.panels: List[List[L.TokenPanel]] = []
        for s in range(N):
            per = []
            for k in range(self.S):
                m = (shard == s) & (slc == k)
                per.append(L.build_token_panel(docs[m], local[m], z[m], self.B, max_run=self.max_run,
                                               pos=pos[m] if pos is not None else None))
            self.panels.append(per)
        cdt = torch.int32
        self.num_docs_local = int(num_docs_local)
        self._n_dk = None
        if not self.tl:
            self._n_dk = torch.zeros(num_docs_local, self.K, dtype=cdt, device=self.dev)
            for s in range(N):
                for k in range(self.S):
                    p = self.panels[s][k]
                    L.count_into(p.tok_doc, p.tok_z, self.K, self._n_dk)
            if self.dk16:
                self._n_dk = self._n_dk.to(torch.int16)
        self.ring = None
        self._cur = [0] * S
        self.sched = [RingSchedule(self.N, None) for _ in range(S)]     # plain ring (LDAMPCollectiveMapper.java:255-259)
        if cuda:
            # token-list sampler: every slice carries, behind its word rows, the exact per-32-topic block counts of
            # those rows (they rotate together); a run then reads only the non-empty blocks of its word's row
            self.nb = (self.K + 31) // 32 if (self.tl and os.environ.get("HB_LDA_TL_BLOCKS", "1") != "0") else 0
            self.ring = RingState(self.dev, self.cap * (self.K + self.nb), cdt, S, symm=self.symm, rank=self.me,
                                  world=self.N, loopback=loopback)
            self.Wbuf = [[b_[:self.cap * self.K].view(self.cap, self.K) for b_ in pair] for pair in self.ring.buf]
            self.Bbuf = [[b_[self.cap * self.K:].view(self.cap, self.nb) for b_ in pair] for pair in self.ring.buf] if self.nb else None
        else:
            self.Wbuf = [[torch.zeros(self.cap, self.K, dtype=cdt, device=self.dev) for _ in range(2 if N > 1 else 1)]
                         for _ in range(S)]
        if self.dev.type == "cuda":
            self.comm_stream = torch.cuda.Stream(device=self.dev)
            self.ev_compute = [torch.cuda.Event(), torch.cuda.Event()]
            self.ev_comm = [torch.cuda.Event(), torch.cuda.Event()]
            for e in self.ev_comm:
                e.record()
        # word-topic table: one full rotation in which every worker adds its tokens to the visiting shard
        for j in range(N):
            for k in range(self.S):
                self._wait_slice(k)
                p = self.panels[j if self.Nv else self._visiting(k)][k]
                L.count_into(p.tok_word, p.tok_z, self.K, self.Wbuf[k][self.cur[k]])
                if not self.Nv:
                    self._rotate(k)
        if cuda and self.nb:
            for k in range(self.S):
                self._wait_slice(k)
                self._rebuild_block_counts(k)
        self.topic_sums = torch.zeros(self.K, dtype=torch.int64, device=self.dev)
        self.commons = torch.zeros(self.K, dtype=torch.float32, device=self.dev)
        self.trained = torch.zeros(1, dtype=torch.int64, device=self.dev)
        self.epochs_fused = 0
        if cuda:
            flat = [self.panels[s][k] for s in range(N) for k in range(S)]
            self.desc = L.tl_panel_descriptors(flat) if self.tl else L.panel_descriptors(flat)
        self._refresh_topic_sums()
        if dist.is_initialized() and N > 1:
            dist.barrier(group=group)

    @property
    def n_dk(self) -> torch.Tensor:
        """Document-topic counts.  The token-list sampler keeps no such table: it is rebuilt from the documents' token
        lists on demand (tests, model dumps)."""
        if self._n_dk is not None:
            return self._n_dk
        t = torch.zeros(self.num_docs_local, self.K, dtype=torch.int32, device=self.dev)
        t.view(-1).index_add_(0, self.doc_of_pos * self.K + self.doc_z.long(), torch.ones(self.ntok, dtype=torch.int32, device=self.dev))
        return t

    @property
    def cur(self) -> List[int]:
        return self.ring.cur if self.ring is not None else self._cur

    @property
    def rot(self) -> List[int]:
        return [s.rotations for s in self.sched]

    def _visiting(self, k: int) -> int:
        return self.sched[k].shard_at[self.me]

    def _wait_slice(self, k: int) -> None:
        if self.N > 1 and self.symm is not None and self.dev.type == "cuda":
            torch.cuda.current_stream().wait_event(self.ev_comm[k])

    def _rotate(self, k: int) -> None:
        srcs = self.sched[k].rotate()
        if self.N == 1:
            return
        src, dst = self.Wbuf[k][self.cur[k]], self.Wbuf[k][1 - self.cur[k]]
        pred, succ = srcs[self.me], srcs.index(self.me)
        if self.symm is not None and self.dev.type == "cuda":
            # on the stream of every other collective of this workspace (its flags / tickets are per workspace);
            # the per-panel path is the A/B arm, the whole-epoch kernel is what overlaps the transfer
            self.symm.rotate_(dst, src, pred)
        else:
            ops = [dist.P2POp(dist.isend, src.contiguous(), succ, group=self.group),
                   dist.P2POp(dist.irecv, dst, pred, group=self.group)]
            for r in dist.batch_isend_irecv(ops):
                r.wait()
        self.cur[k] ^= 1

    def _rebuild_block_counts(self, k: int) -> None:
        """Block counts of slice k's current buffer from its word rows (the kernel keeps them exact afterwards)."""
        W, Bc = self.Wbuf[k][self.cur[k]], self.Bbuf[k][self.cur[k]]
        step = max(1, (1 << 28) // max(self.K, 1))              # bounded scratch: 1 GB of rows at a time
        for r0 in range(0, self.cap, step):
            w = W[r0:r0 + step]
            full = (self.K // 32) * 32
            Bc[r0:r0 + step, :self.K // 32] = w[:, :full].view(w.shape[0], -1, 32).sum(-1, dtype=torch.int32)
            if full < self.K:
                Bc[r0:r0 + step, self.K // 32] = w[:, full:].sum(-1, dtype=torch.int32)

    def _refresh_topic_sums(self) -> None:
        self.topic_sums.zero_()
        for k in range(self.S):
            self._wait_slice(k)
            L.topic_sums(self.Wbuf[k][self.cur[k]], self.topic_sums)
        if self.N > 1:
            dist.all_reduce(self.topic_sums, group=self.group)
        self.commons.copy_((1.0 / (self.topic_sums.double() + self.beta * self.V)).float())

    def iterate(self, deadline_ms: Optional[float] = None) -> torch.Tensor:
        """One iteration = N rotations x S slices, then the global topic-sum refresh."""
        self.trained.zero_()
        self.iteration += 1
        if self.fused and self.dev.type == "cuda":
            N, S = self.N, self.S
            pos_l, src_l = [], []
            if self.Nv:
                N = self.Nv
                pos_l = [j * S + k for j in range(N) for k in range(S)]
                src_l = [0] * (N * S)
            else:
                for _ in range(N):
                    for k in range(S):
                        pos_l.append(self.sched[k].shard_at[self.me] * S + k)
                        src_l.append(self.sched[k].rotate()[self.me])
            for k in range(S):
                self._wait_slice(k)
            seed_it = self.seed * 1000003 + self.iteration * 7919 + 104729 * self.me
            if self.tl:
                L.lda_tl_epoch(self.doc_z, self.d_ptr, self.commons, self.K, self.desc, pos_l, S, self.B, self.cs, self.alpha,
                               self.beta, seed_it, self.epochs_fused, deadline_ms, self.trained, self.ring, src_l,
                               wb_off=self.cap * self.K if self.nb else 0)
            else:
                L.lda_epoch(self._n_dk, self.commons, self.desc, pos_l, S, self.B, self.cs, self.alpha, self.beta,
                            seed_it, self.epochs_fused, deadline_ms, self.trained, self.ring, src_l)
            self.epochs_fused += 1
            self.ring.advance(N)
            tb = (self.epochs_fused * self.B) & 0x7FFFFFFF
            for per in self.panels:
                for p in per:
                    p.ticket_base = tb
            self._refresh_topic_sums()
            return self.trained
        for j in range(self.N):
            for k in range(self.S):
                self._wait_slice(k)
                p = self.panels[self._visiting(k)][k]
                L.lda_pass(p, self.n_dk, self.Wbuf[k][self.cur[k]], self.commons, self.alpha, self.beta,
                           self.seed * 1000003 + self.iteration * 7919 + j * 2 + k + 104729 * self.me, deadline_ms, self.trained)
                self._rotate(k)
        self._refresh_topic_sums()
        return self.trained

    def log_likelihood(self) -> float:
        """Word-side model log-likelihood (MALLET ``modelLogLikelihood`` form used by the reference):
        ``Σ_k Σ_w lgamma(β + n_wk) − Σ_k lgamma(Vβ + n_k) + K lgamma(Vβ) − K V lgamma(β)`` plus the
        document side ``Σ_d [Σ_k lgamma(α + n_dk) − lgamma(Kα + n_d)] + D lgamma(Kα) − D K lgamma(α)``."""
        b, a, K, V = self.beta, self.alpha, self.K, self.V
        acc = torch.zeros(1, dtype=torch.float64, device=self.dev)
        for k in range(self.S):
            self._wait_slice(k)
            # K11: one fused pass over the int32 table (rows beyond the vocabulary are all zero and add nothing);
            # Σ lgamma(n + β) − (rows K) lgamma(β) = Σ_{n>0} [lgamma(n + β) − lgamma(β)]
            sm, nz = L.lgamma_sum(self.Wbuf[k][self.cur[k]], b)
            acc += sm - nz * math.lgamma(b)
        if self.tl:
            # sparse document side from the token lists: counts of the distinct (document, topic) pairs
            _, cnt = torch.unique(self.doc_of_pos * K + self.doc_z.long(), return_counts=True)
            lens = (self.d_ptr[1:] - self.d_ptr[:-1]).double()
            doc = ((torch.lgamma(cnt.double() + a) - math.lgamma(a)).sum() - torch.lgamma(lens + K * a).sum()
                   + lens.numel() * math.lgamma(K * a))
        else:
            nd = self._n_dk.double()
            doc = (torch.lgamma(nd + a).sum() - torch.lgamma(nd.sum(1) + K * a).sum()
                   + nd.shape[0] * (math.lgamma(K * a) - K * math.lgamma(a)))
        acc += doc
        if self.N > 1:
            dist.all_reduce(acc, group=self.group)
        acc += -(torch.lgamma(self.topic_sums.double() + V * b)).sum() + K * math.lgamma(V * b)
        return float(acc.item())

    def close(self) -> None:
        """Gives the engine's block of the symmetric workspace back (every rank, same order); unusable afterwards."""
        self.synchronize()
        if self.ring is not None:
            self.ring.release()

    def synchronize(self) -> None:
        if self.dev.type == "cuda":
            for k in range(self.S):
                self._wait_slice(k)
            torch.cuda.synchronize(self.dev)
            if self.ring is not None and self.fused:
                self.ring.check()

    def home_word_rows(self):
        """``(global word ids, count rows)`` of the word-topic rows currently resident."""
        ids, rows = [], []
        for k in range(self.S):
            self._wait_slice(k)
            s = self._visiting(k)
            gid = torch.arange(self.cap, device=self.dev) * (self.S * self.N) + k * self.N + s
            keep = gid < self.V
            ids.append(gid[keep]); rows.append(self.Wbuf[k][self.cur[k]][keep])
        return torch.cat(ids), torch.cat(rows)
