"""Graph / embedding applications: colour-coding subgraph counting (``edu.iu.subgraph`` and the three SAHAD
rotation variants ``edu.iu.sahad.rotation{,2,3}``), WDA-MDS (``edu.iu.wdamds``) and the SGX-cost-model K-means of
``experimental/`` (``edu.iu.kmeans.sgxsimu``).

Subgraph counting (FASCIA colour coding, ``ml/java/.../subgraph/colorcount_HJ.java:861-925``): the tree template is cut
recursively into (active, passive) sub-templates; for a random colouring with k colours
``cnt_S[v][C] = Σ_{C = Ca ⊔ Cp} cnt_a[v][Ca] · Σ_{u ∈ N(v)} cnt_p[u][Cp]``, and the estimate is
``Σ_v cnt_T[v][all] / (P(colourful) · |Aut(T)|)`` (``SCCollectiveMapper.java:203-214``).  Tables are dense
``[vertices, 2^k]`` tensors; the neighbour sum is a sparse-matrix product, so one DP step is one SpMM plus
gathers — GPU shaped.  Remote neighbours' counts arrive either by ``allgather`` of the passive table
(``comm="allgather"``, the regroup-all-at-once mode, ``:1303-1432``) or by **rotating** the passive shards
round the ring and accumulating partial neighbour sums (``comm="rotation"``, SAHAD's model,
``sahad/rotation/SCCollectiveMapper.java:490``).
"""
from __future__ import annotations

import itertools
import math
import time
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

from ..io import formats as F
from ..mapper import CollectiveMapper, Context, KeyValReader, LOG
from ..partition import DoubleArrPlus, FnCombiner, Partition, Partitioner, Table
from .kmeans import KMeansRegroupAllgatherMapper


# ---------------------------------------------------------------------------------------
# template handling
# ---------------------------------------------------------------------------------------
def automorphisms(nv: int, edges: Sequence[Tuple[int, int]]) -> int:
    es = {frozenset(e) for e in edges}
    cnt = 0
    for perm in itertools.permutations(range(nv)):
        if all(frozenset((perm[a], perm[b])) in es for a, b in edges):
            cnt += 1
    return cnt


class SubTemplate:
    """Rooted sub-template with its (active, passive) split."""

    def __init__(self, root: int, verts: frozenset, adj: Dict[int, List[int]]):
        self.root, self.verts, self.size = root, verts, len(verts)
        self.active: Optional["SubTemplate"] = None
        self.passive: Optional["SubTemplate"] = None
        if self.size > 1:
            child = min(c for c in adj[root] if c in verts)          # cut the edge (root, first child)
            sub = self._component(child, verts - {root}, adj)
            self.passive = SubTemplate(child, sub, adj)
            self.active = SubTemplate(root, verts - sub, adj)

    @staticmethod
    def _component(start, allowed, adj):
        seen, stack = {start}, [start]
        while stack:
            v = stack.pop()
            for u in adj[v]:
                if u in allowed and u not in seen:
                    seen.add(u); stack.append(u)
        return frozenset(seen)


def colour_coding_pass(nbr_ptr: torch.Tensor, nbr_idx: torch.Tensor, colours: torch.Tensor, template: SubTemplate, k: int,
                       fetch_rows=None, presummed=None) -> torch.Tensor:
    """One colouring.  ``nbr_ptr / nbr_idx``: CSR of the locally owned vertices; ``nbr_idx`` names the ROW of every
    neighbour in the table that ``fetch_rows(local table)`` returns (local rows first, then the remote rows this rank
    needs; default: the local table itself).  ``presummed(local table) -> [n_local, cols]``, when given, supplies the
    neighbour sums instead (rotation mode: the shards visit one after the other).  Tables are compressed: a
    sub-template of size s has one column per colour set of size s (``ops/subgraph.py``); returns the ``[n, 1]``
    table of the whole template (the only colour set of size k is all colours)."""
    from ..ops import subgraph as SG
    n = colours.numel()
    dev = colours.device
    dt = torch.float32 if dev.type == "cuda" else torch.float64
    self_ptr = torch.arange(n + 1, dtype=torch.int64, device=dev)
    self_idx = torch.arange(n, dtype=torch.int32, device=dev)

    def solve(t: SubTemplate) -> torch.Tensor:
        if t.size == 1:
            tab = torch.zeros(n, k, dtype=dt, device=dev)            # colour sets of size 1 in mask order = colours 0..k-1
            tab[torch.arange(n, device=dev), colours.long()] = 1.0
            return tab
        act = solve(t.active)
        pas = solve(t.passive)
        sa, sp = SG.split_tables(k, t.size, t.active.size, dev)
        if presummed is not None:
            return SG.combine(self_ptr, self_idx, act, presummed(pas), sa, sp)
        return SG.combine(nbr_ptr, nbr_idx, act, pas if fetch_rows is None else fetch_rows(pas), sa, sp)
    return solve(template)


class SubgraphMapper(CollectiveMapper):
    """conf: ``template_file``, ``num_iterations`` (colourings), ``comm`` (``exchange`` = default, also selected by the
    old name ``allgather``: every rank receives exactly the passive-table rows of its remote neighbours, compressed;
    ``rotation``: the passive shards visit every worker in turn), ``seed``; input = adjacency files
    ``v\\tn1,n2,...,``; vertices are owned by ``v % N``."""

    def setup(self, context: Context) -> None:
        c = context.get_configuration()
        self.nv, self.edges = F.read_template(c["template_file"])
        self.iters = int(c.get("num_iterations", 10))
        self.mode = c.get("comm", "exchange")
        self.seed = int(c.get("seed", 0))

    def map_collective(self, reader: KeyValReader, context: Context) -> None:
        N, me, dev = self.get_num_workers(), self.get_self_id(), self.device
        g: Dict[int, np.ndarray] = {}
        while reader.next_key_value():
            g.update(F.read_adjacency(reader.get_current_value()))
        # every vertex's adjacency list goes to its owner (regroup with list-append combiner)
        t = Table(0, FnCombiner(lambda a, b: {**a, **b}))
        for w in range(N):
            part = {v: nb for v, nb in g.items() if v % N == w}
            if part:
                t.add_partition(Partition(w, part))
        self.regroup("sc", "regroup-graph", t, Partitioner(N))
        mine = t.get_partition(me).get() if t.get_partition(me) is not None else {}
        max_v = max([max([v] + nb.tolist()) for v, nb in mine.items()] + [-1])
        n_total = max(self.comm.all_gather_object(max_v)) + 1
        own = torch.arange(me, n_total, N)
        n_loc = own.numel()
        rows, cols = [], []
        for v, nb in mine.items():
            rows.extend([v // N] * len(nb)); cols.extend(nb.tolist())
        k = self.nv
        adj_t: Dict[int, List[int]] = {i: [] for i in range(k)}
        for a, b in self.edges:
            adj_t[a].append(b); adj_t[b].append(a)
        tmpl = SubTemplate(0, frozenset(range(k)), adj_t)
        aut = automorphisms(k, self.edges)
        p_colourful = math.factorial(k) / float(k ** k)
        ri, ci = torch.tensor(rows, dtype=torch.long), torch.tensor(cols, dtype=torch.long)
        order = torch.argsort(ri, stable=True)
        ri, ci = ri[order], ci[order]
        nbr_ptr = torch.zeros(n_loc + 1, dtype=torch.int64)
        nbr_ptr[1:] = torch.cumsum(torch.bincount(ri, minlength=n_loc), 0)
        # remote neighbours: the sorted distinct ids per owner; their rows follow the local rows in the fetched table
        need: Dict[int, torch.Tensor] = {}
        row_of = torch.empty(ci.numel(), dtype=torch.long)
        local = (ci % N) == me
        row_of[local] = ci[local] // N
        base = n_loc
        for r in range(N):
            if r == me:
                continue
            m = (ci % N) == r
            ids = torch.unique(ci[m])
            need[r] = ids
            if ids.numel():
                row_of[m] = base + torch.searchsorted(ids, ci[m])
            base += ids.numel()
        # owners learn once which of their rows each peer wants (the adjacency never changes)
        serve: Dict[int, torch.Tensor] = {}
        if N > 1:
            got = self.comm.exchange({r: [(0, need[r].to(dev))] for r in need if need[r].numel()})
            for src, _pid, ids in got:
                serve[src] = (ids.to(dev).long() // N)
        nbr_ptr_d, nbr_idx_d = nbr_ptr.to(dev), row_of.to(torch.int32).to(dev)
        # per-shard adjacency blocks for the rotation mode: columns owned by worker s, re-indexed locally
        A_shard = []
        if self.mode == "rotation":
            vals = torch.ones(ri.numel(), dtype=torch.float32 if dev.type == "cuda" else torch.float64)
            for s_ in range(N):
                m = (ci % N) == s_
                ns = (n_total - s_ + N - 1) // N if n_total > s_ else 0
                A_shard.append(torch.sparse_coo_tensor(torch.stack([ri[m], ci[m] // N]), vals[m], (n_loc, max(ns, 1))).coalesce().to(dev))
        self.estimates: List[float] = []
        self.comm_time = 0.0
        self.exchanged_bytes = 0
        step = [0]

        def fetch_rows(tab: torch.Tensor) -> torch.Tensor:
            if N == 1:
                return tab
            t0 = time.perf_counter()
            got = self.comm.exchange({r: [(0, tab[idx])] for r, idx in serve.items()})
            step[0] += 1
            by_src = {src: payload.to(dev) for src, _pid, payload in got}
            parts = [tab] + [by_src[r] for r in range(N) if r != me and need[r].numel()]
            self.exchanged_bytes += sum(int(p_.numel()) * p_.element_size() for p_ in parts[1:])
            self.comm_time += time.perf_counter() - t0
            return torch.cat(parts, 0)

        def nsum_rotation(tab: torch.Tensor) -> torch.Tensor:
            acc = torch.zeros(n_loc, tab.shape[1], dtype=tab.dtype, device=dev)
            tbl = Table(1, FnCombiner(lambda a, b: a))
            tbl.add_partition(Partition(me, tab.clone()))
            for j in range(N):
                (p,) = tbl.get_partitions()
                s_ = p.id()
                pt = p.get().to(dev)
                if pt.shape[0]:
                    acc += torch.sparse.mm(A_shard[s_][:, :pt.shape[0]] if A_shard[s_].shape[1] != pt.shape[0] else A_shard[s_], pt)
                t0 = time.perf_counter()
                self.rotate("sc", f"rotate-{step[0]}", tbl, None); step[0] += 1
                self.comm_time += time.perf_counter() - t0
            return acc
        for it in range(self.iters):
            gcol = torch.Generator().manual_seed(self.seed * 7919 + it)           # same colouring on all workers
            colours_all = torch.randint(0, k, (n_total,), generator=gcol)
            tab = colour_coding_pass(nbr_ptr_d, nbr_idx_d, colours_all[own].to(dev), tmpl, k,
                                     fetch_rows=fetch_rows if self.mode != "rotation" else None,
                                     presummed=nsum_rotation if (self.mode == "rotation" and N > 1) else None)
            tot = Table(2, DoubleArrPlus())
            tot.add_partition(Partition(0, tab[:, 0].double().sum().view(1)))
            self.allreduce("sc", f"allreduce-count-{it}", tot)
            self.estimates.append(float(tot.get_partition(0).get().item()) / (p_colourful * aut))
            context.progress()
        self.count = float(np.mean(self.estimates))
        if self.is_master():
            context.write("subgraph_count", self.count)


# ---------------------------------------------------------------------------------------
# WDA-MDS (weighted deterministic-annealing SMACOF)
# ---------------------------------------------------------------------------------------
class WDAMDSMapper(CollectiveMapper):
    """conf: ``n`` (points), ``target_dim``, ``max_iter``, ``threshold``, ``alpha`` (cooling, default 0.95), ``cg_iter``;
    input = CSV row blocks of the full distance matrix: each line ``row_id, d_0, ..., d_{n-1}``.

    Outer loop anneals the temperature, inner loop runs SMACOF: ``B(X) X`` on the local rows
    (``BCCalcTask``), the weighted Laplacian system solved by conjugate gradients whose matrix-vector
    products are row-distributed with an ``allgather`` of X per step (``MMCalcTask``,
    ``WDAMDSMapper.java:585-725``), stress by ``allreduce`` (``StressCalcTask.java:44-101``)."""

    def setup(self, context: Context) -> None:
        c = context.get_configuration()
        self.n = int(c["n"]); self.dim = int(c.get("target_dim", 3))
        self.max_iter = int(c.get("max_iter", 100)); self.threshold = float(c.get("threshold", 1e-6))
        self.alpha = float(c.get("alpha", 0.95)); self.cg_iter = int(c.get("cg_iter", 20))
        self.seed = int(c.get("seed", 0))

    def _gather_rows(self, local_rows: torch.Tensor, ids: torch.Tensor, op: str) -> torch.Tensor:
        tbl = Table(0, FnCombiner(lambda a, b: a))
        tbl.add_partition(Partition(self.get_self_id(), torch.cat([ids.double()[:, None], local_rows], 1)))
        self.allgather("mds", op, tbl)
        full = torch.zeros(self.n, local_rows.shape[1], dtype=torch.float64, device=local_rows.device)
        for p in tbl.get_partitions():
            blk = p.get().to(local_rows.device)
            full[blk[:, 0].long()] = blk[:, 1:]
        return full

    def _allsum(self, v: torch.Tensor, op: str) -> torch.Tensor:
        t = Table(0, DoubleArrPlus()); t.add_partition(Partition(0, v.clone()))
        self.allreduce("mds", op, t)
        return t.get_partition(0).get()

    def map_collective(self, reader: KeyValReader, context: Context) -> None:
        dev = self.device
        mats = []
        while reader.next_key_value():
            m = F.read_dense_csv(reader.get_current_value())
            if m.size:
                mats.append(m)
        blk = np.concatenate(mats) if mats else np.empty((0, self.n + 1))
        ids = torch.from_numpy(blk[:, 0].copy()).long().to(dev)
        D = torch.from_numpy(blk[:, 1:].copy()).to(dev)                      # [rows_local, n]
        W = torch.ones_like(D)
        W[torch.arange(ids.numel(), device=dev), ids] = 0.0
        n, d = self.n, self.dim
        g = torch.Generator().manual_seed(self.seed)
        X = torch.rand(n, d, generator=g, dtype=torch.float64).to(dev)         # replicated
        step = 0
        # temperature schedule: start where all annealed distances vanish, cool geometrically to 0
        dmax = float(self._allsum(torch.tensor([float(D.max()) if D.numel() else 0.0], dtype=torch.float64, device=dev), "max0")[0])
        # (at T = dmax / sqrt(2d) every annealed distance is zero and SMACOF has only the trivial fixed point,
        #  so the schedule starts at half of that)
        T = 0.5 * dmax / math.sqrt(2.0 * d) if dmax > 0 else 0.0
        vdiag_local = W.sum(1)                                                  # V = diag(rowsum W) - W
        self.stress_history = []

        def v_times(Y_full: torch.Tensor) -> torch.Tensor:                     # rows of V·Y owned here
            return vdiag_local[:, None] * Y_full[ids] - W @ Y_full
        it_total = 0
        from ..ops import mds as KM
        D32 = D.float() if dev.type == "cuda" else D                            # the n_local x n matrix is read in fp32 there
        while True:
            shift = T * math.sqrt(2.0 * d) if T > 0 else 0.0
            prev = float("inf")
            for _ in range(self.max_iter):
                # B(X) X on local rows (b_ii = Σ_j ratio_ij, b_ij = -ratio_ij): one fused pass over the distance rows
                BX = KM.bc_stress(D32, ids, X, shift)[0].double()
                # 1'B(X)X = 0 exactly; the fp32 pass leaves ~1e-7 of it, which lies in the null space of V and would
                # blow the conjugate gradient up: remove the global column mean before solving
                BX = BX - self._allsum(BX.sum(0), f"bx-mean-{step}") / float(n); step += 1
                # solve V X_new = BX by CG (V is singular along constants: work in the centred subspace)
                Xn = X.clone()
                R_loc = BX - v_times(Xn)
                R = self._gather_rows(R_loc, ids, f"cg-r-{step}"); step += 1
                Pm = R.clone()
                rs = (R * R).sum()
                for _cg in range(self.cg_iter):
                    if float(rs) < 1e-20:
                        break
                    VP = self._gather_rows(v_times(Pm), ids, f"cg-vp-{step}"); step += 1
                    a = rs / (Pm * VP).sum().clamp(min=1e-300)
                    Xn = Xn + a * Pm
                    R = R - a * VP
                    rs_new = (R * R).sum()
                    Pm = R + (rs_new / rs) * Pm
                    rs = rs_new
                X = Xn - Xn.mean(0)
                if float(X.abs().max()) < 1e-12:                     # collapsed: restart from a small random layout
                    X = (torch.rand(n, d, generator=g, dtype=torch.float64).to(dev) - 0.5) * 1e-2
                st = self._allsum(KM.bc_stress(D32, ids, X, shift)[1], f"stress-{step}"); step += 1
                stress = float(st[0] / st[1].clamp(min=1e-300))
                it_total += 1
                if abs(prev - stress) < self.threshold:
                    break
                prev = stress
            self.stress_history.append((T, stress))
            context.progress()
            if T <= 0:
                break
            T = T * self.alpha if T * self.alpha > 1e-3 * dmax else 0.0
        self.X = X
        self.stress = self.stress_history[-1][1]
        self.iterations = it_total


# ---------------------------------------------------------------------------------------
# K-means with the SGX enclave cost model (experimental/.../kmeans/sgxsimu)
# ---------------------------------------------------------------------------------------
class SGXSimuKMeansMapper(KMeansRegroupAllgatherMapper):
    """Ordinary regroup-allgather K-means plus the reference's *cost model* of running inside SGX enclaves:
    Ecall/Ocall, cross-enclave copy per KB, enclave creation, attestation and page-swap penalties, in
    kilo-cycles converted with ``ms_per_kcycle`` (``Constants.java``, ``KMeansCollectiveMapper.java:279-345``).
    The overhead is accounted (and optionally slept, ``conf["sgx_sleep"]``) per iteration."""

    ECALL, OCALL, CROSS_PER_KB = 8.5, 9.0, 1.4
    CREATE_FIX, CREATE_PER_KB = 221000.0, 22.677
    LOCAL_ATT, REMOTE_ATT, SWAP_PAGE = 80.0, 27200.0, 40.0
    MS_PER_KCYCLE = 0.0002941

    def setup(self, context: Context) -> None:
        super().setup(context)
        c = context.get_configuration()
        self.enclave_total_mb = float(c.get("enclave_total_mb", 96.0))
        self.sleep = bool(c.get("sgx_sleep", False))
        self.sgx_overhead_ms = {"init": 0.0, "comm": 0.0, "compute": 0.0}

    def _simu(self, key: str, ms: float) -> None:
        self.sgx_overhead_ms[key] += ms
        if self.sleep and ms > 0:
            time.sleep(ms / 1e3)

    def run_kmeans(self, files, context) -> None:
        N = self.get_num_workers()
        cen_kb = self.num_centroids * self.cen_vec_size * 8 / 1024.0
        # enclave creation + attestation once per worker
        self._simu("init", (self.CREATE_FIX + self.CREATE_PER_KB * self.enclave_total_mb * 1024 + self.LOCAL_ATT
                            + (self.REMOTE_ATT if N > 1 else 0.0)) * self.MS_PER_KCYCLE)
        super().run_kmeans(files, context)
        pts_kb = getattr(self, "_points_kb", 0.0)
        for _ in range(self.num_iterations):
            # regroup: (N-1) sends + (N-1) receives of 1/N of the table; allgather: the whole table
            self._simu("comm", ((N - 1) * (self.OCALL + self.ECALL) + cen_kb / max(N, 1) * (2 * N - 2) * self.CROSS_PER_KB) * self.MS_PER_KCYCLE)
            self._simu("comm", ((self.OCALL + self.ECALL * (N - 1)) + cen_kb * self.CROSS_PER_KB) * self.MS_PER_KCYCLE)
            # data larger than the enclave page cache is swapped page by page (4 KB) every pass
            over_kb = max(0.0, pts_kb - self.enclave_total_mb * 1024)
            self._simu("compute", (self.ECALL + over_kb / 4.0 * self.SWAP_PAGE) * self.MS_PER_KCYCLE)

    def _load_points(self, files):
        pts = super()._load_points(files)
        self._points_kb = pts.numel() * 8 / 1024.0
        return pts
