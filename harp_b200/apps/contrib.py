"""Tutorial applications of ``contrib/`` (besides K-means, see ``apps/kmeans.py``).

* :class:`PageRankMapper` — ``edu.iu.simplepagerank`` (``PageRankMapper.java:60-190``): ``Long2DoubleKVTable``,
  allgather of the initial ranks, then per iteration local contributions → ``allreduce`` → damping.
* :class:`IterativeSVMMapper` — ``edu.iu.svm`` (``SVMMapper.java:78-209``): every worker trains a C-SVC on
  (local data ∪ global support vectors), support-vector sets are merged with ``allreduce`` of a
  concatenating combiner (``HarpStringPlus``), repeated.
* :class:`MLRMapper` — ``edu.iu.mlr`` (``MLRMapper.java``, ``GDtask.java:26-51``, ``EVtask.java:27-59``): one
  binary logistic regressor per topic; the weight table (partition id = topic) **rotates** so every
  worker's documents update every topic once per iteration.
* :class:`LDACVBMapper` — ``edu.iu.lda`` (contrib): collapsed variational Bayes (Mr.LDA style) with a
  log-domain φ table combined by log-add (``DoubleArrLogPlus``); static (allreduce) and dynamic
  (push/pull) variants.
* :class:`NNMapper` — ``edu.iu.NN``: MLP trained locally, weights averaged every ``sync_iter`` mini-batches
  (``NNMapper.java:143-179``).
* :class:`RandomForestMapper` — ``edu.iu.rf`` / ``edu.iu.randomforest`` / ``com.rf.fast``: local bagged trees,
  test-set votes reduced (``RFMapper.java:117``).
"""
from __future__ import annotations

import math
import os
from typing import Dict, List

import numpy as np
import torch

from ..daal import classify as DC
from ..daal.models import MLP
from ..io import formats as F
from ..keyval import Long2DoubleKVTable, TypeCombiner
from ..mapper import CollectiveMapper, Context, KeyValReader
from ..partition import DoubleArrPlus, FnCombiner, LogSumExpCombiner, Partition, Partitioner, Table


def _files(reader: KeyValReader) -> List[str]:
    out = []
    while reader.next_key_value():
        out.append(reader.get_current_value())
    return out


# ---------------------------------------------------------------------------------------
class PageRankMapper(CollectiveMapper):
    """conf: ``num_urls``, ``num_iterations``, ``tables`` (``kv`` | ``dense`` | ``auto`` = dense on GPUs); input lines
    ``src dst dst ...`` (no targets = dangling)."""

    def setup(self, context: Context) -> None:
        c = context.get_configuration()
        self.num_urls = int(c["num_urls"])
        self.num_iterations = int(c["num_iterations"])
        self.tables = c.get("tables", "auto")          # "kv": the reference's key-value tables; "dense": rank vector + push kernel

    def _run_dense(self, graph: Dict[int, np.ndarray], context: Context) -> None:
        """Device formulation: dense rank vector, CSR of the local pages, one push kernel + one allreduce per iteration
        (``csrc/graph.cu::pagerank_push_kernel``); same arithmetic as the key-value tables of the reference."""
        from ..ops import graphops as GO
        dev, n = self.device, self.num_urls
        srcs = sorted(graph)
        deg = [graph[s_].size for s_ in srcs]
        ptr = torch.zeros(len(srcs) + 1, dtype=torch.int64)
        ptr[1:] = torch.cumsum(torch.tensor(deg, dtype=torch.int64), 0) if srcs else ptr[1:]
        idx = torch.from_numpy(np.concatenate([graph[s_] for s_ in srcs]).astype(np.int32)) if srcs and sum(deg) else torch.zeros(0, dtype=torch.int32)
        ptr, idx, src_ids = ptr.to(dev), idx.to(dev), torch.tensor(srcs, dtype=torch.int64, device=dev)
        # initial table: 1/n on every page that appears as a source somewhere (allgather_initPRTable)
        pr = torch.zeros(n, dtype=torch.float64, device=dev)
        pr[src_ids] = 1.0 / n
        t0 = Table(0, DoubleArrPlus()); t0.add_partition(Partition(0, pr))
        self.allreduce("harp-pagerank", "allgather_initPRTable", t0)
        pr = t0.get_partition(0).get()
        for it in range(self.num_iterations):
            buf = torch.zeros(n + 1, dtype=torch.float64, device=dev)          # [ranks | dangling mass]
            GO.pagerank_push(ptr, idx, src_ids, pr, buf[:n], buf[n:])
            tt = Table(1, DoubleArrPlus()); tt.add_partition(Partition(0, buf))
            self.allreduce("harp-pagerank", f"allgather_{it}", tt)
            buf = tt.get_partition(0).get()
            pr = 0.85 * (buf[:n] + buf[n] / n) + 0.15 / n
            context.progress()
        self.ranks = pr.cpu()
        if self.is_master():
            for i in range(n):
                context.write(i, repr(float(self.ranks[i])))

    def map_collective(self, reader: KeyValReader, context: Context) -> None:
        graph: Dict[int, np.ndarray] = {}
        for f in _files(reader):
            graph.update(F.read_pagerank_graph(f))
        n = self.num_urls
        if self.tables == "dense" or (self.tables == "auto" and self.device.type == "cuda"):
            return self._run_dense(graph, context)
        local = Long2DoubleKVTable(1, TypeCombiner("sum"))
        if graph:
            local.add_batch(torch.tensor(sorted(graph)), torch.full((len(graph),), 1.0 / n, dtype=torch.float64))
        self.allgather("harp-pagerank", "allgather_initPRTable", local)
        src = torch.tensor(sorted(graph), dtype=torch.long)
        for it in range(self.num_iterations):
            glob = Long2DoubleKVTable(2, TypeCombiner("sum"))
            keys, vals = [], []
            for s in sorted(graph):
                targets = graph[s]
                pr = local.get_val(s)
                if targets.size == 0:                       # dangling page: spread over everybody
                    keys.append(torch.arange(n)); vals.append(torch.full((n,), pr / n, dtype=torch.float64))
                else:
                    keys.append(torch.from_numpy(targets)); vals.append(torch.full((targets.size,), pr / targets.size, dtype=torch.float64))
            if keys:
                glob.add_batch(torch.cat(keys), torch.cat(vals))
            self.allreduce("harp-pagerank", f"allgather_{it}", glob)
            # damping (updatePRTable): pages nobody links to still get the teleport share
            new = Long2DoubleKVTable(1, TypeCombiner("sum"))
            ks = torch.arange(n)
            got = torch.zeros(n, dtype=torch.float64)
            for k, v in glob.items():
                if 0 <= k < n:
                    got[k] = v
            new.add_batch(ks, 0.85 * got + 0.15 / n)
            local = new
            context.progress()
        self.ranks = torch.tensor([local.get_val(i) for i in range(n)], dtype=torch.float64)
        if self.is_master():
            for i in range(n):
                context.write(i, repr(float(self.ranks[i])))


# ---------------------------------------------------------------------------------------
class IterativeSVMMapper(CollectiveMapper):
    """conf: ``num_iterations``, ``C`` (default 1), ``kernel`` (``linear``); input = libsvm files, binary ±1 labels."""

    def setup(self, context: Context) -> None:
        c = context.get_configuration()
        self.num_iterations = int(c.get("num_iterations", 5))
        self.C = float(c.get("C", 1.0))
        self.kernel = c.get("kernel", "linear")
        self.num_features = c.get("num_features")

    def map_collective(self, reader: KeyValReader, context: Context) -> None:
        Xs, ys = [], []
        for f in _files(reader):
            X, y = F.read_libsvm(f, int(self.num_features) if self.num_features else None)
            Xs.append(X); ys.append(y)
        nf = max((x.shape[1] for x in Xs), default=0)
        nf = max(self.comm.all_gather_object(nf))
        Xl = torch.from_numpy(np.concatenate([np.pad(x, ((0, 0), (0, nf - x.shape[1]))) for x in Xs])) if Xs else torch.zeros(0, nf, dtype=torch.float64)
        yl = torch.from_numpy(np.concatenate(ys)) if ys else torch.zeros(0, dtype=torch.float64)
        sv = torch.zeros(0, nf + 1, dtype=torch.float64)          # rows [label, features]
        self.history = []
        for it in range(self.num_iterations):
            data = torch.cat([torch.cat([yl[:, None], Xl], 1), sv], 0)
            data = torch.unique(data, dim=0)
            model = DC.svm_train_binary(data[:, 1:], data[:, 0], self.C, self.kernel, tol=1e-3)
            labels = torch.sign(model.coef)
            mine = torch.cat([labels[:, None], model.sv], 1)
            t = Table(0, FnCombiner(lambda a, b: torch.cat([a, b], 0)))      # HarpStringPlus: concatenate SV sets
            t.add_partition(Partition(0, mine))
            self.allreduce("main", f"allreduce-sv-{it}", t)
            sv = torch.unique(t.get_partition(0).get(), dim=0)
            self.history.append(int(sv.shape[0]))
            self.model = model
            context.progress()
        acc = torch.tensor([float((torch.sign(self.model.decision(Xl)) == yl).sum()) if yl.numel() else 0.0, float(yl.numel())],
                           dtype=torch.float64)
        tt = Table(1, DoubleArrPlus()); tt.add_partition(Partition(0, acc))
        self.allreduce("main", "allreduce-acc", tt)
        a = tt.get_partition(0).get()
        self.train_accuracy = float(a[0] / max(float(a[1]), 1.0))
        self.support_vectors = sv


# ---------------------------------------------------------------------------------------
class MLRMapper(CollectiveMapper):
    """conf: ``num_topics``, ``num_terms``, ``num_iterations``, ``alpha`` (learning rate); input = libsvm-like
    ``topic[,topic] idx:val ...`` reduced here to one topic id per row (``label idx:val``)."""

    def setup(self, context: Context) -> None:
        c = context.get_configuration()
        self.T = int(c["num_topics"]); self.D = int(c["num_terms"])
        self.iters = int(c.get("num_iterations", 10)); self.lr = float(c.get("alpha", 0.5))

    def map_collective(self, reader: KeyValReader, context: Context) -> None:
        Xs, ys = [], []
        for f in _files(reader):
            X, y = F.read_libsvm(f, self.D)
            Xs.append(X); ys.append(y)
        X = torch.from_numpy(np.concatenate(Xs)) if Xs else torch.zeros(0, self.D, dtype=torch.float64)
        y = torch.from_numpy(np.concatenate(ys)).long() if ys else torch.zeros(0, dtype=torch.long)
        Xb = torch.cat([X, torch.ones(X.shape[0], 1, dtype=torch.float64)], 1)          # bias term
        N, me = self.get_num_workers(), self.get_self_id()
        table = Table(0, DoubleArrPlus())
        for t in range(self.T):
            if t % N == me:
                table.add_partition(Partition(t, torch.zeros(self.D + 1, dtype=torch.float64)))
        for it in range(self.iters):
            for j in range(N):
                for p in table.get_partitions():                 # GDtask: one pass of SGD per visiting topic
                    w = p.get()
                    if X.shape[0] == 0:
                        continue
                    target = (y == p.id()).double()
                    z = Xb @ w
                    grad = Xb.t() @ (torch.sigmoid(z) - target) / X.shape[0]
                    w -= self.lr * grad
                self.rotate("mlr", f"rotate-{it}-{j}", table, None)
            context.progress()
        self.allgather("mlr", "allgather-final", table)
        self.W = torch.stack([table.get_partition(t).get() for t in range(self.T)])
        pred = (Xb @ self.W.t()).argmax(1) if X.shape[0] else torch.zeros(0, dtype=torch.long)
        # macro-F1 over all workers: confusion counts are summed (EVtask rotates them; the sum is the same)
        cm = torch.zeros(self.T, 3, dtype=torch.float64)
        for t in range(self.T):
            cm[t, 0] = float(((pred == t) & (y == t)).sum()); cm[t, 1] = float(((pred == t) & (y != t)).sum())
            cm[t, 2] = float(((pred != t) & (y == t)).sum())
        tt = Table(1, DoubleArrPlus()); tt.add_partition(Partition(0, cm.flatten()))
        self.allreduce("mlr", "allreduce-eval", tt)
        cm = tt.get_partition(0).get().view(self.T, 3)
        f1 = 2 * cm[:, 0] / (2 * cm[:, 0] + cm[:, 1] + cm[:, 2]).clamp(min=1)
        self.macro_f1 = float(f1.mean())


# ---------------------------------------------------------------------------------------
class LDACVBMapper(CollectiveMapper):
    """conf: ``num_topics``, ``num_terms``, ``num_iterations``, ``alpha``, ``beta``, ``dynamic`` (push/pull variant),
    ``update_alpha`` (default on); input = ``docName term:count term:count ...`` (``contrib/.../lda/README.md``).

    The topic prior is an asymmetric Dirichlet whose vector is re-estimated after every pass from the all-reduced
    statistics sum_d [psi(gamma_dk) - psi(sum_k gamma_dk)] (``contrib/.../lda/LDAMapper.java:345-360,428-447``)."""

    def setup(self, context: Context) -> None:
        c = context.get_configuration()
        self.K, self.V = int(c["num_topics"]), int(c["num_terms"])
        self.iters = int(c.get("num_iterations", 10))
        self.alpha, self.beta = float(c.get("alpha", 0.1)), float(c.get("beta", 0.01))
        self.dynamic = bool(c.get("dynamic", False))
        self.update_alpha = bool(c.get("update_alpha", True))

    @staticmethod
    def update_vector_alpha(alpha: torch.Tensor, num_docs: int, suff: torch.Tensor, *, decay_factor: float = 0.8,
                            max_decay: int = 10, max_iter: int = 1000, tol: float = 1e-6) -> torch.Tensor:
        """Newton iteration on the Dirichlet prior with the diagonal-plus-rank-one Hessian inverted in closed form
        (``LDAMapper.java:476-600``; constants from ``lda/Constants.java:24-31``).  gradient_k = D (psi(sum a) -
        psi(a_k)) + s_k, h_k = -D psi'(a_k), z = D psi'(sum a), c = sum(g/h) / (1/z + sum(1/h)); the step
        (g_k - c) / h_k is shrunk by ``decay_factor`` until every component stays positive."""
        a = alpha.clone().double()
        D = float(num_docs)
        decay = 0
        for _ in range(max_iter + 1):
            asum = a.sum()
            grad = D * (torch.digamma(asum) - torch.digamma(a)) + suff
            if not bool(torch.isfinite(grad).all()):
                raise ArithmeticError("alpha gradient is not finite")
            hess = -D * torch.polygamma(1, a)
            z = D * torch.polygamma(1, asum)
            cc = (grad / hess).sum() / (1.0 / z + (1.0 / hess).sum())
            new = None
            while decay <= max_decay:
                step = (decay_factor ** decay) * (grad - cc) / hess
                if bool((a <= step).any()):
                    decay += 1                                  # would leave the positive orthant: shrink the step
                    continue
                new = a - step
                break
            if new is None:
                break
            done = bool((((new - a) / a).abs() < tol).all())
            a = new
            if done:
                break
        return a

    def map_collective(self, reader: KeyValReader, context: Context) -> None:
        docs = []
        for f in _files(reader):
            with open(f) as fh:
                for line in fh:
                    toks = line.split()
                    if len(toks) < 2:
                        continue
                    ids, cnts = zip(*[(int(a), float(b)) for a, b in (t.split(":") for t in toks[1:])])
                    docs.append((torch.tensor(ids), torch.tensor(cnts, dtype=torch.float64)))
        K, V = self.K, self.V
        g = torch.Generator().manual_seed(7)
        log_phi = torch.log(torch.rand(V, K, generator=g, dtype=torch.float64) + 1.0 / V)
        log_phi -= torch.logsumexp(log_phi, 0, keepdim=True)           # each topic is a distribution over terms
        self.likelihood_history = []
        N = self.get_num_workers()
        alpha = torch.full((K,), self.alpha, dtype=torch.float64)
        num_docs = int(sum(self.comm.all_gather_object(len(docs))))
        for it in range(self.iters):
            new = torch.full((V, K), -float("inf"), dtype=torch.float64)
            ll = 0.0
            suff = torch.zeros(K, dtype=torch.float64)
            for ids, cnt in docs:
                gamma = alpha + float(cnt.sum()) / K
                lp = log_phi[ids]                                       # [terms, K]
                for _ in range(20):                                      # per-document fixed point (digamma updates)
                    log_r = lp + torch.digamma(gamma)
                    log_r = log_r - torch.logsumexp(log_r, 1, keepdim=True)
                    g_new = alpha + (cnt[:, None] * torch.exp(log_r)).sum(0)
                    if float((g_new - gamma).abs().max()) < 1e-4:
                        gamma = g_new
                        break
                    gamma = g_new
                contrib = torch.log(cnt)[:, None] + log_r
                new[ids] = torch.logaddexp(new[ids], contrib)
                theta = gamma / gamma.sum()
                ll += float((cnt * torch.logsumexp(lp + torch.log(theta), 1)).sum())
                suff += torch.digamma(gamma) - torch.digamma(gamma.sum())
            table = Table(0, LogSumExpCombiner())
            if self.dynamic:
                # LDAMapperDyn: log-phi sharded by term block with push / pull
                blocks = 4 * N
                for b in range(blocks):
                    table.add_partition(Partition(b, new[b::blocks].contiguous()))
                glob = Table(1, LogSumExpCombiner())
                self.push("lda", f"push-{it}", table, glob, Partitioner(N))
                local = Table(2, LogSumExpCombiner())
                for b in range(blocks):
                    local.add_partition(Partition(b, torch.full_like(new[b::blocks], -float("inf")).contiguous()))
                self.pull("lda", f"pull-{it}", local, glob, True)
                for b in range(blocks):
                    new[b::blocks] = local.get_partition(b).get()
            else:
                table.add_partition(Partition(0, new))
                self.allreduce("lda", f"allreduce-phi-{it}", table)
                new = table.get_partition(0).get()
            log_phi = torch.logaddexp(new, torch.full_like(new, math.log(self.beta)))
            log_phi = log_phi - torch.logsumexp(log_phi, 0, keepdim=True)
            if self.update_alpha:
                st = Table(4, DoubleArrPlus()); st.add_partition(Partition(0, suff))
                self.allreduce("lda", f"allreduce-alpha-suff-{it}", st)
                alpha = self.update_vector_alpha(alpha, num_docs, st.get_partition(0).get())
            lt = Table(3, DoubleArrPlus()); lt.add_partition(Partition(0, torch.tensor([ll], dtype=torch.float64)))
            self.allreduce("lda", f"allreduce-ll-{it}", lt)
            self.likelihood_history.append(float(lt.get_partition(0).get().item()))
            context.progress()
        self.log_phi = log_phi
        self.alpha_vector = alpha


# ---------------------------------------------------------------------------------------
class NNMapper(CollectiveMapper):
    """conf: ``sizes`` (hidden layer widths), ``num_classes``, ``epochs``, ``batch_size``, ``sync_iter``, ``lr``;
    input = CSV rows ``features..., label``."""

    def setup(self, context: Context) -> None:
        self.c = context.get_configuration()

    def map_collective(self, reader: KeyValReader, context: Context) -> None:
        c = self.c
        mats = [F.read_dense_csv(f) for f in _files(reader)]
        m = np.concatenate(mats) if mats else np.empty((0, int(c["num_features"]) + 1))
        X = torch.from_numpy(m[:, :-1].copy()); y = torch.from_numpy(m[:, -1].copy()).long()
        sizes = [X.shape[1]] + list(c.get("sizes", [16])) + [int(c["num_classes"])]
        net = MLP(sizes, seed=0, dtype=torch.float64)
        N = self.get_num_workers()
        bs, sync, lr = int(c.get("batch_size", 32)), int(c.get("sync_iter", 5)), float(c.get("lr", 0.1))
        steps_per_epoch = max(self.comm.all_gather_object((X.shape[0] + bs - 1) // bs))
        g = torch.Generator().manual_seed(self.get_self_id())
        step = 0
        for ep in range(int(c.get("epochs", 10))):
            for b in range(steps_per_epoch):
                if X.shape[0]:
                    idx = torch.randint(0, X.shape[0], (bs,), generator=g)
                    _, gW, gb = net.loss_and_grads(X[idx], y[idx])
                    for t, gr in zip(net.W + net.b, gW + gb):
                        t -= lr * gr / bs
                step += 1
                if step % sync == 0:                              # average the weights (NNMapper.java:143-179)
                    flat = torch.cat([t.flatten() for t in net.W + net.b])
                    tab = Table(0, DoubleArrPlus()); tab.add_partition(Partition(0, flat))
                    self.allreduce("nn", f"allreduce-weights-{step}", tab)
                    flat = tab.get_partition(0).get() / N
                    off = 0
                    for t in net.W + net.b:
                        t.copy_(flat[off:off + t.numel()].view_as(t)); off += t.numel()
            context.progress()
        self.net = net
        acc = torch.tensor([float((net.predict(X) == y).sum()) if X.shape[0] else 0.0, float(X.shape[0])], dtype=torch.float64)
        tt = Table(1, DoubleArrPlus()); tt.add_partition(Partition(0, acc))
        self.allreduce("nn", "allreduce-acc", tt)
        a = tt.get_partition(0).get()
        self.train_accuracy = float(a[0] / max(float(a[1]), 1.0))


# ---------------------------------------------------------------------------------------
class RandomForestMapper(CollectiveMapper):
    """conf: ``num_trees`` (per worker), ``num_classes``, ``test_file``; input = CSV rows ``features..., label``;
    every worker votes on the shared test file and the per-class vote counts are reduced to the master."""

    def setup(self, context: Context) -> None:
        self.c = context.get_configuration()

    def map_collective(self, reader: KeyValReader, context: Context) -> None:
        c = self.c
        mats = [F.read_dense_csv(f) for f in _files(reader)]
        m = np.concatenate(mats) if mats else None
        test = F.read_dense_csv(c["test_file"])
        Xt, yt = torch.from_numpy(test[:, :-1].copy()), torch.from_numpy(test[:, -1].copy()).long()
        nc = int(c["num_classes"])
        votes = torch.zeros(Xt.shape[0], nc, dtype=torch.float64)
        if m is not None and m.shape[0]:
            X, y = torch.from_numpy(m[:, :-1].copy()), torch.from_numpy(m[:, -1].copy()).long()
            forest = DC.DecisionForest("classification", n_trees=int(c.get("num_trees", 5)), max_depth=int(c.get("max_depth", 8)),
                                       seed=self.get_self_id()).fit(X, y)
            for tree in forest.trees:
                pr = tree._predict_value(Xt)
                lab = pr.argmax(1)
                votes[torch.arange(Xt.shape[0]), lab] += 1
        t = Table(0, DoubleArrPlus()); t.add_partition(Partition(0, votes.flatten()))
        self.reduce("rf", "reduce-votes", t, self.get_master_id())
        self.test_accuracy = None
        if self.is_master():
            v = t.get_partition(0).get().view(-1, nc)
            pred = v.argmax(1)
            self.test_accuracy = float((pred == yt).double().mean())
            for i, p in enumerate(pred.tolist()):
                context.write(i, p)
