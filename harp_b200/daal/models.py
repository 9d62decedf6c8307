"""Harp-DAAL algorithms, part 3: K-means (DAAL variants), EM for Gaussian mixtures, implicit ALS
(distributed 4-step and batch), association rules (Apriori), neural-network training with gradient
allreduce, and the iterative optimisation solvers (``ml/daal/src/main/java/edu/iu/daal_*``).
"""
from __future__ import annotations

import itertools
import math
from typing import Dict, FrozenSet, NamedTuple, Optional, Sequence

import torch
import torch.distributed as dist

from ..ops.kmeans import kmeans_step
from .linalg import _allreduce_


# ---------------------------------------------------------------------------------------
# K-means (daal_kmeans/regroupallgather with 4 selectable comm patterns, daal_kmeans/allreducecsr)
# ---------------------------------------------------------------------------------------
def kmeans_init(X: torch.Tensor, k: int, method: str = "random", seed: int = 0, comm=None) -> torch.Tensor:
    """DAAL ``kmeans.init`` for dense and CSR input (``daal_kmeans/allreducecsr/KMeansDaalCollectiveMapper.java``:
    ``init.Method.randomCSR`` / ``deterministicCSR``; dense twins ``randomDense`` / ``deterministicDense`` /
    ``plusPlusDense``): ``deterministic`` = the first k rows of the global data set, ``random`` = k distinct rows drawn
    with a seed shared by all ranks, ``plusPlus`` = D^2 seeding.  Row-distributed: every rank contributes the rows it
    holds (global row order = rank order) and all ranks return the same [k, d] dense float64 centroids."""
    N = 1 if comm is None else comm.world
    me = 0 if comm is None else comm.rank
    n_loc = X.shape[0]
    counts = [n_loc] if N == 1 else comm.all_gather_object(n_loc)
    start = sum(counts[:me])
    total = sum(counts)
    assert total >= k, "fewer rows than centroids"

    def rows(global_ids: torch.Tensor) -> torch.Tensor:
        loc = global_ids - start
        mine = (loc >= 0) & (loc < n_loc)
        out = torch.zeros(global_ids.numel(), X.shape[1], dtype=torch.float64)
        if bool(mine.any()):
            sel = loc[mine]
            blk = X.index_select(0, sel.to(X.device)) if X.layout == torch.strided else \
                torch.stack([X[int(i)].to_dense() for i in sel])          # CSR: only the chosen rows are densified
            out[mine] = blk.double().cpu()
        if N > 1:
            _allreduce_(out, comm)
        return out
    g = torch.Generator().manual_seed(seed)
    if method == "deterministic":
        return rows(torch.arange(k))
    if method == "random":
        return rows(torch.randperm(total, generator=g)[:k].sort().values)
    if method == "plusPlus":
        C = rows(torch.randint(0, total, (1,), generator=g))
        for _ in range(1, k):
            d2 = _sq_dist_to_nearest(X, C)                          # local rows
            tot = d2.sum().double().reshape(1).cpu()
            if N > 1:
                _allreduce_(tot, comm)
            u = float(torch.rand(1, generator=g)) * float(tot)
            # the rank whose cumulative share contains u picks the row
            shares = [float(d2.sum())] if N == 1 else comm.all_gather_object(float(d2.sum()))
            acc, pick = 0.0, torch.full((1,), -1, dtype=torch.long)
            for r_, sh in enumerate(shares):
                if u < acc + sh or r_ == N - 1:
                    if r_ == me:
                        idx = int(torch.searchsorted(torch.cumsum(d2.double().cpu(), 0), torch.tensor(u - acc)).clamp(max=n_loc - 1))
                        pick[0] = start + idx
                    break
                acc += sh
            if N > 1:
                pk = pick.double()
                _allreduce_(pk, comm, "max")
                pick = pk.long()
            C = torch.cat([C, rows(pick)], 0)
        return C
    raise ValueError(method)


def _sq_dist_to_nearest(X: torch.Tensor, C: torch.Tensor) -> torch.Tensor:
    """min_j ||x - c_j||^2 for dense or CSR rows (CSR: x.c through a sparse product, nothing densified)."""
    Cd = C.to(X.device, torch.float64)
    if X.layout == torch.strided:
        return torch.cdist(X.double(), Cd).pow(2).min(1).values
    Xd = X.to(torch.float64)
    xn = torch.sparse.sum(Xd * Xd, 1).to_dense() if Xd.layout != torch.sparse_csr else \
        torch.zeros(X.shape[0], dtype=torch.float64, device=X.device).index_add_(
            0, torch.repeat_interleave(torch.arange(X.shape[0], device=X.device), Xd.crow_indices()[1:] - Xd.crow_indices()[:-1]),
            Xd.values() ** 2)
    cross = torch.sparse.mm(Xd, Cd.t())
    return (xn[:, None] - 2 * cross + (Cd * Cd).sum(1)[None, :]).clamp(min=0).min(1).values


def kmeans_csr_step(X: torch.Tensor, C: torch.Tensor):
    """One Lloyd pass over CSR rows without densifying them (``daal_kmeans/allreducecsr``): returns
    ``(sums [k, d], counts [k], objective)`` in float64; x.c and the per-cluster sums are sparse products."""
    Xd = X.to(torch.float64)
    Cd = C.to(X.device, torch.float64)
    n, k = X.shape[0], C.shape[0]
    crow = Xd.crow_indices()
    row_of = torch.repeat_interleave(torch.arange(n, device=X.device), crow[1:] - crow[:-1])
    xn = torch.zeros(n, dtype=torch.float64, device=X.device).index_add_(0, row_of, Xd.values() ** 2)
    d2 = (xn[:, None] - 2 * torch.sparse.mm(Xd, Cd.t()) + (Cd * Cd).sum(1)[None, :]).clamp(min=0)
    best_d, best = d2.min(1)
    onehot = torch.sparse_coo_tensor(torch.stack([best, torch.arange(n, device=X.device)]), torch.ones(n, dtype=torch.float64, device=X.device), (k, n))
    sums = torch.sparse.mm(onehot, Xd.to_sparse_coo()).to_dense()
    counts = torch.bincount(best, minlength=k).double()
    return sums, counts, best_d.sum()


def kmeans_daal(X: torch.Tensor, centroids: torch.Tensor, n_iter: int, comm=None, comm_mode: str = "allreduce"):
    """DAAL ``kmeans.DistributedStep1Local`` partial results (partialSums, nObservations) per iteration, combined
    with one of the reference's four patterns (``KMeansDaalCollectiveMapper.java:435-532``).  With tensors all four
    move the same fixed-shape ``[K, d+1]`` block; they differ in who divides:
    ``allreduce`` (everyone), ``regroup_allgather`` (owners of row ranges), ``broadcastreduce`` (rank 0),
    ``push_pull`` (owners, then pulled)."""
    sparse = X.layout == torch.sparse_csr
    if X.layout != torch.strided and not sparse:
        X = X.to_dense()
    C = centroids.clone().to(X.device, torch.float64 if (sparse or X.device.type != "cuda") else X.dtype)
    if X.device.type != "cuda" and not sparse:
        X = X.double()
    K, d = C.shape
    N = 1 if comm is None else comm.world
    me = 0 if comm is None else comm.rank
    hist = []
    for _ in range(n_iter):
        if sparse:                         # allreducecsr: the CSR rows are never densified
            sums_, counts_, obj_ = kmeans_csr_step(X, C)
            part = torch.cat([sums_, counts_[:, None]], 1)
            obj = obj_.reshape(1).clone()
        else:
            res = kmeans_step(X, C, want_obj=True, faithful=True)
            part = torch.cat([res.sums.double(), res.counts.double()[:, None]], 1)     # [K, d+1]
            obj = res.obj.clone()
        if N > 1:
            if comm_mode in ("allreduce",):
                _allreduce_(part, comm)
            elif comm_mode in ("regroup_allgather", "push_pull"):
                work = part.to(comm.device)
                rows = (K + N - 1) // N
                padded = torch.zeros(rows * N, d + 1, dtype=work.dtype, device=work.device)
                padded[:K] = work
                mine = torch.zeros(rows, d + 1, dtype=work.dtype, device=work.device)
                dist.reduce_scatter_tensor(mine, padded, group=comm.group)            # regroup / push
                nz = mine[:, d:] > 0
                mine[:, :d] = torch.where(nz, mine[:, :d] / torch.where(nz, mine[:, d:], torch.ones_like(mine[:, d:])), mine[:, :d])
                dist.all_gather_into_tensor(padded, mine, group=comm.group)           # allgather / pull
                part = padded[:K].to(part.device)
                part[:, :d] = part[:, :d] * torch.where(part[:, d:] > 0, part[:, d:], torch.ones_like(part[:, d:]))
            elif comm_mode == "broadcastreduce":
                work = part.to(comm.device)
                dist.reduce(work, 0, group=comm.group)
                dist.broadcast(work, 0, group=comm.group)
                part = work.to(part.device)
            else:
                raise ValueError(comm_mode)
            _allreduce_(obj, comm)
        cnt = part[:, d:]
        newC = torch.where(cnt > 0, part[:, :d] / torch.where(cnt > 0, cnt, torch.ones_like(cnt)), torch.zeros_like(part[:, :d]))
        C = newC.to(C.dtype)
        hist.append(float(obj.item()))
    return C, hist


# ---------------------------------------------------------------------------------------
# EM for Gaussian mixtures (daal_em)
# ---------------------------------------------------------------------------------------
class GMM(NamedTuple):
    weights: torch.Tensor
    means: torch.Tensor
    covariances: torch.Tensor
    log_likelihood: float
    n_iter: int


def em_gmm(X: torch.Tensor, n_components: int, n_iter: int = 100, tol: float = 1e-6, seed: int = 0, reg: float = 1e-6) -> GMM:
    Xd = X.double()
    n, d = Xd.shape
    g = torch.Generator(device="cpu").manual_seed(seed)
    idx = torch.randperm(n, generator=g)[:n_components].to(X.device)
    mu = Xd[idx].clone()
    cov = torch.cov(Xd.t()).expand(n_components, d, d).clone() + reg * torch.eye(d, dtype=Xd.dtype, device=X.device)
    w = torch.full((n_components,), 1.0 / n_components, dtype=Xd.dtype, device=X.device)
    prev = -float("inf")
    ll = prev
    it = 0
    for it in range(1, n_iter + 1):
        Lc = torch.linalg.cholesky(cov)
        diff = Xd[None, :, :] - mu[:, None, :]                                   # [k, n, d]
        sol = torch.linalg.solve_triangular(Lc, diff.transpose(1, 2), upper=False)   # [k, d, n]
        maha = (sol * sol).sum(1)                                                # [k, n]
        logdet = 2 * torch.log(torch.diagonal(Lc, dim1=1, dim2=2)).sum(1)
        logp = -0.5 * (maha + logdet[:, None] + d * math.log(2 * math.pi)) + torch.log(w)[:, None]
        lse = torch.logsumexp(logp, 0)
        ll = float(lse.sum())
        r = torch.exp(logp - lse)                                                # responsibilities [k, n]
        nk = r.sum(1)
        w = nk / n
        mu = (r @ Xd) / nk[:, None]
        diff = Xd[None] - mu[:, None]
        cov = torch.einsum("kn,knd,kne->kde", r, diff, diff) / nk[:, None, None] + reg * torch.eye(d, dtype=Xd.dtype, device=X.device)
        if abs(ll - prev) < tol * max(1.0, abs(ll)):
            break
        prev = ll
    return GMM(w, mu, cov, ll, it)


# ---------------------------------------------------------------------------------------
# implicit ALS (daal_als 4-step distributed, daal_als_batch)
# ---------------------------------------------------------------------------------------
def _als_solve_side(ptr, idx, val, other: torch.Tensor, n_rows: int, lam: float, alpha: float) -> torch.Tensor:
    """One half-iteration of implicit-feedback ALS (Hu–Koren–Volinsky): for every row u,
    ``x_u = (YᵀY + Yᵀ(Cu − I)Y + λI)^{-1} Yᵀ Cu p(u)`` with confidence ``c = 1 + α r``."""
    f = other.shape[1]
    YtY = other.t() @ other
    out = torch.zeros(n_rows, f, dtype=other.dtype, device=other.device)
    eye = torch.eye(f, dtype=other.dtype, device=other.device)
    ptr_l = ptr.tolist()
    for u in range(n_rows):
        a, b = ptr_l[u], ptr_l[u + 1]
        if a == b:
            continue
        Yu = other[idx[a:b].long()]
        cu = 1.0 + alpha * val[a:b].to(other.dtype)
        A = YtY + (Yu * (cu - 1.0)[:, None]).t() @ Yu + lam * eye
        rhs = (Yu * cu[:, None]).sum(0)
        out[u] = torch.linalg.solve(A, rhs)
    return out


def _csr_from_coo(rows, cols, vals, n_rows):
    order = torch.argsort(rows, stable=True)
    rows, cols, vals = rows[order], cols[order], vals[order]
    ptr = torch.zeros(n_rows + 1, dtype=torch.long, device=rows.device)
    ptr[1:] = torch.cumsum(torch.bincount(rows, minlength=n_rows), 0)
    return ptr, cols, vals


def als_implicit(rows: torch.Tensor, cols: torch.Tensor, vals: torch.Tensor, n_users: int, n_items: int, n_factors: int = 10,
                 n_iter: int = 10, lam: float = 0.01, alpha: float = 40.0, seed: int = 0, comm=None):
    """``daal_als_batch`` (comm=None) and the distributed 4-step protocol of ``daal_als``
    (``ALSDaalCollectiveMapper.java:354-492``): users and items are block-partitioned by ``id % N``; every
    half-iteration the *other* side's factors are allgathered (step 1-3: partial cross-products and factor
    exchange) and each rank solves for the rows it owns (step 4)."""
    dev = rows.device
    N = 1 if comm is None else comm.world
    me = 0 if comm is None else comm.rank
    g = torch.Generator(device="cpu").manual_seed(seed)
    X = (torch.rand(n_users, n_factors, generator=g, dtype=torch.float64) * 0.1).to(dev)
    Y = (torch.rand(n_items, n_factors, generator=g, dtype=torch.float64) * 0.1).to(dev)
    rows, cols = rows.long(), cols.long()
    if N > 1:
        # each rank needs, for the users/items it owns, ALL of their ratings: exchange by owner
        def exchange(keys, a, b, v):
            owner = keys % N
            send = [torch.stack([a[owner == r].double(), b[owner == r].double(), v[owner == r].double()], 1) for r in range(N)]
            gathered = comm.all_gather_object([s.cpu() for s in send])
            return torch.cat([gathered[src][me] for src in range(N)], 0).to(dev)
        by_user = exchange(rows, rows, cols, vals)
        by_item = exchange(cols, rows, cols, vals)
    else:
        by_user = torch.stack([rows.double(), cols.double(), vals.double()], 1)
        by_item = by_user
    up, ui, uv = _csr_from_coo(by_user[:, 0].long(), by_user[:, 1].long(), by_user[:, 2], n_users)
    ip, ii, iv = _csr_from_coo(by_item[:, 1].long(), by_item[:, 0].long(), by_item[:, 2], n_items)
    own_u = torch.arange(me, n_users, N, device=dev)
    own_i = torch.arange(me, n_items, N, device=dev)

    def share(M, own):
        if N == 1:
            return M
        part = torch.zeros_like(M)
        part[own] = M[own]
        _allreduce_(part, comm)
        return part
    from ..ops import linalg as KL
    if dev.type == "cuda" and KL.als_pad(n_factors):
        # K22: steps 1-2 = cross-product of the owned factor rows, summed over ranks; step 3 = the factor rows
        # travel; step 4 = one warp per owned row builds, factors and solves its normal matrix (fp32)
        F = KL.als_pad(n_factors)
        Xf = torch.zeros(n_users, F, dtype=torch.float32, device=dev); Xf[:, :n_factors] = X.float()
        Yf = torch.zeros(n_items, F, dtype=torch.float32, device=dev); Yf[:, :n_factors] = Y.float()
        su = KL.ALSSide(up, ui, uv, own_u if N > 1 else None)
        si = KL.ALSSide(ip, ii, iv, own_i if N > 1 else None)

        def half(side, other, own_other, out):
            blk = other[own_other] if N > 1 else other
            G = blk.t() @ blk
            if N > 1:
                _allreduce_(G, comm)
                out.zero_()
            KL.als_solve(side, other, out, n_factors, lam, alpha, G)
            if N > 1:
                _allreduce_(out, comm)
        for _ in range(n_iter):
            half(su, Yf, own_i, Xf)
            half(si, Xf, own_u, Yf)
        bad = int(su.info.item()) + int(si.info.item())
        if bad:
            raise RuntimeError(f"als_implicit: {bad} normal matrices were not positive definite")
        return Xf[:, :n_factors].double(), Yf[:, :n_factors].double()
    for _ in range(n_iter):
        Xn = _als_solve_side(up, ui, uv, Y, n_users, lam, alpha)
        X = share(Xn, own_u) if N > 1 else Xn
        Yn = _als_solve_side(ip, ii, iv, X, n_items, lam, alpha)
        Y = share(Yn, own_i) if N > 1 else Yn
    return X, Y


# ---------------------------------------------------------------------------------------
# association rules (daal_ar: Apriori)
# ---------------------------------------------------------------------------------------
def apriori(transactions: Sequence[Sequence[int]], min_support: float = 0.1, min_confidence: float = 0.6):
    """Returns (large itemsets with supports, rules ``(antecedent, consequent, confidence)``)."""
    n = len(transactions)
    tsets = [frozenset(t) for t in transactions]
    items = sorted({i for t in tsets for i in t})
    # bitmap [n, items] lets supports be counted with tensor ops
    col = {it: j for j, it in enumerate(items)}
    M = torch.zeros(n, len(items), dtype=torch.bool)
    for r, t in enumerate(tsets):
        for it in t:
            M[r, col[it]] = True
    support: Dict[FrozenSet[int], float] = {}
    level = []
    for it in items:
        s = float(M[:, col[it]].double().mean())
        if s >= min_support:
            support[frozenset([it])] = s
            level.append(frozenset([it]))
    k = 2
    while level:
        cands = set()
        for a, b in itertools.combinations(level, 2):
            u = a | b
            if len(u) == k and all(frozenset(s) in support for s in itertools.combinations(u, k - 1)):
                cands.add(u)
        level = []
        for cnd in cands:
            s = float(M[:, [col[i] for i in cnd]].all(1).double().mean())
            if s >= min_support:
                support[cnd] = s
                level.append(cnd)
        k += 1
    rules = []
    for iset, s in support.items():
        if len(iset) < 2:
            continue
        for r in range(1, len(iset)):
            for ante in itertools.combinations(sorted(iset), r):
                a = frozenset(ante)
                conf = s / support[a]
                if conf >= min_confidence:
                    rules.append((tuple(sorted(a)), tuple(sorted(iset - a)), conf))
    return support, sorted(rules, key=lambda x: (-x[2], x[0], x[1]))


# ---------------------------------------------------------------------------------------
# optimisation solvers (daal_optimization_solvers: SGD, mini-batch SGD, momentum, AdaGrad, LBFGS on
# the MSE objective) — explicit tensor implementations, no torch.optim
# ---------------------------------------------------------------------------------------
class MSEObjective:
    """``optimization_solver.mse``: f(θ) = 1/(2n) Σ (θ0 + xᵀθ − y)²."""

    def __init__(self, X: torch.Tensor, y: torch.Tensor):
        self.X = torch.cat([torch.ones(X.shape[0], 1, dtype=torch.float64, device=X.device), X.double()], 1)
        self.y = y.double().flatten()
        self.n = self.X.shape[0]

    def value(self, theta, idx=None):
        X, y = (self.X, self.y) if idx is None else (self.X[idx], self.y[idx])
        r = X @ theta - y
        return float((r * r).sum() / (2 * X.shape[0]))

    def gradient(self, theta, idx=None):
        X, y = (self.X, self.y) if idx is None else (self.X[idx], self.y[idx])
        return X.t() @ (X @ theta - y) / X.shape[0]

    def dim(self):
        return self.X.shape[1]


def solver_sgd(obj: MSEObjective, n_iter: int = 1000, lr: float = 0.01, batch_size: int = 1, momentum: float = 0.0,
               seed: int = 0, theta0: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Plain (batch_size=1), mini-batch and momentum SGD (``SGDDenseBatch``, ``SGDMiniDenseBatch``,
    ``SGDMomentDenseBatch``)."""
    g = torch.Generator(device="cpu").manual_seed(seed)
    theta = torch.zeros(obj.dim(), dtype=torch.float64, device=obj.X.device) if theta0 is None else theta0.clone()
    vel = torch.zeros_like(theta)
    for _ in range(n_iter):
        idx = torch.randint(0, obj.n, (batch_size,), generator=g).to(obj.X.device)
        grad = obj.gradient(theta, idx)
        vel = momentum * vel - lr * grad
        theta = theta + vel
    return theta


def solver_adagrad(obj: MSEObjective, n_iter: int = 1000, lr: float = 0.5, batch_size: int = 16, eps: float = 1e-8,
                   seed: int = 0) -> torch.Tensor:
    g = torch.Generator(device="cpu").manual_seed(seed)
    theta = torch.zeros(obj.dim(), dtype=torch.float64, device=obj.X.device)
    acc = torch.zeros_like(theta)
    for _ in range(n_iter):
        idx = torch.randint(0, obj.n, (batch_size,), generator=g).to(obj.X.device)
        grad = obj.gradient(theta, idx)
        acc += grad * grad
        theta = theta - lr * grad / (acc.sqrt() + eps)
    return theta


def solver_lbfgs(obj: MSEObjective, n_iter: int = 100, m: int = 10, tol: float = 1e-10) -> torch.Tensor:
    """Limited-memory BFGS with two-loop recursion and Armijo backtracking (``LBFGSDenseBatch``)."""
    theta = torch.zeros(obj.dim(), dtype=torch.float64, device=obj.X.device)
    S, Yh = [], []
    grad = obj.gradient(theta)
    for _ in range(n_iter):
        q = grad.clone()
        alphas = []
        for s, y in zip(reversed(S), reversed(Yh)):
            a = float(s @ q) / float(y @ s)
            alphas.append(a)
            q -= a * y
        if S:
            q *= float(S[-1] @ Yh[-1]) / float(Yh[-1] @ Yh[-1])
        for (s, y), a in zip(zip(S, Yh), reversed(alphas)):
            b = float(y @ q) / float(y @ s)
            q += (a - b) * s
        d = -q
        f0, slope = obj.value(theta), float(grad @ d)
        step = 1.0
        while obj.value(theta + step * d) > f0 + 1e-4 * step * slope and step > 1e-12:
            step *= 0.5
        new_theta = theta + step * d
        new_grad = obj.gradient(new_theta)
        s, y = new_theta - theta, new_grad - grad
        if float(y @ s) > 1e-14:
            S.append(s); Yh.append(y)
            if len(S) > m:
                S.pop(0); Yh.pop(0)
        theta, grad = new_theta, new_grad
        if float(grad.norm()) < tol:
            break
    return theta


# ---------------------------------------------------------------------------------------
# neural network (daal_nn distributed: local gradients, gather+bcast per batch → here allreduce)
# ---------------------------------------------------------------------------------------
class MLP:
    """Fully connected network with explicit forward/backward (no autograd), tanh hidden units, softmax
    output; the gradient of every mini-batch is allreduced so all ranks apply the same update
    (``NNDaalCollectiveMapper.java``; contrib NN averages weights instead, see ``apps/contrib.py``)."""

    def __init__(self, sizes: Sequence[int], seed: int = 0, device="cpu", dtype=torch.float32):
        g = torch.Generator(device="cpu").manual_seed(seed)
        self.W = [(torch.randn(a, b, generator=g, dtype=torch.float64) * math.sqrt(1.0 / a)).to(device, dtype)
                  for a, b in zip(sizes[:-1], sizes[1:])]
        self.b = [torch.zeros(b, device=device, dtype=dtype) for b in sizes[1:]]

    def forward(self, X):
        acts = [X]
        for i, (W, b) in enumerate(zip(self.W, self.b)):
            z = acts[-1] @ W + b
            acts.append(torch.tanh(z) if i < len(self.W) - 1 else z)
        return acts

    def loss_and_grads(self, X, y):
        acts = self.forward(X)
        logits = acts[-1]
        logp = logits - torch.logsumexp(logits, 1, keepdim=True)
        n = X.shape[0]
        loss = -logp[torch.arange(n, device=X.device), y.long()].sum()
        delta = torch.exp(logp)
        delta[torch.arange(n, device=X.device), y.long()] -= 1.0
        gW, gb = [], []
        for i in range(len(self.W) - 1, -1, -1):
            gW.append(acts[i].t() @ delta)
            gb.append(delta.sum(0))
            if i > 0:
                delta = (delta @ self.W[i].t()) * (1 - acts[i] ** 2)
        return loss, gW[::-1], gb[::-1]

    def train(self, X, y, epochs: int = 10, batch_size: int = 64, lr: float = 0.1, comm=None, seed: int = 0):
        n = X.shape[0]
        g = torch.Generator(device="cpu").manual_seed(seed)
        hist = []
        nb = (n + batch_size - 1) // batch_size
        if comm is not None and comm.world > 1:
            nb_t = torch.tensor([float(nb)], dtype=torch.float64, device=X.device)
            _allreduce_(nb_t, comm, "max")
            nb = int(nb_t.item())
        for _ in range(epochs):
            perm = torch.randperm(n, generator=g).to(X.device)
            tot = torch.zeros(2, dtype=torch.float64, device=X.device)
            for bi in range(nb):
                idx = perm[(bi * batch_size) % max(n, 1):][:batch_size]
                loss, gW, gb = self.loss_and_grads(X[idx], y[idx])
                flat = torch.cat([t.flatten() for t in gW + gb] + [torch.tensor([float(idx.numel())], device=X.device, dtype=gW[0].dtype)])
                _allreduce_(flat, comm)
                cnt = float(flat[-1])
                off = 0
                for t in self.W + self.b:
                    k = t.numel()
                    t -= lr * flat[off:off + k].view_as(t) / cnt
                    off += k
                tot += torch.stack([loss.double(), torch.tensor(float(idx.numel()), dtype=torch.float64, device=X.device)])
            _allreduce_(tot, comm)
            hist.append(float(tot[0] / tot[1]))
        return hist

    def predict(self, X):
        return self.forward(X)[-1].argmax(1)
