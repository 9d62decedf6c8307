"""Harp-DAAL algorithms, part 1: the distributed "partial result" family and the dense linear-algebra
batch algorithms (``ml/daal/src/main/java/edu/iu/daal_*``).

The reference pattern is DAAL ``DistributedStep1Local.compute()`` on every mapper →
``harpdaal_gather`` of serialised partial results to the master → ``DistributedStep2Master``
(``COVDaalCollectiveMapper.java:143-176``).  Here every rank computes its fixed-shape partial on
the GPU (tcgen05 Gram kernel, streaming moments kernel) and the partials are **allreduced**, so all
ranks end with the result and nothing is serialised (SURVEY §2.3 HarpDAALComm row).

Every function takes ``comm`` (a :class:`harp_b200.collective.Comm` or None for a single process).
"""
from __future__ import annotations

import math
from typing import Dict, NamedTuple, Optional, Tuple

import os

import torch
import torch.distributed as dist

from ..ops.gram import fused_gram_for, gram, moments


def _allreduce_(t: torch.Tensor, comm, op: str = "sum") -> torch.Tensor:
    """In-place allreduce of a fixed-shape partial result across the communicator."""
    if comm is None or comm.world == 1:
        return t
    if comm.symm is not None and comm.symm.contains(t) and t.dtype in (torch.float32, torch.float64, torch.int32, torch.int64):
        comm.symm.allreduce_(t, op)
        return t
    rop = {"sum": dist.ReduceOp.SUM, "max": dist.ReduceOp.MAX, "min": dist.ReduceOp.MIN}[op]
    work = t if t.device == comm.device else t.to(comm.device)
    dist.all_reduce(work, op=rop, group=comm.group)
    if work is not t:
        t.copy_(work)
    return t


def _f64(t: torch.Tensor) -> torch.Tensor:
    return t.double()


def _fused_for(X: torch.Tensor, comm):
    return fused_gram_for(comm, X.shape[1], X.device) if X.dtype == torch.float32 and X.shape[0] < (1 << 31) else None


def _gram_allreduced(X: torch.Tensor, comm) -> torch.Tensor:
    """Global ``XᵀX`` in float64.  On CUDA with a tensor-core-sized ``d`` the partial Gram and its allreduce
    are ONE kernel launch (``ops.gram.FusedGram``); otherwise Gram + ``_allreduce_``."""
    fg = _fused_for(X, comm)
    if fg is not None:
        return _f64(fg(X))
    G = _f64(gram(X))
    _allreduce_(G, comm)
    return G


def _gram_and_moments(X: torch.Tensor, comm):
    """(global XᵀX in float64, this rank's column ``Moments``): the fused Gram launch plus the moments kernel, or ONE
    launch for both with ``HB_GRAM_FUSED_STATS=1`` (statistics warps inside the Gram kernel)."""
    fg = _fused_for(X, comm)
    if fg is not None:
        # statistics inside the Gram launch save the second HBM pass but not time: the kernel is power-capped and the
        # extra work lowers its clocks by as much as the separate pass costs (profiles/EXPERIMENTS.md), so opt-in
        if os.environ.get("HB_GRAM_FUSED_STATS", "0") == "1":
            G, m = fg(X, want_moments=True)
            return _f64(G), m
        return _f64(fg(X)), moments(X)
    G = _f64(gram(X))
    _allreduce_(G, comm)
    return G, moments(X)


# ---------------------------------------------------------------------------------------
# low-order moments (daal_mom/{dense,csr}distri)
# ---------------------------------------------------------------------------------------
class MomentsResult(NamedTuple):
    n: int
    minimum: torch.Tensor
    maximum: torch.Tensor
    sum: torch.Tensor
    sum_squares: torch.Tensor
    sum_squares_centered: torch.Tensor
    mean: torch.Tensor
    second_order_raw_moment: torch.Tensor
    variance: torch.Tensor
    standard_deviation: torch.Tensor
    variation: torch.Tensor


def low_order_moments(X: torch.Tensor, comm=None) -> MomentsResult:
    """DAAL ``low_order_moments`` results (``MOMDaalCollectiveMapper.java``); sparse input is accepted as a
    torch sparse CSR tensor (``daal_mom/csrdistri``) and densified block-wise."""
    if X.layout != torch.strided:
        X = X.to_dense()
    m = moments(X)
    n = torch.tensor([float(m.n)], dtype=torch.float64, device=X.device)
    s, ss = m.sum.clone(), m.sumsq.clone()
    mn, mx = m.min.double().clone(), m.max.double().clone()
    _allreduce_(n, comm); _allreduce_(s, comm); _allreduce_(ss, comm)
    _allreduce_(mn, comm, "min"); _allreduce_(mx, comm, "max")
    N = float(n.item())
    mean = s / N
    ssc = ss - N * mean * mean
    var = ssc / max(N - 1.0, 1.0)
    sd = var.clamp(min=0).sqrt()
    return MomentsResult(int(N), mn, mx, s, ss, ssc, mean, ss / N, var, sd, sd / mean)


# ---------------------------------------------------------------------------------------
# covariance / PCA (daal_cov, daal_pca)
# ---------------------------------------------------------------------------------------
class CovResult(NamedTuple):
    covariance: torch.Tensor
    mean: torch.Tensor
    n: int
    correlation: torch.Tensor


def _cross_product_partials(X: torch.Tensor, comm) -> Tuple[torch.Tensor, torch.Tensor, float]:
    """(XᵀX, Σx, n) allreduced — DAAL covariance partial results (nObservations, sum, crossProduct)."""
    if X.layout != torch.strided:
        X = X.to_dense()
    G, m = _gram_and_moments(X, comm)
    s = m.sum.clone()
    n = torch.tensor([float(X.shape[0])], dtype=torch.float64, device=X.device)
    _allreduce_(s, comm); _allreduce_(n, comm)
    return G, s, float(n.item())


def covariance(X: torch.Tensor, comm=None, center: Optional[bool] = None) -> CovResult:
    """``daal_cov/densedistri`` and ``csrdistri`` (``COVDaalCollectiveMapper.java:143-176``).

    ``G - N mean meanᵀ`` loses everything to cancellation once the mean is large against the spread (and TF32
    products of the raw values have already dropped most of the spread by then), so such data takes a two-pass
    route: exact float64 column sums first (one streaming kernel), then the tensor-core Gram of the CENTRED rows,
    whose partial results are allreduced like the plain ones.  ``center=None`` decides from the first pass
    (any column with |mean| > 4 sd); True / False force either route."""
    if X.layout != torch.strided:
        X = X.to_dense()
    if center is not False and X.is_floating_point():
        from ..ops.gram import centered_gram
        G1 = None
        guess = center
        if center is None:
            ms = moments(X[:1 << 16])                  # a 64 K-row prefix predicts the route; the full moments confirm it
            mu = ms.sum / max(ms.n, 1)
            flag = (mu.abs() > 4.0 * (ms.sumsq / max(ms.n, 1) - mu * mu).clamp(min=0).sqrt()).any().double().reshape(1)
            _allreduce_(flag, comm, "max")            # every rank must take the same route (the routes differ in collectives)
            guess = bool(flag.item() > 0)
        if center is None and not guess:
            G1, mom = _gram_and_moments(X, comm)      # usually the answer, always the decision
        else:
            mom = moments(X)
        st = torch.stack([mom.sum.double(), mom.sumsq.double(),
                          torch.full_like(mom.sum.double(), float(X.shape[0]))])
        _allreduce_(st, comm)
        N = float(st[2, 0].item())
        mean = st[0] / N
        var = (st[1] / N - mean * mean).clamp(min=0)
        if center is None:
            center = bool((mean.abs() > 4.0 * var.sqrt()).any().item())
        if center:
            m32 = mean.to(torch.float32)
            Gc = _f64(centered_gram(X if X.dtype == torch.float32 or X.device.type != "cuda" else X.float(), m32))
            _allreduce_(Gc, comm)
            delta = mean - m32.double()                      # the shift was the float32-rounded mean
            cov = (Gc - N * torch.outer(delta, delta)) / max(N - 1.0, 1.0)
            sd = cov.diag().clamp(min=1e-300).sqrt()
            return CovResult(cov, mean, int(N), cov / torch.outer(sd, sd))
        if G1 is not None:
            cov = (G1 - N * torch.outer(mean, mean)) / max(N - 1.0, 1.0)
            sd = cov.diag().clamp(min=1e-300).sqrt()
            return CovResult(cov, mean, int(N), cov / torch.outer(sd, sd))
    G, s, N = _cross_product_partials(X, comm)
    mean = s / N
    cov = (G - N * torch.outer(mean, mean)) / max(N - 1.0, 1.0)
    sd = cov.diag().clamp(min=1e-300).sqrt()
    return CovResult(cov, mean, int(N), cov / torch.outer(sd, sd))


class PCAResult(NamedTuple):
    eigenvalues: torch.Tensor      # descending
    eigenvectors: torch.Tensor     # rows = components (DAAL layout)
    mean: torch.Tensor


def pca_correlation(X: torch.Tensor, comm=None) -> PCAResult:
    """``daal_pca/cordensedistr`` / ``corcsrdistr``: eigen-decomposition of the correlation matrix
    (``PCADaalCollectiveMapper.java:114-161``)."""
    c = covariance(X, comm)
    w, v = torch.linalg.eigh(c.correlation)
    idx = torch.argsort(w, descending=True)
    vec = v[:, idx].t().contiguous()
    # deterministic sign: largest-magnitude entry of each component positive
    sgn = torch.sign(vec.gather(1, vec.abs().argmax(1, keepdim=True)))
    return PCAResult(w[idx], vec * sgn, c.mean)


def tsqr_r(X: torch.Tensor, comm=None) -> torch.Tensor:
    """R factor of the row-distributed matrix by communication-avoiding TSQR: local QR, gather the small
    R blocks, QR of the stack (DAAL qr/svd step1Local → step2Master)."""
    Xd = X.to_dense() if X.layout != torch.strided else X
    from ..ops import linalg as KL
    if KL.tsqr_supported(Xd):
        R = KL.tsqr_r(Xd).double()           # K21: one streaming Householder pass over the local rows
    else:
        R = torch.linalg.qr(Xd.double(), mode="r").R
    if comm is None or comm.world == 1:
        return _fix_r_sign(R)
    blocks = [torch.zeros_like(R) for _ in range(comm.world)]
    work = R if R.device == comm.device else R.to(comm.device)
    blocks = [torch.zeros_like(work) for _ in range(comm.world)]
    dist.all_gather(blocks, work.contiguous(), group=comm.group)
    R2 = torch.linalg.qr(torch.cat(blocks, 0), mode="r").R
    return _fix_r_sign(R2.to(R.device))


def _fix_r_sign(R: torch.Tensor) -> torch.Tensor:
    s = torch.sign(R.diag())
    s[s == 0] = 1
    return R * s[:, None]


def pca_svd(X: torch.Tensor, comm=None) -> PCAResult:
    """``daal_pca/svddensedistr`` (``…/svddensedistr/PCADaalCollectiveMapper.java:116-164``): SVD of the
    z-scored data through TSQR."""
    mom = low_order_moments(X, comm)
    if X.device.type == "cuda" and X.dtype == torch.float32:
        Z = (X - mom.mean.to(X)) / mom.standard_deviation.clamp(min=1e-300).to(X)     # stays fp32 for the K21 kernel
    else:
        Z = (X.double() - mom.mean) / mom.standard_deviation.clamp(min=1e-300)
    R = tsqr_r(Z, comm)
    _, S, Vh = torch.linalg.svd(R)
    ev = S * S / max(mom.n - 1, 1)
    sgn = torch.sign(Vh.gather(1, Vh.abs().argmax(1, keepdim=True)))
    return PCAResult(ev, Vh * sgn, mom.mean)


# ---------------------------------------------------------------------------------------
# QR / SVD (daal_qr, daal_svd: step1 local → step2 master → bcast → step3 local)
# ---------------------------------------------------------------------------------------
def qr_distributed(X: torch.Tensor, comm=None) -> Tuple[torch.Tensor, torch.Tensor]:
    """Returns (local rows of Q, R) — ``SVDDaalCollectiveMapper.java:136-219`` 3-step protocol."""
    R = tsqr_r(X, comm)
    if X.device.type == "cuda" and X.dtype == torch.float32:
        # keep the n x d product in fp32 on the device: Q = X R^-1 as one library GEMM against the small inverse
        Rinv = torch.linalg.inv(R).to(torch.float32)
        return X @ Rinv, R
    Q = torch.linalg.solve_triangular(R, X.double(), upper=True, left=False)
    return Q, R


def svd_distributed(X: torch.Tensor, comm=None) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """Returns (local rows of U, singular values, Vᵀ)."""
    Q, R = qr_distributed(X, comm)
    Ur, S, Vh = torch.linalg.svd(R)
    return Q @ Ur.to(Q.dtype), S, Vh


def cholesky(A: torch.Tensor) -> torch.Tensor:
    """``daal_cholesky``: lower-triangular factor."""
    return torch.linalg.cholesky(A.double())


def pivoted_qr(A: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """``daal_pivoted_qr``: Householder QR with column pivoting, returns (Q, R, permutation) with
    ``A[:, perm] = Q R``."""
    A = A.double().clone()
    m, n = A.shape
    k = min(m, n)
    perm = torch.arange(n, device=A.device)
    Q = torch.eye(m, dtype=A.dtype, device=A.device)
    norms = (A * A).sum(0)
    for j in range(k):
        p = j + int(torch.argmax(norms[j:]))
        if p != j:
            A[:, [j, p]] = A[:, [p, j]]
            perm[[j, p]] = perm[[p, j]]
            norms[[j, p]] = norms[[p, j]]
        x = A[j:, j]
        nx = torch.linalg.norm(x)
        if nx > 0:
            v = x.clone()
            v[0] += torch.sign(x[0]) * nx if x[0] != 0 else nx
            v = v / torch.linalg.norm(v)
            A[j:, j:] -= 2.0 * torch.outer(v, v @ A[j:, j:])
            Q[:, j:] -= 2.0 * torch.outer(Q[:, j:] @ v, v)
        norms[j + 1:] = (A[j + 1:, j + 1:] ** 2).sum(0) if j + 1 < n else norms[j + 1:]
    R = torch.triu(A[:k, :])
    return Q[:, :k], R, perm


# ---------------------------------------------------------------------------------------
# linear / ridge regression (daal_linreg/{normaleq,qrdense}, daal_ridgereg)
# ---------------------------------------------------------------------------------------
class LinearModel(NamedTuple):
    beta: torch.Tensor      # [n_dep, n_features + 1], column 0 = intercept (DAAL layout)

    def predict(self, X: torch.Tensor) -> torch.Tensor:
        b = self.beta.to(X.device)
        return X.double() @ b[:, 1:].t() + b[:, 0]


def _normal_eq_partials(X: torch.Tensor, Y: torch.Tensor, comm):
    """ZᵀZ of Z = [1 | X | Y] assembled from its blocks WITHOUT building Z: XᵀX from the (fused, allreduced) Gram of X
    itself, [Σx ; YᵀX] from one streaming pass (K14, ``ops.gram.xty``, float64 accumulation — no TF32 in the intercept
    and Xᵀy terms), and the tiny Y blocks in float64."""
    n, d = X.shape
    k = Y.shape[1]
    width = 1 + d + k
    if X.device.type == "cuda" and X.dtype == torch.float32 and d % 32 == 0 and k <= 8:
        from ..ops.gram import centered_gram, xty
        small = torch.zeros(k + 2, max(d, k + 1), dtype=torch.float64, device=X.device)
        small[:k + 1, :d] = xty(X, Y)
        Yd = Y.double()
        small[k + 1, 0] = float(n)
        small[k + 1, 1:1 + k] = Yd.sum(0)
        yty = Yd.t() @ Yd
        _allreduce_(small, comm); _allreduce_(yty, comm)
        G = _gram_allreduced(X, comm)                                   # [d, d] float64, already global
        # The tensor core truncates its inputs to TF32; on columns with a non-zero mean that one-sided error is a
        # rank-one bias ~ -7e-4 N m mᵀ whose norm grows with d |m|² and soon rivals the covariance the solve depends on.
        # Such data gets XᵀX rebuilt, in float64, from the Gram of the CENTRED rows and the exact column sums.
        N = float(small[k + 1, 0].item())
        mean = small[0, :d] / N
        var = (G.diag() / N - mean * mean).clamp(min=0)
        if bool((mean.abs() > 0.05 * var.sqrt()).any().item()):
            m32 = mean.to(torch.float32)
            Gc = _f64(centered_gram(X, m32))
            _allreduce_(Gc, comm)
            sx, m = small[0, :d], m32.double()
            G = Gc + torch.outer(sx, m) + torch.outer(m, sx) - N * torch.outer(m, m)
        Z = torch.zeros(width, width, dtype=torch.float64, device=X.device)
        Z[0, 0] = small[k + 1, 0]
        Z[0, 1:1 + d] = small[0, :d]; Z[1:1 + d, 0] = small[0, :d]
        Z[1:1 + d, 1:1 + d] = G
        Z[1 + d:, 1:1 + d] = small[1:k + 1, :d]; Z[1:1 + d, 1 + d:] = small[1:k + 1, :d].t()
        Z[0, 1 + d:] = small[k + 1, 1:1 + k]; Z[1 + d:, 0] = small[k + 1, 1:1 + k]
        Z[1 + d:, 1 + d:] = yty
        return Z
    pad = (-width) % 32 if X.device.type == "cuda" else 0          # tensor-core path wants d % 32 == 0
    Z = torch.zeros(n, width + pad, dtype=X.dtype, device=X.device)
    Z[:, 0] = 1
    Z[:, 1:1 + d] = X
    Z[:, 1 + d:width] = Y
    return _gram_allreduced(Z, comm)[:width, :width].contiguous()


def linear_regression_normeq(X: torch.Tensor, Y: torch.Tensor, comm=None, ridge: float = 0.0) -> LinearModel:
    """``daal_linreg/normaleq`` (``LinRegDaalCollectiveMapper.java:159-205``) and, with ``ridge > 0``,
    ``daal_ridgereg`` (the intercept is not penalised)."""
    if Y.dim() == 1:
        Y = Y[:, None]
    d = X.shape[1]
    G = _normal_eq_partials(X, Y, comm)
    A = G[:1 + d, :1 + d].clone()
    B = G[:1 + d, 1 + d:]
    if ridge > 0:
        A[1:, 1:] += ridge * torch.eye(d, dtype=A.dtype, device=A.device)
    beta = torch.linalg.solve(A, B)          # [(1+d), k]
    return LinearModel(beta.t().contiguous())


def ridge_regression(X, Y, comm=None, ridge: float = 1.0) -> LinearModel:
    return linear_regression_normeq(X, Y, comm, ridge=ridge)


def linear_regression_qr(X: torch.Tensor, Y: torch.Tensor, comm=None) -> LinearModel:
    """``daal_linreg/qrdense``: QR of [1 | X | Y] through TSQR, then back-substitution."""
    if Y.dim() == 1:
        Y = Y[:, None]
    n, d = X.shape
    Z = torch.cat([torch.ones(n, 1, dtype=torch.float64, device=X.device), X.double(), Y.double()], 1)
    R = tsqr_r(Z, comm)
    R11, R12 = R[:1 + d, :1 + d], R[:1 + d, 1 + d:]
    beta = torch.linalg.solve_triangular(R11, R12, upper=True)
    return LinearModel(beta.t().contiguous())


# ---------------------------------------------------------------------------------------
# normalisation, quantiles, sorting, outliers, quality metrics (batch algorithms)
# ---------------------------------------------------------------------------------------
def zscore(X: torch.Tensor) -> torch.Tensor:
    """``daal_normalization/ZscoreDenseBatch``."""
    Xd = X.double()
    return (Xd - Xd.mean(0)) / Xd.std(0, unbiased=True)


def minmax(X: torch.Tensor, lower: float = 0.0, upper: float = 1.0) -> torch.Tensor:
    """``daal_normalization/MinMaxDenseBatch``."""
    Xd = X.double()
    mn, mx = Xd.min(0).values, Xd.max(0).values
    return lower + (Xd - mn) * (upper - lower) / (mx - mn).clamp(min=1e-300)


def quantiles(X: torch.Tensor, orders=(0.25, 0.5, 0.75)) -> torch.Tensor:
    """``daal_quantile``: per-column quantiles, DAAL's definition = order statistic ``x_(ceil(q n))``
    (no interpolation).  Returns ``[n_columns, n_orders]``."""
    Xs = torch.sort(X.double(), dim=0).values
    n = X.shape[0]
    idx = [min(n - 1, max(0, math.ceil(q * n) - 1)) for q in orders]
    return Xs[idx].t().contiguous()


def sort_columns(X: torch.Tensor) -> torch.Tensor:
    """``daal_sorting``: every column sorted ascending."""
    return torch.sort(X, dim=0).values


def outlier_univariate(X: torch.Tensor, threshold: float = 3.0) -> torch.Tensor:
    """``daal_outlier/OutUniDenseBatch``: weight 0 where ``|x - mean| / sd > threshold`` else 1."""
    Xd = X.double()
    z = (Xd - Xd.mean(0)).abs() / Xd.std(0, unbiased=True).clamp(min=1e-300)
    return (z <= threshold).to(torch.float64)


def outlier_multivariate(X: torch.Tensor, threshold: float = 3.0) -> torch.Tensor:
    """``daal_outlier/OutMultDenseBatch``: Mahalanobis distance to the sample mean / covariance."""
    Xd = X.double()
    mu = Xd.mean(0)
    C = torch.cov(Xd.t())
    diff = Xd - mu
    m = (diff @ torch.linalg.inv(C) * diff).sum(1).clamp(min=0).sqrt()
    return (m <= threshold).to(torch.float64)[:, None]


def outlier_bacon(X: torch.Tensor, alpha: float = 0.05, init_frac: float = 0.5, max_iter: int = 50) -> torch.Tensor:
    """``daal_outlier/OutBaconDenseBatch``: BACON blocked-adaptive outlier nominator (median start)."""
    Xd = X.double()
    n, p = Xd.shape
    med = Xd.median(0).values
    d0 = torch.linalg.norm(Xd - med, dim=1)
    m = max(p + 1, int(init_frac * n))
    basic = torch.zeros(n, dtype=torch.bool, device=X.device)
    basic[torch.argsort(d0)[:m]] = True
    try:
        from scipy.stats import chi2
        chi = math.sqrt(chi2.ppf(1 - alpha / n, p))
    except Exception:  # noqa: BLE001
        chi = math.sqrt(p) + 3.0
    for _ in range(max_iter):
        sub = Xd[basic]
        mu = sub.mean(0)
        C = torch.cov(sub.t()) + 1e-12 * torch.eye(p, dtype=Xd.dtype, device=X.device)
        diff = Xd - mu
        dist_m = (diff @ torch.linalg.inv(C) * diff).sum(1).clamp(min=0).sqrt()
        r = int(basic.sum())
        h = (n + p + 1) // 2
        c_np = 1 + (p + 1) / (n - p) + 2 / (n - 1 - 3 * p) if n - 1 - 3 * p > 0 else 1.0
        c_hr = max(0.0, (h - r) / (h + r))
        new_basic = dist_m < (c_np + c_hr) * chi
        if torch.equal(new_basic, basic):
            break
        basic = new_basic
    return basic.to(torch.float64)[:, None]


def confusion_matrix(y_true: torch.Tensor, y_pred: torch.Tensor, n_classes: int) -> torch.Tensor:
    idx = y_true.long() * n_classes + y_pred.long()
    return torch.bincount(idx, minlength=n_classes * n_classes).view(n_classes, n_classes)


def quality_metrics_classification(y_true, y_pred, n_classes: int) -> Dict[str, float]:
    """``daal_quality_metrics`` (binary / multi-class confusion matrix metrics, macro averages)."""
    cm = confusion_matrix(y_true, y_pred, n_classes).double()
    tp = cm.diag()
    fp = cm.sum(0) - tp
    fn = cm.sum(1) - tp
    prec = tp / (tp + fp).clamp(min=1)
    rec = tp / (tp + fn).clamp(min=1)
    f1 = 2 * prec * rec / (prec + rec).clamp(min=1e-300)
    return {"accuracy": float(tp.sum() / cm.sum()), "precision": float(prec.mean()), "recall": float(rec.mean()),
            "fscore": float(f1.mean()), "error_rate": float(1 - tp.sum() / cm.sum()),
            "specificity": float(((cm.sum() - tp - fp - fn) / (cm.sum() - tp - fn).clamp(min=1)).mean())}


def quality_metrics_regression(y_true: torch.Tensor, y_pred: torch.Tensor) -> Dict[str, float]:
    """Linear-regression quality metrics (single beta group): RMSE, variance, R²."""
    yt, yp = y_true.double().flatten(), y_pred.double().flatten()
    res = yt - yp
    ss_res = float((res * res).sum())
    ss_tot = float(((yt - yt.mean()) ** 2).sum())
    return {"rms": math.sqrt(ss_res / yt.numel()), "variance": ss_res / max(yt.numel() - 1, 1),
            "r2": 1.0 - ss_res / max(ss_tot, 1e-300)}
