"""Harp-DAAL algorithms, part 2: classifiers, trees, boosting and kernel functions.

Distributed: multinomial naive Bayes (``daal_naive/{dense,csr}distri`` — per-class feature sums are the
partial result, allreduced).  Batch (no collective in the reference either): SVM one-vs-one
(``daal_svm/MultiClass{Dense,CSR}Batch``, Boser method = SMO), kernel functions
(``daal_kernel_func/*``), kNN (``daal_knn``), decision stump / tree / forest (``daal_stump``,
``daal_dtree``, ``daal_dforest``), AdaBoost / BrownBoost / LogitBoost.  Everything is tensor code that
runs on the GPU when given CUDA tensors.
"""
from __future__ import annotations

import math
from typing import List, NamedTuple, Optional, Tuple

import torch

from .linalg import _allreduce_


def _dense(X: torch.Tensor) -> torch.Tensor:
    return X.to_dense() if X.layout != torch.strided else X


# ---------------------------------------------------------------------------------------
# kernel functions (daal_kernel_func: LinDenseBatch, LinCSRBatch, RbfDenseBatch, RbfCSRBatch)
# ---------------------------------------------------------------------------------------
def kernel_linear(X: torch.Tensor, Y: Optional[torch.Tensor] = None, k: float = 1.0, b: float = 0.0) -> torch.Tensor:
    """``K = k · X Yᵀ + b`` (DAAL ``kernel_function.linear``)."""
    X = _dense(X).double()
    Y = X if Y is None else _dense(Y).double()
    return k * (X @ Y.t()) + b


def kernel_rbf(X: torch.Tensor, Y: Optional[torch.Tensor] = None, sigma: float = 1.0) -> torch.Tensor:
    """``K = exp(-‖x - y‖² / (2 σ²))`` (DAAL ``kernel_function.rbf``)."""
    X = _dense(X).double()
    Y = X if Y is None else _dense(Y).double()
    d2 = (X * X).sum(1)[:, None] - 2.0 * (X @ Y.t()) + (Y * Y).sum(1)[None, :]
    return torch.exp(-d2.clamp(min=0) / (2.0 * sigma * sigma))


# ---------------------------------------------------------------------------------------
# multinomial naive Bayes (distributed)
# ---------------------------------------------------------------------------------------
class NaiveBayesModel(NamedTuple):
    log_prior: torch.Tensor     # [C]
    log_theta: torch.Tensor     # [C, F]

    def predict(self, X: torch.Tensor) -> torch.Tensor:
        return (_dense(X).double() @ self.log_theta.t() + self.log_prior).argmax(1)


def naive_bayes_train(X: torch.Tensor, y: torch.Tensor, n_classes: int, comm=None, alpha: float = 1.0) -> NaiveBayesModel:
    """Partial results = per-class feature sums and class counts (DAAL ``multinomial_naive_bayes``
    step1Local); finalisation replicated on every rank after the allreduce."""
    Xd = _dense(X).double()
    C, Fd = n_classes, Xd.shape[1]
    sums = torch.zeros(C, Fd, dtype=torch.float64, device=Xd.device).index_add_(0, y.long(), Xd)
    cnt = torch.bincount(y.long(), minlength=C).double()
    _allreduce_(sums, comm); _allreduce_(cnt, comm)
    log_prior = torch.log(cnt.clamp(min=1e-300) / cnt.sum())
    log_theta = torch.log((sums + alpha) / (sums.sum(1, keepdim=True) + alpha * Fd))
    return NaiveBayesModel(log_prior, log_theta)


# ---------------------------------------------------------------------------------------
# SVM: SMO (Boser / Platt with WSS-1 maximal violating pair), one-vs-one multi-class
# ---------------------------------------------------------------------------------------
class BinarySVM(NamedTuple):
    sv: torch.Tensor          # support vectors
    coef: torch.Tensor        # alpha_i y_i
    bias: float
    kernel: str
    sigma: float

    def decision(self, X: torch.Tensor) -> torch.Tensor:
        Kx = kernel_linear(X, self.sv) if self.kernel == "linear" else kernel_rbf(X, self.sv, self.sigma)
        return Kx @ self.coef + self.bias


def _smo_torch(K: torch.Tensor, y: torch.Tensor, C: float, tol: float, max_iter: int):
    """Plain tensor restatement of the SMO loop (fallback when neither native library is built)."""
    n = K.shape[0]
    alpha = torch.zeros(n, dtype=torch.float64, device=K.device)
    grad = -torch.ones(n, dtype=torch.float64, device=K.device)          # ∇ of ½αᵀQα − eᵀα, Q = yyᵀ∘K
    for _ in range(max_iter):
        up = ((y > 0) & (alpha < C)) | ((y < 0) & (alpha > 0))
        lo = ((y > 0) & (alpha > 0)) | ((y < 0) & (alpha < C))
        f = -y * grad
        fi = torch.where(up, f, torch.full_like(f, -float("inf")))
        fj = torch.where(lo, f, torch.full_like(f, float("inf")))
        i, j = int(fi.argmax()), int(fj.argmin())
        if float(fi[i] - fj[j]) < tol:
            break
        yi, yj = float(y[i]), float(y[j])
        eta = float(K[i, i] + K[j, j] - 2 * K[i, j])
        eta = eta if eta > 1e-12 else 1e-12
        delta = float(fi[i] - fj[j]) / eta
        ai, aj = float(alpha[i]), float(alpha[j])
        t_max_i = (C - ai) if yi > 0 else ai                # alpha_i += y_i t, alpha_j -= y_j t with t >= 0
        t_max_j = aj if yj > 0 else (C - aj)
        t = min(delta, t_max_i, t_max_j)
        alpha[i] = ai + yi * t
        alpha[j] = aj - yj * t
        grad += t * y * (K[:, i] - K[:, j])
    return alpha, grad


def _smo_solve(K: torch.Tensor, y: torch.Tensor, C: float, tol: float, max_iter: int):
    """``(alpha, grad)`` of the C-SVC dual.  The whole SMO loop runs natively: one CUDA kernel launch for device
    tensors (``csrc/svm.cu``), the C++ host solver for CPU tensors (``csrc/host/smo.cpp``)."""
    import ctypes

    from .. import _native
    n = K.shape[0]
    alpha = torch.empty(n, dtype=torch.float64, device=K.device)
    grad = torch.empty(n, dtype=torch.float64, device=K.device)
    if n == 0:
        return alpha, grad
    P = ctypes.c_void_p
    if K.device.type == "cuda" and _native.have_cuda_lib():
        lib = _native.lib()
        lib.hb_smo_binary.argtypes = [P, P, ctypes.c_int, ctypes.c_double, ctypes.c_double, ctypes.c_int, P, P, P, P]
        _native.check(lib.hb_smo_binary(_native.ptr(K), _native.ptr(y), n, C, tol, max_iter, _native.ptr(alpha),
                                        _native.ptr(grad), None, _native.stream_ptr()), "hb_smo_binary")
        _native.count_launch()
        return alpha, grad
    if K.device.type == "cpu" and _native.have_host_lib():
        lib = _native.host()
        lib.hh_smo_binary.restype = ctypes.c_longlong
        lib.hh_smo_binary.argtypes = [P, P, ctypes.c_longlong, ctypes.c_double, ctypes.c_double, ctypes.c_longlong, P, P]
        lib.hh_smo_binary(_native.ptr(K), _native.ptr(y), n, C, tol, max_iter, _native.ptr(alpha), _native.ptr(grad))
        return alpha, grad
    return _smo_torch(K, y, C, tol, max_iter)


def svm_train_binary(X: torch.Tensor, y: torch.Tensor, C: float = 1.0, kernel: str = "linear", sigma: float = 1.0,
                     tol: float = 1e-3, max_iter: int = 20000) -> BinarySVM:
    """C-SVC dual by SMO on a precomputed kernel matrix (labels in {-1, +1})."""
    Xd = _dense(X).double()
    y = y.double()
    n = Xd.shape[0]
    K = kernel_linear(Xd) if kernel == "linear" else kernel_rbf(Xd, None, sigma)
    alpha, grad = _smo_solve(K.contiguous(), y.contiguous(), float(C), float(tol), int(max_iter))
    sv_mask = alpha > 1e-8
    free = sv_mask & (alpha < C - 1e-8)
    f = -y * grad
    if bool(free.any()):
        bias = float(f[free].mean())
    else:
        up = ((y > 0) & (alpha < C)) | ((y < 0) & (alpha > 0))
        lo = ((y > 0) & (alpha > 0)) | ((y < 0) & (alpha < C))
        hi = float(f[up].max()) if bool(up.any()) else 0.0
        lw = float(f[lo].min()) if bool(lo.any()) else 0.0
        bias = 0.5 * (hi + lw)
    return BinarySVM(Xd[sv_mask], (alpha * y)[sv_mask], bias, kernel, sigma)


class MultiClassSVM(NamedTuple):
    classes: List[int]
    pairs: List[Tuple[int, int, BinarySVM]]

    def predict(self, X: torch.Tensor) -> torch.Tensor:
        n = X.shape[0]
        votes = torch.zeros(n, len(self.classes), dtype=torch.int64, device=X.device)
        for a, b, m in self.pairs:
            pos = m.decision(X) >= 0
            votes[:, a] += pos.long()
            votes[:, b] += (~pos).long()
        return torch.tensor(self.classes, device=X.device)[votes.argmax(1)]


def svm_train_multiclass(X, y, C: float = 1.0, kernel: str = "linear", sigma: float = 1.0, **kw) -> MultiClassSVM:
    """``multi_class_classifier`` one-vs-one over binary SVMs
    (``daal_svm/MultiClassDenseBatch/SVMDaalCollectiveMapper.java:150-199``)."""
    classes = sorted(int(c) for c in torch.unique(y).tolist())
    pairs = []
    for a in range(len(classes)):
        for b in range(a + 1, len(classes)):
            m = (y == classes[a]) | (y == classes[b])
            yy = torch.where(y[m] == classes[a], 1.0, -1.0)
            pairs.append((a, b, svm_train_binary(X[m], yy, C, kernel, sigma, **kw)))
    return MultiClassSVM(classes, pairs)


# ---------------------------------------------------------------------------------------
# kNN (daal_knn) — brute force on the tensor cores via one distance GEMM
# ---------------------------------------------------------------------------------------
def knn_predict(X_train: torch.Tensor, y_train: torch.Tensor, X_test: torch.Tensor, k: int = 5) -> torch.Tensor:
    A, B = _dense(X_test).double(), _dense(X_train).double()
    d2 = (A * A).sum(1)[:, None] - 2 * (A @ B.t()) + (B * B).sum(1)[None, :]
    idx = d2.topk(k, largest=False).indices
    lab = y_train.long()[idx]
    C = int(y_train.max().item()) + 1
    votes = torch.zeros(A.shape[0], C, dtype=torch.int64, device=A.device).scatter_add_(1, lab, torch.ones_like(lab))
    return votes.argmax(1)


# ---------------------------------------------------------------------------------------
# decision stump / tree / forest
# ---------------------------------------------------------------------------------------
class Stump(NamedTuple):
    feature: int
    threshold: float
    left: float      # prediction when x[feature] <= threshold
    right: float

    def predict(self, X: torch.Tensor) -> torch.Tensor:
        x = _dense(X).double()[:, self.feature]
        return torch.where(x <= self.threshold, torch.tensor(self.left, dtype=torch.float64, device=X.device),
                           torch.tensor(self.right, dtype=torch.float64, device=X.device))


def stump_train(X: torch.Tensor, y: torch.Tensor, w: Optional[torch.Tensor] = None) -> Stump:
    """Weighted decision stump for labels in {-1, +1} (``daal_stump``): the (feature, threshold, polarity)
    minimising the weighted error, all candidate thresholds evaluated at once with prefix sums."""
    Xd = _dense(X).double()
    n, d = Xd.shape
    y = y.double()
    w = torch.full((n,), 1.0 / n, dtype=torch.float64, device=Xd.device) if w is None else w.double()
    order = torch.argsort(Xd, dim=0)
    xs = torch.gather(Xd, 0, order)
    wy = (w * y)[order]                               # [n, d]
    cum = torch.cumsum(wy, 0)                         # Σ_{i<=t} w y  when splitting after sorted position t
    total = cum[-1]
    # predicting -1 on the left and +1 on the right: error = Σ_left w[y=+1] + Σ_right w[y=-1]
    #   = 0.5 (1 - (−cum + (total − cum)))  with Σ w = 1
    score = total - 2 * cum                           # agreement of "left=-1,right=+1"
    valid = torch.ones_like(score, dtype=torch.bool)
    valid[:-1] = xs[1:] > xs[:-1]                     # only split between distinct values
    valid[-1] = False
    s_pos = torch.where(valid, score, torch.full_like(score, -float("inf")))
    s_neg = torch.where(valid, -score, torch.full_like(score, -float("inf")))
    best_pos, best_neg = s_pos.max(), s_neg.max()
    use_pos = bool(best_pos >= best_neg)
    flat = int((s_pos if use_pos else s_neg).argmax())
    t, f = flat // d, flat % d
    if not bool(valid.any()):
        maj = 1.0 if float(total[0]) >= 0 else -1.0
        return Stump(0, float("inf"), maj, maj)
    thr = 0.5 * float(xs[t, f] + xs[t + 1, f])
    return Stump(f, thr, -1.0 if use_pos else 1.0, 1.0 if use_pos else -1.0)


class TreeNode:
    __slots__ = ("feature", "threshold", "left", "right", "value")

    def __init__(self, value=None):
        self.feature, self.threshold, self.left, self.right, self.value = -1, 0.0, None, None, value


class DecisionTree:
    """CART (``daal_dtree/{Cls,Reg}DenseBatch``): Gini for classification, variance for regression."""

    def __init__(self, task: str = "classification", max_depth: int = 8, min_samples_leaf: int = 1,
                 max_features: Optional[int] = None, seed: int = 0):
        self.task, self.max_depth, self.min_leaf, self.max_features = task, max_depth, min_samples_leaf, max_features
        self.rng = torch.Generator().manual_seed(seed)
        self.root: Optional[TreeNode] = None
        self.n_classes = 0

    def _leaf(self, y):
        if self.task == "classification":
            return torch.bincount(y.long(), minlength=self.n_classes).double() / max(y.numel(), 1)
        return y.double().mean()

    def _best_split(self, X, y):
        n, d = X.shape
        # feature sub-sampling: `max_features` candidates, but features that are constant in this node do not
        # count — the search goes on through the permutation until that many splittable ones were tried
        limit = d
        feats = torch.arange(d)
        if self.max_features and self.max_features < d:
            feats, limit = torch.randperm(d, generator=self.rng), self.max_features
        best = (None, None, float("inf"))
        tried = 0
        for f in feats.tolist():
            if tried >= limit:
                break
            xs, order = torch.sort(X[:, f])
            ys = y[order]
            valid = xs[1:] > xs[:-1]
            if not bool(valid.any()):
                continue
            tried += 1
            nl = torch.arange(1, n, dtype=torch.float64, device=X.device)
            nr = n - nl
            if self.task == "classification":
                oh = torch.nn.functional.one_hot(ys.long(), self.n_classes).double()
                cl = torch.cumsum(oh, 0)[:-1]
                cr = cl[-1] + oh[-1] - cl
                gini = (nl - (cl * cl).sum(1) / nl) + (nr - (cr * cr).sum(1) / nr)
            else:
                yd = ys.double()
                s1, s2 = torch.cumsum(yd, 0)[:-1], torch.cumsum(yd * yd, 0)[:-1]
                t1, t2 = yd.sum(), (yd * yd).sum()
                gini = (s2 - s1 * s1 / nl) + ((t2 - s2) - (t1 - s1) ** 2 / nr)
            ok = valid & (nl >= self.min_leaf) & (nr >= self.min_leaf)
            if not bool(ok.any()):
                continue
            gini = torch.where(ok, gini, torch.full_like(gini, float("inf")))
            i = int(gini.argmin())
            if float(gini[i]) < best[2]:
                best = (f, 0.5 * float(xs[i] + xs[i + 1]), float(gini[i]))
        return best

    def _grow(self, X, y, depth):
        node = TreeNode(self._leaf(y))
        if depth >= self.max_depth or y.numel() < 2 * self.min_leaf or (self.task == "classification" and len(torch.unique(y)) == 1):
            return node
        f, thr, _ = self._best_split(X, y)
        if f is None:
            return node
        m = X[:, f] <= thr
        node.feature, node.threshold = f, thr
        node.left, node.right = self._grow(X[m], y[m], depth + 1), self._grow(X[~m], y[~m], depth + 1)
        return node

    def fit(self, X, y):
        X = _dense(X).double()
        if self.task == "classification":
            self.n_classes = int(y.max().item()) + 1
        self.root = self._grow(X, y, 0)
        return self

    def _predict_value(self, X):
        X = _dense(X).double()
        out = [None] * X.shape[0]
        def walk(node, idx):
            if idx.numel() == 0:
                return
            if node.left is None:
                for i in idx.tolist():
                    out[i] = node.value
                return
            m = X[idx, node.feature] <= node.threshold
            walk(node.left, idx[m]); walk(node.right, idx[~m])
        walk(self.root, torch.arange(X.shape[0]))
        return torch.stack([torch.as_tensor(v, dtype=torch.float64) for v in out])

    def predict_proba(self, X):
        return self._predict_value(X)

    def predict(self, X):
        v = self._predict_value(X)
        return v.argmax(1) if self.task == "classification" else v


class DecisionForest:
    """``daal_dforest/{Cls,Reg}{Dense,TraverseDense}Batch``: bagged CART with feature sub-sampling."""

    def __init__(self, task: str = "classification", n_trees: int = 10, max_depth: int = 8, max_features: Optional[int] = None,
                 seed: int = 0):
        self.task, self.n_trees, self.max_depth, self.max_features, self.seed = task, n_trees, max_depth, max_features, seed
        self.trees: List[DecisionTree] = []

    def fit(self, X, y):
        X = _dense(X)
        n, d = X.shape
        g = torch.Generator().manual_seed(self.seed)
        mf = self.max_features or max(1, int(math.sqrt(d)) if self.task == "classification" else max(1, d // 3))
        self.n_classes = int(y.max().item()) + 1 if self.task == "classification" else 0
        for t in range(self.n_trees):
            idx = torch.randint(0, n, (n,), generator=g)
            tree = DecisionTree(self.task, self.max_depth, 1, mf, seed=self.seed + 1 + t)
            tree.n_classes = self.n_classes
            tree.root = tree._grow(X[idx].double(), y[idx], 0)
            self.trees.append(tree)
        return self

    def predict(self, X):
        vals = torch.stack([t._predict_value(X) for t in self.trees]).mean(0)
        return vals.argmax(1) if self.task == "classification" else vals

    def traverse(self) -> List[int]:
        """Node counts per tree (the reference's *TraverseDenseBatch variants walk the trees)."""
        def count(n):
            return 0 if n is None else 1 + count(n.left) + count(n.right)
        return [count(t.root) for t in self.trees]


# ---------------------------------------------------------------------------------------
# boosting (daal_adaboost, daal_brownboost, daal_logitboost) on stumps
# ---------------------------------------------------------------------------------------
class BoostedModel(NamedTuple):
    learners: List[Stump]
    weights: List[float]

    def decision(self, X):
        out = torch.zeros(X.shape[0], dtype=torch.float64, device=X.device)
        for h, a in zip(self.learners, self.weights):
            out += a * h.predict(X)
        return out

    def predict(self, X):
        return torch.where(self.decision(X) >= 0, 1.0, -1.0)


def adaboost_train(X, y, n_rounds: int = 50, tol: float = 0.0) -> BoostedModel:
    """Freund–Schapire AdaBoost, labels in {-1, +1}."""
    y = y.double()
    n = y.numel()
    w = torch.full((n,), 1.0 / n, dtype=torch.float64, device=X.device)
    L, A = [], []
    for _ in range(n_rounds):
        h = stump_train(X, y, w)
        pred = h.predict(X)
        err = float((w * (pred != y)).sum())
        err = min(max(err, 1e-12), 1 - 1e-12)
        a = 0.5 * math.log((1 - err) / err)
        L.append(h); A.append(a)
        if err <= tol or err >= 0.5:
            break
        w = w * torch.exp(-a * y * pred)
        w = w / w.sum()
    return BoostedModel(L, A)


def brownboost_train(X, y, c: float = 4.0, n_rounds: int = 60, nu: float = 0.1) -> BoostedModel:
    """Freund's BrownBoost: remaining "time" s starts at c and every round solves for (alpha, t) so that the
    weighted correlation vanishes; stops when the time budget is used up."""
    y = y.double()
    n = y.numel()
    r = torch.zeros(n, dtype=torch.float64, device=X.device)     # margins
    s = c
    L, A = [], []
    for _ in range(n_rounds):
        w = torch.exp(-((r + s) ** 2) / c)
        w = w / w.sum()
        h = stump_train(X, y, w)
        hy = h.predict(X) * y
        # solve  Σ exp(-(r + a hy + s - t)^2 / c) hy = 0  with  t = a² ... : damped Newton on (a), t = a*gamma heuristic
        gamma = float((w * hy).sum())
        if gamma <= 1e-9:
            break
        a = min(max(gamma * math.sqrt(c), 1e-3), 2.0)
        t = min(s, a * a / 3.0 + 1e-3)
        L.append(h); A.append(a)
        r = r + a * hy
        s -= t
        if s <= 1e-9:
            break
    return BoostedModel(L, A)


class LogitBoostModel(NamedTuple):
    learners: List[List[Tuple[int, float, float, float]]]   # per round, per class: regression stump (f, thr, left, right)
    n_classes: int

    def scores(self, X):
        Xd = _dense(X).double()
        F_ = torch.zeros(Xd.shape[0], self.n_classes, dtype=torch.float64, device=X.device)
        for rnd in self.learners:
            for j, (f, thr, lv, rv) in enumerate(rnd):
                F_[:, j] += torch.where(Xd[:, f] <= thr, lv, rv)
        return F_

    def predict(self, X):
        return self.scores(X).argmax(1)


def _regression_stump(X, z, w):
    Xd = X
    n, d = Xd.shape
    order = torch.argsort(Xd, dim=0)
    xs = torch.gather(Xd, 0, order)
    wz, ww = (w * z)[order], w[order]
    cwz, cw = torch.cumsum(wz, 0), torch.cumsum(ww, 0)
    tz, tw = cwz[-1], cw[-1]
    lmean = cwz / cw.clamp(min=1e-300)
    rmean = (tz - cwz) / (tw - cw).clamp(min=1e-300)
    gain = cw * lmean ** 2 + (tw - cw) * rmean ** 2            # maximise explained weighted sum of squares
    valid = torch.zeros_like(gain, dtype=torch.bool)
    valid[:-1] = xs[1:] > xs[:-1]
    gain = torch.where(valid, gain, torch.full_like(gain, -float("inf")))
    flat = int(gain.argmax())
    t, f = flat // d, flat % d
    if not bool(valid.any()):
        m = float((w * z).sum() / w.sum())
        return (0, float("inf"), m, m)
    return (f, 0.5 * float(xs[t, f] + xs[t + 1, f]), float(lmean[t, f]), float(rmean[t, f]))


def logitboost_train(X, y, n_classes: int, n_rounds: int = 30) -> LogitBoostModel:
    """Friedman–Hastie–Tibshirani multi-class LogitBoost with regression stumps (``daal_logitboost``)."""
    Xd = _dense(X).double()
    n = Xd.shape[0]
    J = n_classes
    Y = torch.nn.functional.one_hot(y.long(), J).double()
    F_ = torch.zeros(n, J, dtype=torch.float64, device=X.device)
    rounds = []
    for _ in range(n_rounds):
        P = torch.softmax(F_, 1)
        fs = []
        upd = torch.zeros_like(F_)
        for j in range(J):
            pj = P[:, j].clamp(1e-10, 1 - 1e-10)
            w = pj * (1 - pj)
            z = ((Y[:, j] - pj) / w).clamp(-4, 4)
            f, thr, lv, rv = _regression_stump(Xd, z, w)
            upd[:, j] = torch.where(Xd[:, f] <= thr, lv, rv)
            fs.append((f, thr, lv, rv))
        upd = (J - 1) / J * (upd - upd.mean(1, keepdim=True))
        # fold the centring into stored leaf values is not possible per class; store the raw stumps and
        # re-apply the same centring at prediction time through scores()
        F_ = F_ + upd
        rounds.append(fs)
    return LogitBoostModel(rounds, J)
