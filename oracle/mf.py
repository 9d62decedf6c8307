"""NumPy float64 restatements of the reference's matrix-factorisation hot loops, written from the Java sources and
independent of harp_b200's own CPU branches.

``sgd_pass``   ``ml/java/src/main/java/edu/iu/sgd/SGDMPTask.java:60-72``: for every rating (i, j, v), in list order,
               e = v - w_i . h_j;  w_i += eps (e h_j - lambda w_i);  h_j += eps (e w_i_old - lambda h_j).
``ccd_iteration`` ``ml/java/src/main/java/edu/iu/ccd/CCDMPTask.java:67-125`` (doColCCD then doRowCCD for every
               feature, as ``CCDMPCollectiveMapper`` schedules them): one residual array m1 = v - W'H shared by both.
"""
import numpy as np


def sgd_pass(W: np.ndarray, H: np.ndarray, rows, cols, vals, eps: float, lam: float) -> None:
    """In place, sequential, rating by rating (W [n_rows, r], H [n_cols, r])."""
    for i, j, v in zip(rows, cols, vals):
        w = W[i].copy()
        h = H[j]
        e = v - float(w @ h)
        W[i] = w + eps * (e * h - lam * w)
        H[j] = h + eps * (e * w - lam * h)


def rmse(W: np.ndarray, H: np.ndarray, rows, cols, vals) -> float:
    pred = np.einsum("ij,ij->i", W[np.asarray(rows)], H[np.asarray(cols)])
    return float(np.sqrt(np.mean((np.asarray(vals) - pred) ** 2)))


def ccd_iteration(W: np.ndarray, H: np.ndarray, res: np.ndarray, rows, cols, lam: float, inner: int = 1) -> float:
    """W [r, n_rows], H [r, n_cols] feature-major, res = v - sum_f W[f, i] H[f, j] per rating; all updated in place.
    Returns the training RMSE after the iteration."""
    rows, cols = np.asarray(rows), np.asarray(cols)
    by_col = [np.nonzero(cols == j)[0] for j in range(H.shape[1])]
    by_row = [np.nonzero(rows == i)[0] for i in range(W.shape[1])]
    for f in range(W.shape[0]):
        for _ in range(inner):
            for j, ks in enumerate(by_col):                      # doColCCD
                if ks.size:
                    wt, ht = W[f, rows[ks]], H[f, j]
                    s = float(((res[ks] + ht * wt) * wt).sum() / (lam * ks.size + (wt * wt).sum()))
                    res[ks] -= (s - ht) * wt
                    H[f, j] = s
            for i, ks in enumerate(by_row):                      # doRowCCD
                if ks.size:
                    ht, wt = H[f, cols[ks]], W[f, i]
                    z = float(((res[ks] + wt * ht) * ht).sum() / (lam * ks.size + (ht * ht).sum()))
                    res[ks] -= (z - wt) * ht
                    W[f, i] = z
    return float(np.sqrt(np.mean(res ** 2)))
