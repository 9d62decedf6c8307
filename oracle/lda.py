"""NumPy restatement of collapsed Gibbs sampling for LDA as the reference defines it
(``ml/java/src/main/java/edu/iu/lda/LDAMPTask.java:209-254``: remove the token, draw from
p(k) ~ (n_dk + alpha)(n_wk + beta) / (n_k + V beta), put it back) and of its likelihood
(``CalcLikelihoodTask.java:57-77`` + the document side).  Dense and sequential: an oracle for small corpora."""
import math

import numpy as np


def conditional(n_dk_row, n_wk_row, n_k, alpha: float, beta: float, V: int) -> np.ndarray:
    p = (n_dk_row + alpha) * (n_wk_row + beta) / (n_k + V * beta)
    return p / p.sum()


def gibbs_pass(docs, words, z, n_dk, n_wk, n_k, alpha: float, beta: float, rng) -> None:
    """One sweep over all tokens in order; tables and assignments updated in place."""
    V = n_wk.shape[0]
    for t in range(len(docs)):
        d, w, k = docs[t], words[t], z[t]
        n_dk[d, k] -= 1; n_wk[w, k] -= 1; n_k[k] -= 1
        k = int(rng.choice(n_k.size, p=conditional(n_dk[d], n_wk[w], n_k, alpha, beta, V)))
        n_dk[d, k] += 1; n_wk[w, k] += 1; n_k[k] += 1
        z[t] = k


def log_likelihood(n_dk, n_wk, n_k, alpha: float, beta: float) -> float:
    """log p(w, z) of the collapsed model (word side as CalcLikelihoodTask, plus the document side)."""
    lg = np.vectorize(math.lgamma)
    V, K = n_wk.shape
    word = K * (math.lgamma(V * beta) - V * math.lgamma(beta)) + lg(n_wk + beta).sum() - lg(n_k + V * beta).sum()
    D = n_dk.shape[0]
    doc = D * (math.lgamma(K * alpha) - K * math.lgamma(alpha)) + lg(n_dk + alpha).sum() - lg(n_dk.sum(1) + K * alpha).sum()
    return float(word + doc)
