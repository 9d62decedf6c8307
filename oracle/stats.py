"""NumPy float64 restatements of the Harp-DAAL dense statistics the Gram kernels feed
(``ml/daal/src/main/java/edu/iu/daal_cov/densedistri/COVDaalCollectiveMapper.java:143-176`` partial results:
nObservations, sum, crossProduct; ``daal_linreg/normaleq/LinRegDaalCollectiveMapper.java:159-205``)."""
import numpy as np


def covariance_from_partials(blocks):
    """blocks: row blocks of X (one per worker).  Returns (covariance, mean) assembled from the per-block partial
    results exactly as the master step does: N, sum, cross product."""
    n = sum(b.shape[0] for b in blocks)
    s = sum(b.sum(0) for b in blocks)
    cp = sum(b.T @ b for b in blocks)
    mean = s / n
    return (cp - n * np.outer(mean, mean)) / (n - 1), mean


def linear_regression_normal_equations(blocks_x, blocks_y, ridge: float = 0.0):
    """Returns beta [(d + 1), m] with the intercept in row 0, from summed X'X, X'y partials."""
    xtx = sum(np.c_[np.ones(len(x)), x].T @ np.c_[np.ones(len(x)), x] for x in blocks_x)
    xty = sum(np.c_[np.ones(len(x)), x].T @ y for x, y in zip(blocks_x, blocks_y))
    reg = ridge * np.eye(xtx.shape[0]); reg[0, 0] = 0.0
    return np.linalg.solve(xtx + reg, xty)
