"""NumPy float64 restatement of ml/java K-means regroup-allgather
(``CenCalcTask.run`` ``ml/java/.../kmeans/regroupallgather/CenCalcTask.java:67-100``,
``CenMergeTask.java:36-54``, division ``KMeansCollectiveMapper.java:170-183``)."""
import numpy as np


def lloyd_iteration(points: np.ndarray, centroids: np.ndarray):
    """Returns (new_centroids, counts, objective). Empty clusters collapse to the zero vector."""
    k, d = centroids.shape
    best = np.empty(points.shape[0], dtype=np.int64)
    obj = 0.0
    for s in range(0, points.shape[0], 8192):
        blk = points[s:s + 8192]
        d2 = ((blk[:, None, :] - centroids[None, :, :]) ** 2).sum(-1)   # exact (x-c)^2, as the Java loop
        best[s:s + 8192] = d2.argmin(1)                                 # first minimum wins (dist < minDist)
        obj += d2.min(1).sum()
    sums = np.zeros((k, d))
    np.add.at(sums, best, points)
    counts = np.bincount(best, minlength=k).astype(np.float64)
    new = np.zeros_like(centroids)
    nz = counts > 0
    new[nz] = sums[nz] / counts[nz, None]
    return new, counts, obj


def run(points: np.ndarray, centroids: np.ndarray, iterations: int):
    hist = []
    for _ in range(iterations):
        centroids, counts, obj = lloyd_iteration(points, centroids)
        hist.append(obj)
    return centroids, hist
