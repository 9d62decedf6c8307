"""Synthetic data generators of the reference launchers.

* ``generate_kmeans_data``: ``KMUtil.generateData`` / ``DataGenRunnable`` — ``numOfDataPoints`` points
  in ``[0, 1000)^dim`` split over ``numPointFiles`` binary big-endian files ``data_<k>``
  (``ml/java/.../kmeans/regroupallgather/KMUtil.java:99-178``, ``DataGenRunnable.java:47-58``).
* ``generate_centroids``: ``KMUtil.generateCentroids`` — text file of ``numCentroids`` random rows
  (``KMUtil.java:59-97``).
* ``generate_dense_csv``: ``DataGenerator.generateDenseDataMulti`` (``HD/data_gen/DataGenerator.java:56-249``).
* ratings / documents generators for the MF and LDA benchmarks (synthetic, shaped like Netflix /
  a Zipfian corpus) — the reference downloads real datasets instead.
"""
from __future__ import annotations

import os
from typing import List, Tuple

import numpy as np

from . import formats as F


def generate_kmeans_data(num_points: int, dim: int, num_files: int, out_dir: str, seed: int = 0,
                         value_range: float = 1000.0) -> List[str]:
    os.makedirs(out_dir, exist_ok=True)
    rng = np.random.default_rng(seed)
    per = num_points // num_files
    paths = []
    for k in range(num_files):
        n = per + (num_points - per * num_files if k == num_files - 1 else 0)
        pts = rng.random((n, dim)) * value_range
        p = os.path.join(out_dir, f"data_{k}")
        F.write_kmeans_points_bin(p, pts)
        paths.append(p)
    return paths


def generate_centroids(num_centroids: int, dim: int, path: str, seed: int = 1, value_range: float = 1000.0) -> np.ndarray:
    rng = np.random.default_rng(seed)
    c = rng.random((num_centroids, dim)) * value_range
    F.write_centroids_text(path, c)
    return c


def generate_points_text(num_points: int, dim: int, num_files: int, out_dir: str, seed: int = 0,
                         value_range: float = 10.0) -> List[str]:
    """contrib K-means text points (``contrib/.../kmeans/common/Utils.java:26-74``)."""
    os.makedirs(out_dir, exist_ok=True)
    rng = np.random.default_rng(seed)
    per = num_points // num_files
    paths = []
    for k in range(num_files):
        n = per + (num_points - per * num_files if k == num_files - 1 else 0)
        p = os.path.join(out_dir, f"data_{k}")
        F.write_points_text(p, rng.random((n, dim)) * value_range)
        paths.append(p)
    return paths


def generate_dense_csv(num_rows: int, dim: int, num_files: int, out_dir: str, seed: int = 0,
                       label_fn=None, scale: float = 1.0) -> List[str]:
    os.makedirs(out_dir, exist_ok=True)
    rng = np.random.default_rng(seed)
    per = num_rows // num_files
    paths = []
    for k in range(num_files):
        n = per + (num_rows - per * num_files if k == num_files - 1 else 0)
        x = rng.standard_normal((n, dim)) * scale
        if label_fn is not None:
            x = np.concatenate([x, np.asarray(label_fn(x)).reshape(n, -1)], axis=1)
        p = os.path.join(out_dir, f"part_{k}.csv")
        F.write_dense_csv(p, x)
        paths.append(p)
    return paths


def synth_ratings(num_rows: int, num_cols: int, nnz: int, rank: int = 8, seed: int = 0,
                  noise: float = 0.1) -> Tuple[np.ndarray, np.ndarray, np.ndarray]:
    """Low-rank-plus-noise ratings with Zipf-ish row/column popularity."""
    rng = np.random.default_rng(seed)
    w = rng.standard_normal((num_rows, rank)) / np.sqrt(rank)
    h = rng.standard_normal((num_cols, rank)) / np.sqrt(rank)
    r = (rng.random(nnz) ** 2 * num_rows).astype(np.int64) % num_rows
    c = (rng.random(nnz) ** 2 * num_cols).astype(np.int64) % num_cols
    v = np.einsum("ij,ij->i", w[r], h[c]) * 2.0 + 3.0 + noise * rng.standard_normal(nnz)
    return r, c, v


def generate_ratings_files(num_rows: int, num_cols: int, nnz: int, num_files: int, out_dir: str,
                           seed: int = 0) -> List[str]:
    os.makedirs(out_dir, exist_ok=True)
    r, c, v = synth_ratings(num_rows, num_cols, nnz, seed=seed)
    paths = []
    for k, idx in enumerate(np.array_split(np.arange(nnz), num_files)):
        p = os.path.join(out_dir, f"ratings_{k}")
        F.write_coo(p, r[idx], c[idx], v[idx])
        paths.append(p)
    return paths


def synth_corpus(num_docs: int, vocab: int, num_topics: int, avg_len: int, seed: int = 0,
                 alpha: float = 0.1, beta: float = 0.01) -> List[np.ndarray]:
    """Documents drawn from an LDA generative model (so that CGS has structure to find)."""
    rng = np.random.default_rng(seed)
    phi = rng.dirichlet(np.full(vocab, beta), size=num_topics)
    docs = []
    for _ in range(num_docs):
        n = max(1, int(rng.poisson(avg_len)))
        theta = rng.dirichlet(np.full(num_topics, alpha))
        z = rng.choice(num_topics, size=n, p=theta)
        words = np.array([rng.choice(vocab, p=phi[t]) for t in z], dtype=np.int64)
        docs.append(words)
    return docs


def generate_corpus_files(num_docs: int, vocab: int, num_topics: int, avg_len: int, num_files: int,
                          out_dir: str, seed: int = 0) -> List[str]:
    os.makedirs(out_dir, exist_ok=True)
    docs = synth_corpus(num_docs, vocab, num_topics, avg_len, seed)
    paths = []
    for k, idx in enumerate(np.array_split(np.arange(num_docs), num_files)):
        p = os.path.join(out_dir, f"docs_{k}")
        F.write_lda_docs(p, [f"doc{i}" for i in idx], [docs[i] for i in idx])
        paths.append(p)
    return paths
