"""On-disk formats of the reference, read and written bit-compatibly.

===============================  ============================================================
K-means (ml/java) points         raw big-endian IEEE doubles, ``pointsPerFile x dim``, no header
                                 (``MJ/kmeans/regroupallgather/DataGenRunnable.java:47-58``,
                                 reader ``PointLoadTask.java:76-96``)
K-means centroids                text, one centroid per line, space separated
                                 (``KMUtil.java:59-97``, ``:213-251``)
contrib K-means                  text points / centroids, ``evaluation`` = ``MSE : <v>``
                                 (``contrib/.../kmeans/common/Utils.java:26-107``)
MF ratings                       text ``row col value`` (``MJ/sgd/VStore.java:93-111``)
MF model                         ``id : v0 v1 ...`` (``MJ/sgd/SGDCollectiveMapper.java:737-817``)
LDA documents                    ``docName wid wid ...`` (``MJ/lda/DocStore.java:157-168``)
LDA word model                   ``wordId topic:count ...`` (``LDAMPCollectiveMapper.java:593-628``)
Harp-DAAL dense CSV              comma separated doubles (``HD/datasource/HarpDAALDataSource.java:76-151``)
Harp-DAAL CSR                    three lines: row offsets / column indices / values, 1-based
                                 (``HarpDAALDataSource.java:498-560``)
libsvm                           ``label idx:val ...`` (``contrib/.../svm/SVMMapper.java:78-97``)
graph / template                 ``v\\tn1,n2,...,`` and ``#V\\n#E\\nu v`` (``website .../sahad.md:77-84``)
===============================  ============================================================

Bulk numeric parsing goes through the multi-threaded C++ parser in ``csrc/host/parse.cpp`` when the
host library is built; a NumPy fallback keeps everything working without it.
"""
from __future__ import annotations

import ctypes
import os
from typing import Dict, Iterable, List, Optional, Sequence, Tuple

import numpy as np
import torch

from .. import _native


# ---------------------------------------------------------------------------------------
# fast text → float64 parsing (C++ host library, NumPy fallback)
# ---------------------------------------------------------------------------------------
def _parse_numbers(data: bytes, threads: int = 0) -> np.ndarray:
    """All numbers of an ASCII buffer (separators: blank, tab, comma, newline, ':') as float64."""
    if _native.have_host_lib():
        lib = _native.host()
        fn = lib.hh_parse_doubles
        fn.restype = ctypes.c_longlong
        fn.argtypes = [ctypes.c_char_p, ctypes.c_longlong, ctypes.c_void_p, ctypes.c_longlong, ctypes.c_int]
        cap = len(data) // 2 + 2
        out = np.empty(cap, dtype=np.float64)
        n = fn(data, len(data), out.ctypes.data_as(ctypes.c_void_p), cap, threads)
        if n >= 0:
            return out[:n].copy()
    txt = data.decode("ascii", "ignore").replace(",", " ").replace(":", " ")
    return np.array(txt.split(), dtype=np.float64) if txt.strip() else np.empty(0, dtype=np.float64)


# ---------------------------------------------------------------------------------------
# K-means (ml/java): binary big-endian doubles
# ---------------------------------------------------------------------------------------
def write_kmeans_points_bin(path: str, points: np.ndarray) -> None:
    np.asarray(points, dtype=">f8").tofile(path)


def read_kmeans_points_bin(path: str, dim: int) -> np.ndarray:
    raw = np.fromfile(path, dtype=">f8")
    n = raw.size // dim
    return raw[: n * dim].astype(np.float64).reshape(n, dim)


def write_centroids_text(path: str, centroids: np.ndarray, fmt: str = "%s") -> None:
    """One centroid per line, space separated (``KMUtil.generateCentroids`` / ``storeCentroids``)."""
    os.makedirs(os.path.dirname(path) or ".", exist_ok=True)
    with open(path, "w") as f:
        for row in np.asarray(centroids):
            f.write(" ".join(repr(float(v)) for v in row))
            f.write("\n")


def read_centroids_text(path: str, dim: Optional[int] = None) -> np.ndarray:
    rows = []
    with open(path, "rb") as f:
        for line in f:
            vals = _parse_numbers(line)
            if vals.size:
                rows.append(vals if dim is None else vals[:dim])
    return np.stack(rows) if rows else np.empty((0, dim or 0))


# ---------------------------------------------------------------------------------------
# contrib K-means: text points, evaluation file
# ---------------------------------------------------------------------------------------
def write_points_text(path: str, points: np.ndarray, sep: str = " ") -> None:
    os.makedirs(os.path.dirname(path) or ".", exist_ok=True)
    with open(path, "w") as f:
        for row in np.asarray(points):
            f.write(sep.join(repr(float(v)) for v in row))
            f.write("\n")


def read_points_text(path: str, dim: Optional[int] = None) -> np.ndarray:
    with open(path, "rb") as f:
        data = f.read()
    vals = _parse_numbers(data)
    if dim is None:
        first = data.split(b"\n", 1)[0]
        dim = _parse_numbers(first).size
    if dim == 0:
        return np.empty((0, 0))
    return vals[: (vals.size // dim) * dim].reshape(-1, dim)


def write_evaluation(path: str, name: str, value: float) -> None:
    """``<work>/evaluation`` = ``MSE : <v>`` (``contrib/.../kmeans/allreduce/KmeansMapper.java:240-249``)."""
    os.makedirs(os.path.dirname(path) or ".", exist_ok=True)
    with open(path, "w") as f:
        f.write(f"{name} : {value}\n")


def read_evaluation(path: str) -> Tuple[str, float]:
    with open(path) as f:
        name, _, val = f.read().strip().partition(":")
    return name.strip(), float(val)


# ---------------------------------------------------------------------------------------
# dense CSV (Harp-DAAL)
# ---------------------------------------------------------------------------------------
def read_dense_csv(path: str, sep: str = ",") -> np.ndarray:
    with open(path, "rb") as f:
        data = f.read()
    first = data.split(b"\n", 1)[0]
    dim = len([t for t in first.replace(sep.encode(), b" ").split() if t])
    vals = _parse_numbers(data)
    if dim == 0:
        return np.empty((0, 0))
    return vals[: (vals.size // dim) * dim].reshape(-1, dim)


def write_dense_csv(path: str, mat: np.ndarray, sep: str = ",") -> None:
    os.makedirs(os.path.dirname(path) or ".", exist_ok=True)
    np.savetxt(path, np.asarray(mat), delimiter=sep, fmt="%.17g")


def read_dense_csv_files(paths: Sequence[str], sep: str = ",") -> np.ndarray:
    mats = [read_dense_csv(p, sep) for p in paths]
    mats = [m for m in mats if m.size]
    return np.concatenate(mats, axis=0) if mats else np.empty((0, 0))


# ---------------------------------------------------------------------------------------
# CSR (DAAL 3-line, 1-based)
# ---------------------------------------------------------------------------------------
def read_csr(path: str) -> Tuple[np.ndarray, np.ndarray, np.ndarray]:
    """Returns zero-based ``(row_offsets[int64], col_indices[int64], values[float64])``."""
    with open(path, "rb") as f:
        lines = [ln for ln in f.read().split(b"\n") if ln.strip()]
    assert len(lines) >= 3, f"{path}: DAAL CSR needs 3 lines"
    ro = _parse_numbers(lines[0]).astype(np.int64) - 1
    ci = _parse_numbers(lines[1]).astype(np.int64)
    if ci.size == 0 or ci.min() >= 1:      # one-based, as DAAL writes them; a 0 marks a zero-based file (the
        ci = ci - 1                        # reference's own datasets/daal_kernelfunc/csrbatch sample is one)
    va = _parse_numbers(lines[2])
    return ro, ci, va


def write_csr(path: str, row_offsets: np.ndarray, col_indices: np.ndarray, values: np.ndarray) -> None:
    os.makedirs(os.path.dirname(path) or ".", exist_ok=True)
    with open(path, "w") as f:
        f.write(",".join(str(int(v) + 1) for v in row_offsets) + "\n")
        f.write(",".join(str(int(v) + 1) for v in col_indices) + "\n")
        f.write(",".join(repr(float(v)) for v in values) + "\n")


def csr_to_torch(ro, ci, va, ncols: Optional[int] = None, device="cpu", dtype=torch.float32) -> torch.Tensor:
    nrows = len(ro) - 1
    ncols = int(ci.max()) + 1 if ncols is None and len(ci) else (ncols or 0)
    return torch.sparse_csr_tensor(torch.as_tensor(ro), torch.as_tensor(ci), torch.as_tensor(va, dtype=dtype),
                                   size=(nrows, ncols), device=device)


# ---------------------------------------------------------------------------------------
# COO triples: MF ratings ``row col value`` (also ALS, blank or tab separated)
# ---------------------------------------------------------------------------------------
def read_coo(paths: Sequence[str]) -> Tuple[np.ndarray, np.ndarray, np.ndarray]:
    rows, cols, vals = [], [], []
    for p in ([paths] if isinstance(paths, str) else paths):
        with open(p, "rb") as f:
            v = _parse_numbers(f.read())
        v = v[: (v.size // 3) * 3].reshape(-1, 3)
        rows.append(v[:, 0].astype(np.int64)); cols.append(v[:, 1].astype(np.int64)); vals.append(v[:, 2])
    if not rows:
        return np.empty(0, np.int64), np.empty(0, np.int64), np.empty(0)
    return np.concatenate(rows), np.concatenate(cols), np.concatenate(vals)


def write_coo(path: str, rows, cols, vals) -> None:
    os.makedirs(os.path.dirname(path) or ".", exist_ok=True)
    with open(path, "w") as f:
        for r, c, v in zip(rows, cols, vals):
            f.write(f"{int(r)} {int(c)} {float(v)!r}\n")


def write_mf_model(path: str, ids: Iterable[int], mat: np.ndarray) -> None:
    """``id : v0 v1 ... v(r-1)`` per line (``MJ/sgd/SGDCollectiveMapper.java:737-817``)."""
    os.makedirs(os.path.dirname(path) or ".", exist_ok=True)
    with open(path, "w") as f:
        for i, row in zip(ids, np.asarray(mat)):
            f.write(f"{int(i)} : " + " ".join(repr(float(v)) for v in row) + "\n")


def read_mf_model(path: str) -> Tuple[np.ndarray, np.ndarray]:
    ids, rows = [], []
    with open(path) as f:
        for line in f:
            if ":" not in line:
                continue
            a, b = line.split(":", 1)
            ids.append(int(a)); rows.append([float(x) for x in b.split()])
    return np.array(ids, dtype=np.int64), np.array(rows, dtype=np.float64)


# ---------------------------------------------------------------------------------------
# LDA
# ---------------------------------------------------------------------------------------
def read_lda_docs(paths: Sequence[str]) -> Tuple[List[str], List[np.ndarray]]:
    """``docName wid wid ...`` — repeated ids are repeated tokens (``MJ/lda/DocStore.java:157-168``)."""
    names, docs = [], []
    for p in ([paths] if isinstance(paths, str) else paths):
        with open(p) as f:
            for line in f:
                toks = line.split()
                if len(toks) < 2:
                    continue
                names.append(toks[0])
                docs.append(np.array(toks[1:], dtype=np.int64))
    return names, docs


def write_lda_docs(path: str, names: Sequence[str], docs: Sequence[np.ndarray]) -> None:
    os.makedirs(os.path.dirname(path) or ".", exist_ok=True)
    with open(path, "w") as f:
        for nm, d in zip(names, docs):
            f.write(nm + " " + " ".join(str(int(w)) for w in d) + "\n")


def write_lda_word_model(path: str, word_ids: Iterable[int], word_topic: np.ndarray) -> None:
    """``wordId topic:count topic:count ...`` for non-zero counts (``LDAMPCollectiveMapper.java:593-628``)."""
    os.makedirs(os.path.dirname(path) or ".", exist_ok=True)
    with open(path, "w") as f:
        for w, row in zip(word_ids, np.asarray(word_topic)):
            nz = np.nonzero(row)[0]
            if nz.size:
                f.write(f"{int(w)} " + " ".join(f"{int(t)}:{int(row[t])}" for t in nz) + "\n")


def read_lda_word_model(path: str, num_topics: int) -> Dict[int, np.ndarray]:
    out: Dict[int, np.ndarray] = {}
    with open(path) as f:
        for line in f:
            toks = line.split()
            if not toks:
                continue
            row = np.zeros(num_topics, dtype=np.int64)
            for tc in toks[1:]:
                t, c = tc.split(":")
                row[int(t)] = int(c)
            out[int(toks[0])] = row
    return out


# ---------------------------------------------------------------------------------------
# libsvm
# ---------------------------------------------------------------------------------------
def read_libsvm(path: str, num_features: Optional[int] = None) -> Tuple[np.ndarray, np.ndarray]:
    """Dense ``(X[n, f], y[n])`` from ``label idx:val ...`` (1-based indices)."""
    labels, rows = [], []
    maxf = 0
    with open(path) as f:
        for line in f:
            toks = line.split()
            if not toks:
                continue
            labels.append(float(toks[0]))
            feats = []
            for t in toks[1:]:
                i, v = t.split(":")
                feats.append((int(i), float(v)))
                maxf = max(maxf, int(i))
            rows.append(feats)
    nf = num_features or maxf
    X = np.zeros((len(rows), nf), dtype=np.float64)
    for r, feats in enumerate(rows):
        for i, v in feats:
            if i <= nf:
                X[r, i - 1] = v
    return X, np.array(labels)


def write_libsvm(path: str, X: np.ndarray, y: np.ndarray) -> None:
    os.makedirs(os.path.dirname(path) or ".", exist_ok=True)
    with open(path, "w") as f:
        for row, lab in zip(np.asarray(X), np.asarray(y)):
            nz = np.nonzero(row)[0]
            f.write(f"{lab:g} " + " ".join(f"{i + 1}:{row[i]:.17g}" for i in nz) + "\n")


# ---------------------------------------------------------------------------------------
# graphs / templates (subgraph counting) and edge lists (PageRank)
# ---------------------------------------------------------------------------------------
def read_adjacency(path: str) -> Dict[int, np.ndarray]:
    """``v\\tn1,n2,...,`` per line."""
    g: Dict[int, np.ndarray] = {}
    with open(path) as f:
        for line in f:
            line = line.strip()
            if not line:
                continue
            a, _, b = line.partition("\t")
            if not b:
                parts = line.split()
                a, b = parts[0], ",".join(parts[1:])
            nb = [int(x) for x in b.replace(" ", ",").split(",") if x]
            g[int(a)] = np.array(nb, dtype=np.int64)
    return g


def write_adjacency(path: str, g: Dict[int, Iterable[int]]) -> None:
    os.makedirs(os.path.dirname(path) or ".", exist_ok=True)
    with open(path, "w") as f:
        for v in sorted(g):
            f.write(f"{v}\t" + "".join(f"{int(n)}," for n in g[v]) + "\n")


def read_template(path: str) -> Tuple[int, List[Tuple[int, int]]]:
    """``#V`` then ``#E`` then one ``u v`` edge per line (``datasets/daal_subgraph/templates/*.template``)."""
    with open(path) as f:
        toks = [ln.strip() for ln in f if ln.strip()]
    nv, ne = int(toks[0]), int(toks[1])
    edges = [tuple(int(x) for x in ln.split()[:2]) for ln in toks[2:2 + ne]]
    return nv, edges


def write_template(path: str, nv: int, edges: Sequence[Tuple[int, int]]) -> None:
    os.makedirs(os.path.dirname(path) or ".", exist_ok=True)
    with open(path, "w") as f:
        f.write(f"{nv}\n{len(edges)}\n")
        for u, v in edges:
            f.write(f"{u} {v}\n")


def read_pagerank_graph(path: str) -> Dict[int, np.ndarray]:
    """``src dst dst ...`` adjacency lines (``datasets/tutorial/simplepagerank``)."""
    g: Dict[int, np.ndarray] = {}
    with open(path) as f:
        for line in f:
            toks = line.replace(",", " ").split()
            if toks:
                g[int(toks[0])] = np.array([int(x) for x in toks[1:]], dtype=np.int64)
    return g


# ---------------------------------------------------------------------------------------
# rendezvous files (kept for tooling parity; torchrun replaces their function)
# ---------------------------------------------------------------------------------------
def write_rendezvous(job_dir: str, hosts: Sequence[str]) -> None:
    """``<jobID>/{nodes,tasks,lock}`` (``MapCollectiveContainerLauncherImpl.java:293-355``)."""
    os.makedirs(job_dir, exist_ok=True)
    with open(os.path.join(job_dir, "nodes"), "w") as f:
        f.write("#0\n")
        for h in hosts:
            f.write(h + "\n")
    with open(os.path.join(job_dir, "tasks"), "w") as f:
        for i in range(len(hosts)):
            f.write(f"{i}\t{i}\n")
    open(os.path.join(job_dir, "lock"), "w").close()


def read_nodes_file(path: str) -> Dict[int, List[str]]:
    """``#rackID`` lines followed by one host per worker (``HC/worker/Nodes.java:89-118``)."""
    racks: Dict[int, List[str]] = {}
    cur = 0
    with open(path) as f:
        for line in f:
            line = line.strip()
            if not line:
                continue
            if line.startswith("#"):
                cur = int(line[1:] or 0)
                racks.setdefault(cur, [])
            else:
                racks.setdefault(cur, []).append(line)
    return racks
