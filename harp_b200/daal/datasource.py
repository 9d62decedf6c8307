"""``HarpDAALDataSource`` (``core/harp-daal-interface/src/main/java/edu/iu/datasource``): multi-threaded
loaders for dense CSV, CSV split into features/labels, COO triples → grouped → remapped → regrouped →
CSR, and DAAL 3-line CSR files.  Parsing runs in the C++ host library (``csrc/host/parse.cpp``); file
fan-out uses the host ``DynamicScheduler`` like the reference's ``MTReader`` (``MTReader.java:77-285``).
"""
from __future__ import annotations

from typing import Dict, Optional, Sequence, Tuple

import numpy as np
import torch

from ..io import formats as F
from ..partition import FnCombiner, Partition, Partitioner, Table
from ..sched import DynamicScheduler, Task


class _ReadTask(Task):
    def __init__(self, sep: str):
        self.sep = sep

    def run(self, path: str):
        return (path, F.read_dense_csv(path, self.sep))


class HarpDAALDataSource:
    def __init__(self, num_threads: int = 4, conf: Optional[dict] = None, device="cpu", dtype=torch.float32):
        self.num_threads = max(1, int(num_threads))
        self.conf = conf or {}
        self.device = torch.device(device)
        self.dtype = dtype if self.device.type == "cuda" else torch.float64

    # ---- dense -----------------------------------------------------------------------
    def load_dense_csv_files(self, files: Sequence[str], sep: str = ",") -> np.ndarray:
        """``loadDenseCSVFiles`` (``HarpDAALDataSource.java:126-151``): files are parsed by a pool of reader
        tasks, rows keep file order."""
        files = list(files)
        if not files:
            return np.empty((0, 0))
        sched = DynamicScheduler([_ReadTask(sep) for _ in range(min(self.num_threads, len(files)))])
        sched.submit_all(files)
        sched.start()
        got: Dict[str, np.ndarray] = {}
        while sched.has_output():
            out = sched.wait_for_output()
            if out is not None:
                got[out[0]] = out[1]
        sched.stop()
        mats = [got[f] for f in files if got[f].size]
        return np.concatenate(mats, 0) if mats else np.empty((0, 0))

    def create_dense_numeric_table(self, files: Sequence[str], n_features: Optional[int] = None, sep: str = ",") -> torch.Tensor:
        """``createDenseNumericTable`` (``:76-104``) → a device tensor instead of a HomogenNumericTable."""
        m = self.load_dense_csv_files(files, sep)
        if n_features is not None and m.size:
            m = m[:, :n_features]
        return torch.from_numpy(np.ascontiguousarray(m)).to(self.device, self.dtype)

    def create_dense_numeric_table_split(self, files: Sequence[str], n_features: int, n_labels: int, sep: str = ","):
        """``createDenseNumericTableSplit``: leading ``n_features`` columns and trailing ``n_labels`` columns."""
        m = self.load_dense_csv_files(files, sep)
        X = torch.from_numpy(np.ascontiguousarray(m[:, :n_features])).to(self.device, self.dtype)
        Y = torch.from_numpy(np.ascontiguousarray(m[:, n_features:n_features + n_labels])).to(self.device, self.dtype)
        return X, Y

    def load_dense_csv_files_sharding(self, files: Sequence[str], rank: int, world: int, sep: str = ","):
        """``loadDenseCSVFilesSharding``: every worker reads the whole list but keeps rows ``i % world == rank``."""
        m = self.load_dense_csv_files(files, sep)
        return torch.from_numpy(np.ascontiguousarray(m[rank::world])).to(self.device, self.dtype)

    # ---- COO → CSR pipeline (``:317-496``) ----------------------------------------------
    def load_coo_files(self, files: Sequence[str]) -> torch.Tensor:
        r, c, v = F.read_coo(list(files)) if files else (np.empty(0, np.int64), np.empty(0, np.int64), np.empty(0))
        return torch.from_numpy(np.stack([r.astype(np.float64), c.astype(np.float64), v], 1)) if r.size else torch.zeros(0, 3, dtype=torch.float64)

    @staticmethod
    def group_coo_by_ids(coo: torch.Tensor, by_row: bool = True) -> Dict[int, torch.Tensor]:
        """``groupCOOByIDs``: id → its (other id, value) pairs."""
        key = coo[:, 0 if by_row else 1].long()
        out: Dict[int, torch.Tensor] = {}
        order = torch.argsort(key, stable=True)
        ks, cnt = torch.unique_consecutive(key[order], return_counts=True)
        off = 0
        for k, n in zip(ks.tolist(), cnt.tolist()):
            rows = order[off:off + n]
            out[k] = coo[rows][:, [1 if by_row else 0, 2]]
            off += n
        return out

    @staticmethod
    def remap_coo_ids(coo: torch.Tensor, mapper, by_row: bool = True) -> Tuple[torch.Tensor, Dict[int, int]]:
        """``remapCOOIDs``: all workers agree on a dense 0..M-1 numbering of the ids that occur anywhere
        (allgather of the local id sets, then rank in the sorted union)."""
        col = 0 if by_row else 1
        local = torch.unique(coo[:, col].long()).tolist()
        every = sorted({i for ids in mapper.comm.all_gather_object(local) for i in ids})
        remap = {g: j for j, g in enumerate(every)}
        lut = torch.tensor(every, dtype=torch.long)
        new = coo.clone()
        new[:, col] = torch.searchsorted(lut, coo[:, col].long()).double()
        return new, remap

    @staticmethod
    def regroup_coo_list(coo: torch.Tensor, mapper, by_row: bool = True) -> torch.Tensor:
        """``regroupCOOList`` with ``COORegroupPartitioner``: triples travel to worker ``id % N``."""
        N, me = mapper.get_num_workers(), mapper.get_self_id()
        key = coo[:, 0 if by_row else 1].long()
        t = Table(0, FnCombiner(lambda a, b: torch.cat([a, b], 0)))
        for w in range(N):
            part = coo[key % N == w]
            if part.numel():
                t.add_partition(Partition(w, part))
        mapper.regroup("daal", "regroup-coo", t, Partitioner(N))
        p = t.get_partition(me)
        return p.get() if p is not None else coo.new_zeros((0, 3))

    def coo_to_csr(self, coo: torch.Tensor, n_rows: int, n_cols: int, by_row: bool = True) -> torch.Tensor:
        """``COOToCSR`` → torch sparse CSR tensor (DAAL CSR is 1-based on disk, 0-based here)."""
        r = coo[:, 0 if by_row else 1].long()
        c = coo[:, 1 if by_row else 0].long()
        order = torch.argsort(r * max(n_cols, 1) + c, stable=True)
        r, c, v = r[order], c[order], coo[order, 2]
        ptr = torch.zeros(n_rows + 1, dtype=torch.long)
        ptr[1:] = torch.cumsum(torch.bincount(r, minlength=n_rows), 0)
        return torch.sparse_csr_tensor(ptr, c, v.to(self.dtype), size=(n_rows, n_cols)).to(self.device)

    # ---- DAAL CSR files (``:498-560``) ---------------------------------------------------
    def load_csr_numeric_table(self, files: Sequence[str], n_cols: Optional[int] = None) -> torch.Tensor:
        ptrs, idxs, vals = [np.zeros(1, np.int64)], [], []
        for f in files:
            ro, ci, va = F.read_csr(f)
            ptrs.append(ro[1:] + ptrs[-1][-1])
            idxs.append(ci); vals.append(va)
        ptr = np.concatenate(ptrs)
        ci = np.concatenate(idxs) if idxs else np.empty(0, np.int64)
        va = np.concatenate(vals) if vals else np.empty(0)
        nc = n_cols if n_cols is not None else (int(ci.max()) + 1 if ci.size else 0)
        return torch.sparse_csr_tensor(torch.from_numpy(ptr), torch.from_numpy(ci),
                                       torch.from_numpy(va).to(self.dtype), size=(len(ptr) - 1, nc)).to(self.device)

    def load_csr_numeric_table_and_label(self, files: Sequence[str], label_files: Sequence[str], n_cols: Optional[int] = None):
        X = self.load_csr_numeric_table(files, n_cols)
        y = self.load_dense_csv_files(label_files)
        return X, torch.from_numpy(np.ascontiguousarray(y)).to(self.device)

    # camelCase aliases of the reference API
    createDenseNumericTable = create_dense_numeric_table
    createDenseNumericTableSplit = create_dense_numeric_table_split
    loadDenseCSVFiles = load_dense_csv_files
    loadDenseCSVFilesSharding = load_dense_csv_files_sharding
    loadCOOFiles = load_coo_files
    groupCOOByIDs = group_coo_by_ids
    remapCOOIDs = remap_coo_ids
    regroupCOOList = regroup_coo_list
    COOToCSR = coo_to_csr
    loadCSRNumericTable = load_csr_numeric_table
    loadCSRNumericTableAndLabel = load_csr_numeric_table_and_label
