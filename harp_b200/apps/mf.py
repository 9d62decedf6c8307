"""Matrix-factorisation applications: MF-SGD (``edu.iu.sgd``) and MF-CCD (``edu.iu.ccd``).

SGD (``ml/java/src/main/java/edu/iu/sgd/SGDCollectiveMapper.java:154-320``): ratings are regrouped so
that every row lives on one worker (``RandomPartitioner``), W rows stay put, H columns rotate in two
slices; every ``5`` iterations the test RMSE is computed with one extra rotation; at the end
``W-<rank>``, ``H-<rank>`` and ``evaluation`` are written (``:737-817``).  ``<training ratio>`` keeps
its meaning: the fraction (percent) of a panel's ratings a timer-bounded pass should train; after the
first iteration the time budget is re-tuned from what was actually trained (``adjustMiniBatch``,
``:623-668``).

CCD (``ml/java/src/main/java/edu/iu/ccd/CCDMPCollectiveMapper.java``): CCD++ — features are updated
one at a time as rank-one subproblems with residuals kept per rating (``CCDMPTask.java:67-125``).
The reference rotates both factor matrices feature-row by feature-row; here the current feature's
two vectors are allgathered instead (they are only ``rows + cols`` floats), each worker updating
the rows *and* the columns it owns from a row-partitioned and a column-partitioned copy of the
ratings.
"""
from __future__ import annotations

import math
import os
import time
from typing import List, Optional, Tuple

import numpy as np
import torch
import torch.distributed as dist

from ..io import formats as F
from ..mapper import CollectiveMapper, Context, KeyValReader, LOG
from ..models.mfsgd import MFSGDEngine
from ..ops import mf
from ..partition import DoubleArrMax, DoubleArrPlus, FnCombiner, Partition, Partitioner, RandomPartitioner, Table


def _cat_combiner():
    """``VSetCombiner`` (``ml/java/.../sgd/VSetCombiner.java:25-30``): rating lists append."""
    return FnCombiner(lambda a, b: torch.cat([a, b.to(a.device)], 0))


def _regroup_by_owner(mapper: CollectiveMapper, ctx_name: str, op: str, triples: torch.Tensor, owners: torch.Tensor) -> torch.Tensor:
    """Sends each (row, col, val) triple to its owner with ``regroup``; returns the triples this worker owns."""
    N = mapper.get_num_workers()
    table = Table(0, _cat_combiner())
    for w in range(N):
        part = triples[owners == w]
        if part.numel():
            table.add_partition(Partition(w, part))
    mapper.regroup(ctx_name, op, table, Partitioner(N))
    p = table.get_partition(mapper.get_self_id())
    return p.get() if p is not None else triples.new_zeros((0, 3))


def _load_ratings(files: List[str], device) -> torch.Tensor:
    if not files:
        return torch.zeros(0, 3, dtype=torch.float64, device=device)
    r, c, v = F.read_coo(files)
    return torch.from_numpy(np.stack([r.astype(np.float64), c.astype(np.float64), v], 1)).to(device)


class SGDCollectiveMapper(CollectiveMapper):
    """conf: ``r``, ``lambda``, ``epsilon``, ``num_iterations``, ``train_ratio``, ``model_dir``,
    ``test_files`` (list), optional ``seed``, ``blocks``, ``rmse_interval`` (default 5)."""

    def setup(self, context: Context) -> None:
        c = context.get_configuration()
        self.r = int(c["r"])
        self.lam = float(c["lambda"])
        self.eps = float(c["epsilon"])
        self.num_iterations = int(c["num_iterations"])
        self.train_ratio = int(c.get("train_ratio", 100))
        self.model_dir = c.get("model_dir")
        self.test_files = list(c.get("test_files", []))
        self.seed = int(c.get("seed", 0))
        self.blocks = c.get("blocks")
        self.rmse_interval = int(c.get("rmse_interval", 5))

    def map_collective(self, reader: KeyValReader, context: Context) -> None:
        files = []
        while reader.next_key_value():
            files.append(reader.get_current_value())
        self.run_sgd(files, context)

    def run_sgd(self, files: List[str], context: Context) -> None:
        N, me, dev = self.get_num_workers(), self.get_self_id(), self.device
        raw = _load_ratings(files, dev)
        # global maxima and a shared seed (allreduce with LongArrMax, SGDCollectiveMapper.java:373)
        t = Table(0, DoubleArrMax())
        mx = torch.tensor([raw[:, 0].max().item() if raw.numel() else -1.0,
                           raw[:, 1].max().item() if raw.numel() else -1.0, float(self.seed)],
                          dtype=torch.float64, device=dev)
        t.add_partition(Partition(0, mx))
        self.allreduce("sgd", "allreduce-max-ids", t)
        max_row, max_col, seed = (int(x) for x in t.get_partition(0).get().tolist())
        part = RandomPartitioner(N, seed)
        mine = _regroup_by_owner(self, "sgd", "regroup-rows", raw, part.worker_ids(raw[:, 0].long()))
        # compact local row ids
        self.row_ids = torch.unique(mine[:, 0].long())
        lrow = torch.searchsorted(self.row_ids, mine[:, 0].long())
        num_cols = max_col + 1
        # test ratings: same owner rule; rows never seen in training are dropped (reference: skipped)
        test_files = [f for i, f in enumerate(sorted(self.test_files)) if i % N == me]
        traw = _load_ratings(test_files, dev)
        tmine = _regroup_by_owner(self, "sgd", "regroup-test", traw, part.worker_ids(traw[:, 0].long()))
        if tmine.numel() and self.row_ids.numel():
            pos = torch.searchsorted(self.row_ids, tmine[:, 0].long()).clamp(max=self.row_ids.numel() - 1)
            ok = (self.row_ids[pos] == tmine[:, 0].long()) & (tmine[:, 1] < num_cols)
            tmine, tpos = tmine[ok], pos[ok]
        else:
            tmine, tpos = tmine[:0], torch.zeros(0, dtype=torch.long, device=dev)
        wdt = torch.float32 if dev.type == "cuda" else torch.float64
        self.engine = MFSGDEngine(lrow, mine[:, 1].long(), mine[:, 2].to(wdt), self.row_ids.numel(), num_cols,
                                  self.r, self.eps, self.lam, symm=self.symm, blocks=int(self.blocks) if self.blocks else None,
                                  seed=seed)
        eng = self.engine
        deadline_ms: Optional[float] = None
        self.rmse_history: List[Tuple[int, float]] = []
        self.trained_history: List[int] = []
        t_compute = 0.0
        for it in range(1, self.num_iterations + 1):
            t0 = time.perf_counter()
            trained = eng.epoch(deadline_ms)
            eng.synchronize()
            dt = time.perf_counter() - t0
            t_compute += dt
            n_tr = int(trained.item())
            self.trained_history.append(n_tr)
            if it == 1 and self.train_ratio < 100:
                # adjustMiniBatch: pick the per-pass budget that trains ~train_ratio % of a panel
                stats = self.comm.all_gather_object((dt, n_tr, eng.nnz))
                rate = sum(s[1] for s in stats) / max(sum(s[0] for s in stats), 1e-9)      # ratings / s / worker
                per_pass = max(s[2] for s in stats) / (eng.S * N)
                deadline_ms = max(0.01, 1e3 * per_pass * self.train_ratio / 100.0 / max(rate, 1.0))
            if it % self.rmse_interval == 0 or it == self.num_iterations:
                rm = eng.rmse(tpos, tmine[:, 1].long(), tmine[:, 2])
                self.rmse_history.append((it, rm))
                if self.is_master():
                    LOG.info("iteration %d test RMSE %.6f (epoch %.3f s)", it, rm, dt)
            context.progress()
        self.compute_time = t_compute
        self.test_rmse = self.rmse_history[-1][1] if self.rmse_history else float("nan")
        if self.model_dir:
            self.save_models()

    def save_models(self) -> None:
        os.makedirs(self.model_dir, exist_ok=True)
        me = self.get_self_id()
        F.write_mf_model(os.path.join(self.model_dir, f"W-{me}"), self.row_ids.cpu().tolist(),
                         self.engine.W.double().cpu().numpy())
        ids, mats = [], []
        for k, (gid, mat) in self.engine.home_H().items():
            ids.append(gid.cpu()); mats.append(mat.double().cpu())
        F.write_mf_model(os.path.join(self.model_dir, f"H-{me}"), torch.cat(ids).tolist(), torch.cat(mats).numpy())
        if self.is_master():
            with open(os.path.join(self.model_dir, "evaluation"), "w") as f:
                f.write(f"{self.test_rmse}\n")


class CCDCollectiveMapper(CollectiveMapper):
    """conf: ``r``, ``lambda``, ``num_iterations`` (outer), ``inner_iterations`` (default 1), ``model_dir``,
    ``test_files``."""

    def setup(self, context: Context) -> None:
        c = context.get_configuration()
        self.r = int(c["r"])
        self.lam = float(c["lambda"])
        self.num_iterations = int(c["num_iterations"])
        self.inner = int(c.get("inner_iterations", 1))
        self.model_dir = c.get("model_dir")
        self.test_files = list(c.get("test_files", []))
        self.seed = int(c.get("seed", 0))

    def map_collective(self, reader: KeyValReader, context: Context) -> None:
        files = []
        while reader.next_key_value():
            files.append(reader.get_current_value())
        N, me, dev = self.get_num_workers(), self.get_self_id(), self.device
        raw = _load_ratings(files, dev)
        t = Table(0, DoubleArrMax())
        t.add_partition(Partition(0, torch.tensor([raw[:, 0].max().item() if raw.numel() else -1.0,
                                                   raw[:, 1].max().item() if raw.numel() else -1.0],
                                                  dtype=torch.float64, device=dev)))
        self.allreduce("ccd", "allreduce-max-ids", t)
        nrows, ncols = (int(x) + 1 for x in t.get_partition(0).get().tolist())
        by_row = _regroup_by_owner(self, "ccd", "regroup-rows", raw, raw[:, 0].long() % N)      # rows i % N == me
        by_col = _regroup_by_owner(self, "ccd", "regroup-cols", raw, raw[:, 1].long() % N)      # cols j % N == me
        # W and H are SHARDED: rank r keeps the rank-one rows of its own ids (i % N == r) only, like the reference's
        # rotating W / H tables (CCDMPCollectiveMapper.java:546-547, :635-636); the full vectors u, v of the feature
        # being updated exist only transiently, gathered over NVLink when the mapper has a symmetric workspace.  One
        # kernel pass per side and feature, residuals local (models/mfccd.py).
        from ..models.mfccd import MFCCDEngine
        symm = self.symm if (dev.type == "cuda" and N > 1) else None
        eng = MFCCDEngine(by_row[:, 0].long(), by_row[:, 1].long(), by_row[:, 2], by_col[:, 0].long(), by_col[:, 1].long(),
                          by_col[:, 2], nrows, ncols, self.r, self.lam, symm=symm, seed=self.seed, inner=self.inner)
        wdt = eng.dt
        self.objective_history = []
        for it in range(self.num_iterations):
            self.objective_history.append(eng.iterate())
            context.progress()
        W, H = eng.full_factors()
        W_loc, H_loc = eng.W_loc, eng.H_loc
        eng.close()
        self.W, self.H, self.W_loc, self.H_loc = W, H, W_loc, H_loc
        # test RMSE
        test_files = [f for i, f in enumerate(sorted(self.test_files)) if i % N == me]
        traw = _load_ratings(test_files, dev)
        if traw.numel():
            ok = (traw[:, 0] < nrows) & (traw[:, 1] < ncols)
            traw = traw[ok]
        pred = (W[:, traw[:, 0].long()] * H[:, traw[:, 1].long()]).sum(0) if traw.numel() else torch.zeros(0, device=dev)
        tt = Table(0, DoubleArrPlus())
        tt.add_partition(Partition(0, torch.stack([((traw[:, 2] - pred.double()) ** 2).sum() if traw.numel() else torch.zeros((), dtype=torch.float64, device=dev),
                                                   torch.tensor(float(traw.shape[0]), dtype=torch.float64, device=dev)])))
        self.allreduce("ccd", "allreduce-test", tt)
        se, cnt = tt.get_partition(0).get().tolist()
        self.test_rmse = math.sqrt(se / cnt) if cnt > 0 else float("nan")
        if self.model_dir and self.is_master():
            os.makedirs(self.model_dir, exist_ok=True)
            F.write_mf_model(os.path.join(self.model_dir, "W-0"), range(nrows), W.t().double().cpu().numpy())
            F.write_mf_model(os.path.join(self.model_dir, "H-0"), range(ncols), H.t().double().cpu().numpy())
            with open(os.path.join(self.model_dir, "evaluation"), "w") as fh:
                fh.write(f"{self.test_rmse}\n")
