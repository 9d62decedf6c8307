"""K-means applications.

* :class:`KMeansRegroupAllgatherMapper` — ``edu.iu.kmeans.regroupallgather`` (README example):
  centroid table of ``numMappers`` partitions, rows ``[count, x_1..x_d]``; per iteration
  local assignment → ``regroup`` → divide by counts → ``allgather``
  (``ml/java/.../kmeans/regroupallgather/KMeansCollectiveMapper.java:87-209``).
* :class:`ContribKMeansMapper` — the four tutorial variants ``allreduce`` | ``regroup-allgather`` |
  ``broadcast-reduce`` | ``push-pull``: one partition per centroid, rows ``[x_1..x_d, count]``,
  *sparse* tables (a centroid nobody is close to disappears), ``evaluation`` = ``MSE : v``
  (``contrib/.../kmeans/{allreduce,regroupallgather,bcastreduce,pushpull}/KmeansMapper.java``).
* :class:`KMeansRotationMapper` — ``edu.iu.kmeans.rotation``: centroids sharded, E-step and M-step
  each take N ``rotate`` calls (``ml/java/.../kmeans/rotation/KMeansCollectiveMapper.java:150-200``).

The hot loop (``CenCalcTask.run``) is ``ops.kmeans.kmeans_step`` — the tcgen05 kernel on a GPU, the
float64 restatement on a CPU.
"""
from __future__ import annotations

import os
import time
from typing import Dict, List, Optional

import numpy as np
import torch

from ..io import formats as F
from .. import _native
from ..mapper import CollectiveMapper, Context, KeyValReader, LOG
from ..ops.kmeans import kmeans_step
from ..partition import DoubleArrPlus, Partition, Partitioner, Table


def _work_dtype(device: torch.device) -> torch.dtype:
    return torch.float32 if device.type == "cuda" else torch.float64


class KMeansRegroupAllgatherMapper(CollectiveMapper):
    """conf keys (``Constants.java:19-38``): ``num_centroids``, ``vector_size``, ``num_iterations``,
    ``centroid_file``, ``out_file``; input records = point file names (binary big-endian doubles)."""

    def setup(self, context: Context) -> None:
        c = context.get_configuration()
        self.num_centroids = int(c["num_centroids"])
        self.vector_size = int(c["vector_size"])
        self.num_iterations = int(c["num_iterations"])
        self.cen_file = c["centroid_file"]
        self.out_file = c.get("out_file")
        self.cen_vec_size = self.vector_size + 1
        # "table": the reference's structure (partial-sum table -> regroup -> divide -> allgather per iteration);
        # "device": the iterations run in models.kmeans.KMeansEngine — on GPUs one fused launch per iteration with
        # the regroup/update/allgather done in-kernel over NVLink, on CPUs sums + allreduce
        self.engine_mode = str(c.get("engine", "table"))
        self.timings: Dict[str, float] = {"compute": 0.0, "regroup": 0.0, "allgather": 0.0}

    def map_collective(self, reader: KeyValReader, context: Context) -> None:
        files: List[str] = []
        while reader.next_key_value():
            files.append(reader.get_current_value())
        self.run_kmeans(files, context)

    # ------------------------------------------------------------------------------------
    def _load_points(self, files: List[str]) -> torch.Tensor:
        d = self.vector_size
        mats = [F.read_kmeans_points_bin(f, d) for f in files]
        pts = np.concatenate(mats, axis=0) if mats else np.empty((0, d))
        return torch.from_numpy(pts).to(self.device, _work_dtype(self.device))

    def _create_cen_table(self, fill: Optional[np.ndarray]) -> Table:
        """``numMappers`` partitions; partition p holds centroids ``[p*cenParSize, ...)`` as rows
        ``[count, x...]`` (``KMeansCollectiveMapper.java:213-286``)."""
        N, K, v = self.get_num_workers(), self.num_centroids, self.cen_vec_size
        par = K // N
        rest = K % N
        dt = _work_dtype(self.device)
        table = Table(0, DoubleArrPlus())
        off = 0
        for p in range(N):
            rows = par + (1 if p < rest else 0)
            t = torch.zeros(rows * v, dtype=dt, device=self.device)
            if fill is not None:
                blk = torch.from_numpy(fill[off:off + rows]).to(self.device, dt)
                t.view(rows, v)[:, 1:] = blk
            table.add_partition(Partition(p, t))
            off += rows
        return table

    def _centroids_of(self, table: Table) -> torch.Tensor:
        v = self.cen_vec_size
        parts = [table.get_partition(p).get().view(-1, v)[:, 1:] for p in sorted(table.get_partition_ids())]
        return torch.cat(parts, dim=0).contiguous()

    def run_kmeans(self, files: List[str], context: Context) -> None:
        N, K, d, v = self.get_num_workers(), self.num_centroids, self.vector_size, self.cen_vec_size
        # master loads the initial centroids, everybody else receives them
        if self.is_master():
            init = F.read_centroids_text(self.cen_file, d)
            assert init.shape[0] == K, f"centroid file has {init.shape[0]} rows, expected {K}"
            cen_table = self._create_cen_table(init)
        else:
            cen_table = Table(0, DoubleArrPlus())
        if not self.broadcast("main", "broadcast-centroids", cen_table, self.get_master_id(), False):
            raise IOError("Fail to bcast")
        points = self._load_points(files)
        dt = _work_dtype(self.device)
        sizes = [cen_table.get_partition(p).get().numel() // v for p in sorted(cen_table.get_partition_ids())]
        self.objective_history: List[float] = []
        if self.engine_mode == "device":
            from ..models.kmeans import KMeansEngine
            t0 = time.perf_counter()
            # the application keeps fp32-faithful distances (split-precision tensor-core mode); HB_KMEANS_PRECISION overrides
            eng = KMeansEngine(points, K, comm="auto", symm=self.symm,
                               precision=os.environ.get("HB_KMEANS_PRECISION", "fp16x3") if self.device.type == "cuda" else None)
            eng.set_centroids(self._centroids_of(cen_table).to(points.dtype))
            for it in range(self.num_iterations):
                obj = eng.step()
                self.objective_history.append(float(obj.item()))       # global objective (sum over workers)
                context.progress()
            self.timings["compute"] += time.perf_counter() - t0
            self.centroids = eng.C.clone()
            LOG.info("[rank %d] device engine (%s): %.3f s", self.get_self_id(), eng.comm, self.timings["compute"])
            if self.is_master() and self.out_file:
                F.write_centroids_text(self.out_file, self.centroids.double().cpu().numpy())
            return
        for it in range(self.num_iterations):
            t0 = time.perf_counter()
            C = self._centroids_of(cen_table)
            res = kmeans_step(points, C, want_obj=True, faithful=True)
            # new table of local partial sums, same partitioning
            cen_table = Table(0, DoubleArrPlus())
            off = 0
            for p, rows in enumerate(sizes):
                t = torch.empty(rows, v, dtype=dt, device=self.device)
                t[:, 0] = res.counts[off:off + rows].to(dt)
                t[:, 1:] = res.sums[off:off + rows].to(dt)
                cen_table.add_partition(Partition(p, t.view(-1)))
                off += rows
            t1 = time.perf_counter()
            self.regroup("main", f"regroup-{it}", cen_table, Partitioner(N))
            t2 = time.perf_counter()
            for part in cen_table.get_partitions():          # owner divides (…Mapper.java:170-183)
                m = part.get().view(-1, v)
                cnt = m[:, 0:1]
                nz = cnt > 0
                m[:, 1:] = torch.where(nz, m[:, 1:] / torch.where(nz, cnt, torch.ones_like(cnt)), m[:, 1:])
            self.allgather("main", f"allgather-{it}", cen_table)
            t3 = time.perf_counter()
            self.timings["compute"] += t1 - t0
            self.timings["regroup"] += t2 - t1
            self.timings["allgather"] += t3 - t2
            self.objective_history.append(float(res.obj.item()))
            context.progress()
        self.centroids = self._centroids_of(cen_table)
        LOG.info("[rank %d] Compute: %.3f Regroup: %.3f Allgather: %.3f", self.get_self_id(),
                 self.timings["compute"], self.timings["regroup"], self.timings["allgather"])
        if self.is_master() and self.out_file:
            F.write_centroids_text(self.out_file, self.centroids.double().cpu().numpy())


class ContribKMeansMapper(CollectiveMapper):
    """conf keys (``KMeansConstants``): ``num_centroids``, ``vector_size``, ``num_iterations``,
    ``centroid_file``, ``work_dir``, ``comm`` (one of the four variants), ``tables`` (``dense`` on GPUs with a symmetric
    workspace: the centroid table is one slab and the collectives run on the NVLink kernels; ``sparse``: partition
    objects, generic path)."""

    VARIANTS = ("allreduce", "regroup-allgather", "broadcast-reduce", "push-pull")

    def setup(self, context: Context) -> None:
        c = context.get_configuration()
        self.num_centroids = int(c["num_centroids"])
        self.dimension = int(c["vector_size"])
        self.iteration = int(c["num_iterations"])
        self.cen_file = c["centroid_file"]
        self.work_dir = c.get("work_dir")
        self.variant = c.get("comm", "allreduce")
        self.tables = c.get("tables", "dense")         # "sparse": one partition per live centroid (generic object path)
        assert self.variant in self.VARIANTS, self.variant

    def map_collective(self, reader: KeyValReader, context: Context) -> None:
        files = []
        while reader.next_key_value():
            files.append(reader.get_current_value())
        self.run_kmeans(files, context)

    def _load_data(self, files: List[str]) -> torch.Tensor:
        mats = [F.read_points_text(f, self.dimension) for f in files]
        mats = [m for m in mats if m.size]
        pts = np.concatenate(mats, axis=0) if mats else np.empty((0, self.dimension))
        return torch.from_numpy(pts).to(self.device, _work_dtype(self.device))

    def _computation(self, cen_table: Table, prev: Table, points: torch.Tensor) -> float:
        """Assign to the nearest *existing* centroid; build the sparse partial table."""
        d = self.dimension
        ids = sorted(prev.get_partition_ids())
        if not ids or points.shape[0] == 0:
            return 0.0
        C = torch.stack([prev.get_partition(i).get()[:d] for i in ids]).contiguous()
        res = kmeans_step(points, C, want_assign=True, faithful=True)
        dt = _work_dtype(self.device)
        counts = res.counts
        # the tutorial's "calcEucDistSquare" returns the square ROOT (KmeansMapper.java:286-292), so its
        # "MSE" is the mean Euclidean distance; kept for parity with the 7.0 < MSE < 7.8 quality gate
        err = float((points - C[res.assign.long()]).pow(2).sum(1).sqrt().sum().item())
        live = torch.nonzero(counts > 0).flatten().tolist()
        for j in live:
            row = torch.empty(d + 1, dtype=dt, device=self.device)
            row[:d] = res.sums[j].to(dt)
            row[d] = counts[j].to(dt)
            cen_table.add_partition(Partition(ids[j], row))
        return err

    def _calculate_centroids(self, table: Table) -> None:
        d = self.dimension
        for p in table.get_partitions():
            row = p.get()
            if float(row[d]) > 0:
                row[:d] /= row[d]
                row[d] = 0

    def _run_dense(self, files: List[str], context: Context) -> None:
        """The same four variants with the centroid table as ONE dense slab in the symmetric workspace, so that every
        collective of the iteration runs on the NVLink kernels (allreduce; regroup rows + allgather rows; reduce +
        broadcast; push with ``red.add`` + pull with owner broadcast).  A centroid nobody was assigned to drops out of
        later iterations exactly as its partition does in the sparse form: the count column is kept (the reference
        zeroes it) and the rows with count 0 are masked in the next assignment pass."""
        N, d, K = self.get_num_workers(), self.dimension, self.num_centroids
        dt = _work_dtype(self.device)
        symm = self.comm.symm
        slabs = [symm.alloc(K * (d + 1), dt).view(K, d + 1) for _ in range(2)]
        owned = [p for p in range(K) if p % N == self.get_self_id()]
        gslab = symm.alloc(max(len(owned), 1) * (d + 1), dt).view(max(len(owned), 1), d + 1) if self.variant == "push-pull" else None
        cur = slabs[0]
        cur.zero_()
        if self.is_master():
            init = F.read_centroids_text(self.cen_file, d)
            cur[:, :d] = torch.from_numpy(np.asarray(init)[:K, :d]).to(self.device, dt)
            cur[:, d] = 1.0
        tables = [Table.dense(0, DoubleArrPlus(), s_) for s_ in slabs]        # layout check cached per table
        gtable = Table.dense(1, DoubleArrPlus(), gslab[:len(owned)], owned) if gslab is not None else None
        if not self.broadcast("main", "broadcast-centroids", tables[0], self.get_master_id(), False):
            raise IOError("Fail to bcast")
        points = self._load_data(files)
        self.num_points = points.shape[0]
        live = torch.ones(K, dtype=torch.bool, device=self.device)
        mse = 0.0
        self.collective_launches = 0

        def divide(rows: torch.Tensor) -> None:
            cnt = rows[:, d:d + 1]
            rows[:, :d] = torch.where(cnt > 0, rows[:, :d] / cnt.clamp(min=1), rows[:, :d])

        for it in range(self.iteration):
            prev, cur = slabs[it % 2], slabs[(it + 1) % 2]
            table = tables[(it + 1) % 2]
            if table.get_num_partitions() != K:                            # a regroup of an earlier iteration left only owned rows
                table = tables[(it + 1) % 2] = Table.dense(0, DoubleArrPlus(), cur)
            cur.zero_()
            idx = torch.nonzero(live).flatten()
            mse = 0.0
            if points.shape[0] and idx.numel():
                C = prev[idx, :d].contiguous()
                res = kmeans_step(points, C.float() if C.dtype != torch.float32 and self.device.type == "cuda" else C, want_assign=True, faithful=True)
                mse = float((points - C[res.assign.long()]).pow(2).sum(1).sqrt().sum().item())
                cur[idx, :d] = res.sums.to(dt)
                cur[idx, d] = res.counts.to(dt)
            l0 = _native.launches()
            if self.variant == "allreduce":
                self.allreduce("main", f"allreduce_{it}", table)
                divide(cur)
            elif self.variant == "regroup-allgather":
                self.regroup("main", f"regroup_{it}", table, Partitioner(N))
                if owned:
                    rows = cur[owned]
                    divide(rows)
                    cur[owned] = rows
                self.allgather("main", f"allgather_{it}", table)
            elif self.variant == "broadcast-reduce":
                self.reduce("main", f"reduce_{it}", table, self.get_master_id())
                if self.is_master():
                    divide(cur)
                self.broadcast("main", f"bcast_{it}", table, self.get_master_id(), False)
            else:  # push-pull
                gslab.zero_()
                self.push("main", f"push_{it}", table, gtable, Partitioner(N))
                if owned:
                    divide(gslab[:len(owned)])
                cur.zero_()
                self.pull("main", f"pull_{it}", table, gtable, True)
            self.collective_launches += _native.launches() - l0      # NVLink kernels of this iteration's collectives
            live = cur[:, d] > 0
            context.progress()
        mse_table = Table(0, DoubleArrPlus())
        mse_table.add_partition(Partition(0, torch.tensor([float(self.num_points), mse], dtype=torch.float64,
                                                          device=self.device)))
        self.allreduce("main", "allreduce-mse", mse_table)
        arr = mse_table.get_partition(0).get()
        self.final_mse = float(arr[1]) / max(float(arr[0]), 1.0)
        final = Table(0, DoubleArrPlus())
        for j in torch.nonzero(live).flatten().tolist():
            row = cur[j].clone(); row[d] = 0
            final.add_partition(Partition(j, row))
        self.cen_table = final
        if self.is_master():
            if self.work_dir:
                F.write_evaluation(os.path.join(self.work_dir, "evaluation"), "MSE", self.final_mse)
            for pid in sorted(final.get_partition_ids()):
                row = final.get_partition(pid).get()[:d].double().cpu().tolist()
                context.write(None, "\t".join(repr(x) for x in row) + "\t")
        # the slabs go back to the workspace's free list (same order on every rank): a long-lived mapper can run again
        torch.cuda.synchronize(self.device)
        if gslab is not None:
            symm.free(gslab)
        for s_ in reversed(slabs):
            symm.free(s_)

    def run_kmeans(self, files: List[str], context: Context) -> None:
        N, d = self.get_num_workers(), self.dimension
        dt = _work_dtype(self.device)
        if (self.device.type == "cuda" and N > 1 and getattr(self.comm, "symm", None) is not None
                and self.tables != "sparse"):
            return self._run_dense(files, context)
        cen_table = Table(0, DoubleArrPlus())
        if self.is_master():
            init = F.read_centroids_text(self.cen_file, d)
            for i, row in enumerate(init):
                t = torch.zeros(d + 1, dtype=dt, device=self.device)
                t[:d] = torch.from_numpy(row[:d]).to(dt)
                cen_table.add_partition(Partition(i, t))
        if not self.broadcast("main", "broadcast-centroids", cen_table, self.get_master_id(), False):
            raise IOError("Fail to bcast")
        points = self._load_data(files)
        self.num_points = points.shape[0]
        mse = 0.0
        for it in range(self.iteration):
            prev = cen_table
            cen_table = Table(0, DoubleArrPlus())
            mse = self._computation(cen_table, prev, points)
            if self.variant == "allreduce":
                self.allreduce("main", f"allreduce_{it}", cen_table)
                self._calculate_centroids(cen_table)
            elif self.variant == "regroup-allgather":
                self.regroup("main", f"regroup_{it}", cen_table, Partitioner(N))
                self._calculate_centroids(cen_table)
                self.allgather("main", f"allgather_{it}", cen_table)
            elif self.variant == "broadcast-reduce":
                self.reduce("main", f"reduce_{it}", cen_table, self.get_master_id())
                if self.is_master():
                    self._calculate_centroids(cen_table)
                self.broadcast("main", f"bcast_{it}", cen_table, self.get_master_id(), False)
            else:  # push-pull
                global_table = Table(1, DoubleArrPlus())
                self.push("main", f"push_{it}", cen_table, global_table, Partitioner(N))
                self._calculate_centroids(global_table)
                # pull combines into the local table: rebuild it zeroed with every known id
                # (contrib/.../pushpull/KmeansMapper.java:262-275)
                all_ids = sorted({i for ids in self.comm.all_gather_object(sorted(global_table.get_partition_ids()))
                                  for i in ids})
                cen_table = Table(0, DoubleArrPlus())
                for i in all_ids:
                    cen_table.add_partition(Partition(i, torch.zeros(d + 1, dtype=dt, device=self.device)))
                self.pull("main", f"pull_{it}", cen_table, global_table, True)
            context.progress()
        # MSE of the last assignment pass (KmeansMapper.java:222-249)
        mse_table = Table(0, DoubleArrPlus())
        mse_table.add_partition(Partition(0, torch.tensor([float(self.num_points), mse], dtype=torch.float64,
                                                          device=self.device)))
        self.allreduce("main", "allreduce-mse", mse_table)
        arr = mse_table.get_partition(0).get()
        self.final_mse = float(arr[1]) / max(float(arr[0]), 1.0)
        self.cen_table = cen_table
        if self.is_master():
            if self.work_dir:
                F.write_evaluation(os.path.join(self.work_dir, "evaluation"), "MSE", self.final_mse)
            for pid in sorted(cen_table.get_partition_ids()):
                row = cen_table.get_partition(pid).get()[:d].double().cpu().tolist()
                context.write(None, "\t".join(repr(x) for x in row) + "\t")


class KMeansRotationMapper(CollectiveMapper):
    """Centroids sharded over workers (partition id = shard id, rows ``[count, x...]``); the E-step
    rotates the shards N times keeping a per-point (bestDist, bestCentroid); the M-step rotates N more
    times accumulating each point into its visiting shard (``ExpTask.java:58-91``, ``MaxTask.java:50-72``)."""

    def setup(self, context: Context) -> None:
        c = context.get_configuration()
        self.num_centroids = int(c["num_centroids"])
        self.vector_size = int(c["vector_size"])
        self.num_iterations = int(c["num_iterations"])
        self.cen_file = c["centroid_file"]
        self.out_file = c.get("out_file")

    def map_collective(self, reader: KeyValReader, context: Context) -> None:
        files = []
        while reader.next_key_value():
            files.append(reader.get_current_value())
        N, K, d = self.get_num_workers(), self.num_centroids, self.vector_size
        dt = _work_dtype(self.device)
        me = self.get_self_id()
        bounds = [K * r // N for r in range(N + 1)]
        init = F.read_centroids_text(self.cen_file, d)           # every worker reads its own shard
        shard = torch.zeros(bounds[me + 1] - bounds[me], d + 1, dtype=dt, device=self.device)
        shard[:, 1:] = torch.from_numpy(init[bounds[me]:bounds[me + 1]]).to(dt)
        table = Table(0, DoubleArrPlus())
        table.add_partition(Partition(me, shard.view(-1)))
        mats = [F.read_kmeans_points_bin(f, d) for f in files]
        pts = torch.from_numpy(np.concatenate(mats, 0) if mats else np.empty((0, d))).to(self.device, dt)
        n = pts.shape[0]
        self.objective_history = []
        for it in range(self.num_iterations):
            best_d = torch.full((n,), float("inf"), dtype=dt, device=self.device)
            best_c = torch.zeros(n, dtype=torch.long, device=self.device)
            for j in range(N):                                       # E-step
                (part,) = table.get_partitions()
                sid = part.id()
                C = part.get().view(-1, d + 1)[:, 1:].contiguous()
                if n and C.shape[0]:
                    res = kmeans_step(pts, C, want_assign=True, faithful=True)
                    dist = (pts - C[res.assign.long()]).pow(2).sum(1)
                    better = dist < best_d
                    best_d = torch.where(better, dist, best_d)
                    best_c = torch.where(better, res.assign.long() + bounds[sid], best_c)
                self.rotate("main", f"rotate-e-{it}-{j}", table, None)
            (part,) = table.get_partitions()                        # every shard starts the M-step empty
            part.get().zero_()
            for j in range(N):                                       # M-step
                (part,) = table.get_partitions()
                sid = part.id()
                m = part.get().view(-1, d + 1)
                sel = (best_c >= bounds[sid]) & (best_c < bounds[sid + 1])
                if n and bool(sel.any()):
                    idx = best_c[sel] - bounds[sid]
                    m[:, 1:].index_add_(0, idx, pts[sel])
                    m[:, 0].index_add_(0, idx, torch.ones(idx.numel(), dtype=dt, device=self.device))
                self.rotate("main", f"rotate-m-{it}-{j}", table, None)
            (part,) = table.get_partitions()                        # shard is home again: normalise
            m = part.get().view(-1, d + 1)
            nz = m[:, 0:1] > 0
            m[:, 1:] = torch.where(nz, m[:, 1:] / torch.where(nz, m[:, 0:1], torch.ones_like(m[:, 0:1])), m[:, 1:])
            obj = Table(1, DoubleArrPlus())
            obj.add_partition(Partition(0, best_d.sum().double().view(1)))
            self.allreduce("main", f"obj-{it}", obj)
            self.objective_history.append(float(obj.get_partition(0).get().item()))
        # gather the shards for output
        self.allgather("main", "allgather-final", table)
        self.centroids = torch.cat([table.get_partition(p).get().view(-1, d + 1)[:, 1:] for p in
                                    sorted(table.get_partition_ids())], 0)
        if self.is_master() and self.out_file:
            F.write_centroids_text(self.out_file, self.centroids.double().cpu().numpy())
