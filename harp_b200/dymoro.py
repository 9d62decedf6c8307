"""dymoro — dynamic model rotation (``ml/java/src/main/java/edu/iu/dymoro``).

* ``RotationUtil.create_rotation_order`` — randomised ring orders, layout
  ``[P_0..P_{N-1} | S_1..S_{N-1}]`` per row (``RotationUtil.java:58-90``);
  ``get_rotation_sequences`` = master creates, broadcast to everyone (``:34-55``).
* ``RotateTask`` — owns one model slice (a ``Table``), turns the order array into the per-call
  ``rotationMap`` (``RotateTask.java:158-210``), rotates, then re-splits the received partitions into
  ``numColSplits`` lists (round-robin over sorted ids, or random) (``:107-150``).
* ``Rotator`` — one RotateTask per model slice; ``rotate(k)`` is asynchronous, ``get_split_map(k)``
  blocks until slice k's outstanding rotation has finished (``Rotator.java:38-85``): slice k travels
  while slice k±1 is being computed on (2-slice double buffering).
* ``Scheduler`` — timer-bounded dynamic block scheduler over a ``numRowSplits x numColSplits`` grid
  with row/column exclusivity and reservoir-sampled refills (``Scheduler.java:54-237``).
* ``MPTask`` — base of the compute tasks with throughput recording (``MPTask.java:41-79``).

B200 mapping: on the device the same ring is a peer *pull* over NVLink on a side stream
(``SymmetricWorkspace.rotate_``) and block exclusivity is a Latin-square schedule inside one kernel
(``ops/mf.py``); this module keeps the host-level API and semantics for table-based apps.
All slices share ONE communication thread so that every rank issues its collectives in the same
order (the reference relies on unique operation names for that).
"""
from __future__ import annotations

import random as _random
import threading
import time
from typing import Any, Dict, List, Optional, Sequence

import torch

from .partition import IntArrPlus, Partition, Table
from .sched import DynamicScheduler, StaticScheduler, Task


# ---------------------------------------------------------------------------------------
# rotation orders
# ---------------------------------------------------------------------------------------
def create_rotation_order(rng: _random.Random, num_iterations: int, num_workers: int) -> List[int]:
    orders: List[int] = []
    for i in range(num_iterations):
        if i == 0:
            orders.extend(range(num_workers))
        else:
            ids = list(range(num_workers))
            while ids:
                orders.append(ids.pop(rng.randrange(len(ids))))
        shifts = list(range(1, num_workers))
        while shifts:
            orders.append(shifts.pop(rng.randrange(len(shifts))))
    return orders


def get_rotation_sequences(rng: _random.Random, num_workers: int, num_iterations: int, mapper,
                           context_name: str = "sgd") -> List[int]:
    table: Table = Table(0, IntArrPlus())
    if mapper.is_master():
        orders = create_rotation_order(rng, num_iterations, num_workers)
        table.add_partition(Partition(0, torch.tensor(orders, dtype=torch.int32)))
    mapper.broadcast(context_name, "bcast-rotate-orders", table, mapper.get_master_id(), False)
    return [int(x) for x in table.get_partition(0).get().tolist()]


class RotationMapper:
    """The order-array walk of ``RotateTask.updateRotationMap`` as a standalone object, so the
    device-side rings can derive the same per-step ``dest`` / ``source`` tables."""

    def __init__(self, orders: Optional[Sequence[int]], num_workers: int):
        self.orders = list(orders) if orders is not None else None
        self.n = num_workers
        self.row_len = 2 * num_workers - 1
        self.cur = 0
        self.data_worker: Dict[int, int] = {}
        if self.orders is not None:
            for i in range(num_workers):
                self.data_worker[i] = self.orders[self.cur]
                self.cur += 1

    def next_map(self) -> Optional[Dict[int, int]]:
        """rotationMap for the next rotate call (None = plain ring ``next = (self+1) % N``)."""
        if self.orders is None or self.cur >= len(self.orders):
            return None
        o, n = self.orders, self.n
        rmap: Dict[int, int] = {}
        if self.cur % self.row_len == 0:
            last_shift = o[self.cur - 1]
            for i in range(n):
                origin = self.data_worker[i]
                cur_w = (origin + last_shift) % n
                new_w = o[self.cur]
                self.data_worker[i] = new_w
                rmap[cur_w] = new_w
                self.cur += 1
        elif self.cur % self.row_len == n:
            for i in range(n):
                cur_w = self.data_worker[i]
                rmap[cur_w] = (cur_w + o[self.cur]) % n
            self.cur += 1
        else:
            for i in range(n):
                origin = self.data_worker[i]
                rmap[(origin + o[self.cur - 1]) % n] = (origin + o[self.cur]) % n
            self.cur += 1
        return rmap


# ---------------------------------------------------------------------------------------
# RotateTask / Rotator
# ---------------------------------------------------------------------------------------
class RotateTask:
    def __init__(self, table: Table, num_col_splits: int, random_split: bool, mapper,
                 orders: Optional[Sequence[int]], context_name: str, seed: Optional[int] = None):
        self.table = table
        self.num_col_splits = num_col_splits
        self.random_split = random_split
        self.mapper = mapper
        self.context_name = context_name
        self.rng = _random.Random(seed if seed is not None else time.time_ns())
        self.split_map: List[List[Partition]] = [[] for _ in range(num_col_splits)]
        self.operation_id = 0
        self.rot = RotationMapper(orders, mapper.get_num_workers())
        self.comm_time = 0.0
        self._split()

    def _split(self) -> None:
        for lst in self.split_map:
            lst.clear()
        if self.random_split:
            for p in self.table.get_partitions():
                self.split_map[self.rng.randrange(self.num_col_splits)].append(p)
        else:
            for i, pid in enumerate(sorted(self.table.get_partition_ids())):
                self.split_map[i % self.num_col_splits].append(self.table.get_partition(pid))

    def get_split_map(self) -> List[List[Partition]]:
        return self.split_map

    def set_random_split(self, flag: bool) -> None:
        self.random_split = flag

    def run(self, cmd: int = 0) -> List[List[Partition]]:
        t1 = time.perf_counter()
        rmap = self.rot.next_map()
        self.mapper.rotate(self.context_name, f"rotate-{self.table.get_table_id()}-{self.operation_id}",
                           self.table, rmap)
        self._split()
        self.operation_id += 1
        self.comm_time += time.perf_counter() - t1
        return self.split_map

    def reset_comm_time(self) -> float:
        t, self.comm_time = self.comm_time, 0.0
        return t


class _CommWorker(Task):
    def __init__(self, tasks: List[RotateTask]):
        self.tasks = tasks

    def run(self, k: int):
        self.tasks[k].run(0)
        return ("done", k)


class Rotator:
    """``Rotator(tableMap[], numSplits, randomSplit, mapper, orders|null, ctxName)``."""

    def __init__(self, table_map: Sequence[Table], num_splits: int, random_split: bool, mapper,
                 orders: Optional[Sequence[int]], context_name: str, seed: Optional[int] = None):
        self.tasks = [RotateTask(t, num_splits, random_split, mapper, orders, context_name,
                                 None if seed is None else seed + i) for i, t in enumerate(table_map)]
        self.num_slices = len(self.tasks)
        self._sched: StaticScheduler = StaticScheduler([_CommWorker(self.tasks)])
        self._pending = [0] * self.num_slices
        self._lock = threading.Lock()
        self._done_cv = threading.Condition()
        self._collector: Optional[threading.Thread] = None

    def start(self) -> None:
        self._sched.start()

    def get_split_map(self, k: int) -> List[List[Partition]]:
        """Blocks until every outstanding rotation of slice ``k`` has completed."""
        while True:
            with self._lock:
                if self._pending[k] == 0:
                    break
            out = self._sched.wait_for_output(0)
            if out is None:
                time.sleep(0.0005)
                continue
            with self._lock:
                self._pending[out[1]] -= 1
        return self.tasks[k].get_split_map()

    def rotate(self, k: int) -> None:
        with self._lock:
            self._pending[k] += 1
        self._sched.submit(0, k)

    def pause(self) -> None:
        for k in range(self.num_slices):
            self.get_split_map(k)
        self._sched.pause()

    def stop(self) -> None:
        for k in range(self.num_slices):
            self.get_split_map(k)
        self._sched.stop()

    def set_random_split(self, flag: bool) -> None:
        for t in self.tasks:
            t.set_random_split(flag)

    def get_comm_time(self) -> float:
        return sum(t.reset_comm_time() for t in self.tasks)

    getSplitMap, setRandomSplit = get_split_map, set_random_split


# ---------------------------------------------------------------------------------------
# MPTask / Scheduler
# ---------------------------------------------------------------------------------------
class RowColSplit:
    __slots__ = ("row", "col", "r_data", "c_data", "num_items")

    def __init__(self, row: int, col: int, r_data: Any):
        self.row, self.col, self.r_data, self.c_data, self.num_items = row, col, r_data, None, 0


class MPTask(Task):
    """``doRun(cData, rData) -> items processed``; records duration / items when asked
    (``MPTask.java:41-79``)."""

    def __init__(self):
        self._record = False
        self._duration = 0.0
        self._items = 0

    def do_run(self, c_data: List[Partition], r_data: Any) -> int:  # pragma: no cover - interface
        raise NotImplementedError

    def run(self, split: RowColSplit) -> RowColSplit:
        t0 = time.perf_counter()
        n = self.do_run(split.c_data, split.r_data)
        split.num_items = n
        if self._record:
            self._duration += time.perf_counter() - t0
            self._items += n
        return split

    def start_record(self, record: bool) -> None:
        self._record = record
        self._duration, self._items = 0.0, 0

    def get_record_duration(self) -> float:
        return self._duration

    def get_items_recorded(self) -> int:
        return self._items

    doRun, startRecord, getRecordDuration, getItemsRecorded = (
        do_run, start_record, get_record_duration, get_items_recorded)


class Scheduler:
    """Timer-bounded block scheduler: at any time no two running blocks share a row split or a
    column split; freed rows/columns are re-paired by reservoir sampling among not-yet-trained blocks;
    stops submitting when the timer fires or the grid is exhausted (``Scheduler.java:95-237``)."""

    def __init__(self, num_row_splits: int, num_col_splits: int, vwh_map: Sequence[Any], time_ms: float,
                 tasks: Sequence[MPTask], seed: Optional[int] = None):
        self.num_row_splits, self.num_col_splits = num_row_splits, num_col_splits
        self.num_row_limit, self.num_col_limit = num_col_splits, num_row_splits
        self.split_map = [[RowColSplit(i, j, vwh_map[i]) for j in range(num_col_splits)]
                          for i in range(num_row_splits)]
        self.time_ms = time_ms
        self.rng = _random.Random(seed if seed is not None else time.time_ns())
        self.compute: DynamicScheduler = DynamicScheduler(list(tasks))
        self.compute.start()
        self.compute.pause_now()
        self.num_items_trained = 0
        self._running = threading.Event()

    def set_timer(self, time_ms: float) -> None:
        self.time_ms = time_ms

    def get_num_v_items_trained(self) -> int:
        return self.num_items_trained

    def schedule(self, h_map: Sequence[List[Partition]], record: bool = False) -> None:
        R, Cn = self.num_row_splits, self.num_col_splits
        submitted = [[False] * Cn for _ in range(R)]
        row_count, col_count = [0] * R, [0] * Cn
        for i in range(R):
            for j in range(Cn):
                self.split_map[i][j].c_data = h_map[j]
        self.num_items_trained = 0
        free_row, free_col = list(range(R)), list(range(Cn))

        def submit(i: int, j: int) -> None:
            submitted[i][j] = True
            row_count[i] += 1
            col_count[j] += 1
            self.compute.submit(self.split_map[i][j])

        while free_row and free_col:
            i = free_row.pop(self.rng.randrange(len(free_row)))
            j = free_col.pop(self.rng.randrange(len(free_col)))
            submit(i, j)
        self._running.set()
        timer = threading.Timer(self.time_ms / 1000.0, self._running.clear)
        for t in self.compute.get_tasks():
            t.start_record(record)
        timer.start()
        self.compute.start()
        while self.compute.has_output():
            split = self.compute.wait_for_output()
            if split is None:
                break
            self.num_items_trained += split.num_items
            fr = split.row if row_count[split.row] < self.num_row_limit else -1
            fc = split.col if col_count[split.col] < self.num_col_limit else -1
            if not self._running.is_set():
                break
            if fr != -1 and free_col:
                cands = [idx for idx, c in enumerate(free_col) if not submitted[fr][c]]
                if cands:
                    idx = cands[self.rng.randrange(len(cands))]   # uniform choice == reservoir sampling
                    c = free_col.pop(idx)
                    submit(fr, c)
                    fr = -1
            if fc != -1 and free_row:
                cands = [idx for idx, r in enumerate(free_row) if not submitted[r][fc]]
                if cands:
                    idx = cands[self.rng.randrange(len(cands))]
                    r = free_row.pop(idx)
                    submit(r, fc)
                    fc = -1
            if fr != -1:
                free_row.append(fr)
            if fc != -1:
                free_col.append(fc)
        self.compute.pause_now()
        timer.cancel()
        while self.compute.has_output():
            s = self.compute.wait_for_output()
            if s is not None:
                self.num_items_trained += s.num_items
        self.compute.clean_input_queue()

    def stop(self) -> None:
        self.compute.stop()

    setTimer, getNumVItemsTrained = set_timer, get_num_v_items_trained
