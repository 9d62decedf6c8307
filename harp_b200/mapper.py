"""CollectiveMapper — the user-facing programming model.

Reference: ``core/harp-hadoop/src/main/java/org/apache/hadoop/mapred/CollectiveMapper.java``
(``run()`` = init comm → ``setup`` → ``mapCollective(reader, context)`` → ``cleanup`` →
stop client/server, ``:751-790``; the ten collectives + events + identity helpers,
``:317-714``).  Method names and argument order are kept (camelCase aliases included);
``contextName`` / ``operationName`` no longer address a mailbox — ordering on the
communicator replaces it — but they are kept for logging and become NVTX range names.

One process per GPU replaces a Hadoop map task: rank / world size come from ``torchrun``
(``RANK``, ``WORLD_SIZE``, ``LOCAL_RANK``), the rendezvous files ``<jobID>/{nodes,tasks,lock}``
are replaced by ``init_process_group`` and one symmetric-memory handle exchange.
"""
from __future__ import annotations

import logging
import os
import pickle
import time
from contextlib import contextmanager
from typing import Any, Dict, Iterable, Iterator, List, Optional, Tuple

import torch
import torch.distributed as dist

from . import collective as C
from .partition import Partitioner, Table
from .resource import ResourcePool

LOG = logging.getLogger("harp_b200")


# ---------------------------------------------------------------------------------------
# Events (client/Event.java:37-98, EventType.java:25-27, io/EventQueue.java:28-74)
# ---------------------------------------------------------------------------------------
class EventType:
    MESSAGE_EVENT = "MESSAGE_EVENT"        # to one worker
    COLLECTIVE_EVENT = "COLLECTIVE_EVENT"  # to every other worker
    LOCAL_EVENT = "LOCAL_EVENT"            # loops back to the sender


class Event:
    def __init__(self, event_type: str, context_name: str, source_id: int, target_id: int, body: Any):
        self.event_type = event_type
        self.context_name = context_name
        self.source_id = source_id
        self.target_id = target_id
        self.body = body

    def get_event_type(self):
        return self.event_type

    def get_context_name(self):
        return self.context_name

    def get_source_id(self):
        return self.source_id

    def get_target_id(self):
        return self.target_id

    def get_body(self):
        return self.body

    getEventType, getContextName, getSourceID, getTargetID, getBody = (
        get_event_type, get_context_name, get_source_id, get_target_id, get_body)


class _Mailbox:
    """Asynchronous per-worker mailbox on the process group's key-value store: no receiver thread,
    no sockets of our own (replaces SyncClient/SyncQueue/Server/Acceptor/DataReceiver)."""

    def __init__(self, rank: int, world: int):
        self.rank, self.world = rank, world
        self.local: List[Event] = []
        self.consumed = 0
        self.store = None
        if dist.is_available() and dist.is_initialized() and world > 1:
            try:
                self.store = dist.distributed_c10d._get_default_store()
            except Exception:  # noqa: BLE001
                self.store = None

    def _post(self, dest: int, ev: Event) -> None:
        if dest == self.rank or self.store is None:
            self.local.append(ev)
            return
        seq = self.store.add(f"harp/evt/{dest}/n", 1)
        self.store.set(f"harp/evt/{dest}/{seq}", pickle.dumps(ev))

    def send(self, ev: Event) -> bool:
        if ev.event_type == EventType.LOCAL_EVENT:
            self.local.append(ev)
        elif ev.event_type == EventType.MESSAGE_EVENT:
            self._post(ev.target_id, ev)
        elif ev.event_type == EventType.COLLECTIVE_EVENT:
            for d in range(self.world):
                if d != self.rank:
                    self._post(d, ev)
        else:
            return False
        return True

    def poll(self) -> Optional[Event]:
        if self.local:
            return self.local.pop(0)
        if self.store is None:
            return None
        n = self.store.add(f"harp/evt/{self.rank}/n", 0)
        if self.consumed < n:
            self.consumed += 1
            key = f"harp/evt/{self.rank}/{self.consumed}"
            data = self.store.get(key)     # blocks until the sender's set() lands
            try:
                self.store.delete_key(key)
            except Exception:  # noqa: BLE001
                pass
            return pickle.loads(data)
        return None

    def wait(self, timeout_s: float = 1800.0) -> Optional[Event]:
        t0 = time.time()
        while True:
            ev = self.poll()
            if ev is not None:
                return ev
            if time.time() - t0 > timeout_s:   # Constant.DATA_MAX_WAIT_TIME = 1800 s
                return None
            time.sleep(0.001)


# ---------------------------------------------------------------------------------------
# Context / reader
# ---------------------------------------------------------------------------------------
class KeyValReader:
    """Iterates the (key, value) records dealt to this worker (``CollectiveMapper.java:88-130``);
    with ``MultiFileInputFormat`` both key and value are a file path."""

    def __init__(self, records: Iterable[Tuple[Any, Any]]):
        self._it: Iterator[Tuple[Any, Any]] = iter(list(records))
        self._cur: Optional[Tuple[Any, Any]] = None

    def next_key_value(self) -> bool:
        try:
            self._cur = next(self._it)
            return True
        except StopIteration:
            self._cur = None
            return False

    def get_current_key(self):
        return self._cur[0]

    def get_current_value(self):
        return self._cur[1]

    nextKeyValue, getCurrentKey, getCurrentValue = next_key_value, get_current_key, get_current_value


class Context:
    """Stands in for Hadoop's ``Mapper.Context``: configuration + output collector + progress."""

    def __init__(self, conf: Optional[Dict[str, Any]] = None, output_dir: Optional[str] = None, rank: int = 0):
        self.conf: Dict[str, Any] = dict(conf or {})
        self.output: List[Tuple[Any, Any]] = []
        self.output_dir = output_dir
        self.rank = rank
        self._progress = 0

    def get_configuration(self) -> Dict[str, Any]:
        return self.conf

    def write(self, key, value) -> None:
        self.output.append((key, value))

    def progress(self) -> None:
        self._progress += 1

    def flush_output(self) -> Optional[str]:
        """Writes collected (key, value) pairs tab-separated as ``part-m-<rank>`` (Hadoop text output)."""
        if self.output_dir is None or not self.output:
            return None
        os.makedirs(self.output_dir, exist_ok=True)
        path = os.path.join(self.output_dir, f"part-m-{self.rank:05d}")
        with open(path, "w") as f:
            for k, v in self.output:
                f.write(f"{k}\t{v}\n" if k is not None else f"{v}\n")
        return path

    getConfiguration = get_configuration


# ---------------------------------------------------------------------------------------
# CollectiveMapper
# ---------------------------------------------------------------------------------------
class CollectiveMapper:
    """Subclass, override :meth:`setup` / :meth:`map_collective` / :meth:`cleanup`, then ``run(...)``."""

    def __init__(self, device: Optional[torch.device] = None, use_symm: Optional[bool] = None,
                 symm_bytes: int = 256 << 20):
        self._comm: Optional[C.Comm] = None
        self._mailbox: Optional[_Mailbox] = None
        self._device = device
        self._use_symm = use_symm
        self._symm_bytes = symm_bytes
        self._op_times: Dict[str, float] = {}
        self._t_start = time.time()

    # ---- lifecycle (override) -------------------------------------------------------------
    def setup(self, context: Context) -> None:
        pass

    def map_collective(self, reader: KeyValReader, context: Context) -> None:
        raise NotImplementedError

    def cleanup(self, context: Context) -> None:
        pass

    # reference spelling
    def mapCollective(self, reader: KeyValReader, context: Context) -> None:  # noqa: N802
        return self.map_collective(reader, context)

    # ---- bootstrap --------------------------------------------------------------------------
    def init_coll_comm_components(self) -> bool:
        """``initCollCommComponents`` (``CollectiveMapper.java:253-310``): communicator, symmetric
        workspace (GPU), mailbox, then the start-up handshake barrier."""
        symm = None
        if self._device is None:
            if dist.is_initialized() and "nccl" in dist.get_backend():
                self._device = torch.device("cuda", torch.cuda.current_device())
            elif torch.cuda.is_available() and not dist.is_initialized():
                self._device = torch.device("cuda", torch.cuda.current_device())
            else:
                self._device = torch.device("cpu")
        want_symm = self._use_symm
        if want_symm is None:
            want_symm = (self._device.type == "cuda" and dist.is_initialized() and dist.get_world_size() > 1
                         and os.environ.get("HARP_DISABLE_SYMM", "0") != "1")
        if want_symm:
            try:
                from .parallel.symm import SymmetricWorkspace
                symm = SymmetricWorkspace(self._device, self._symm_bytes)
            except Exception as e:  # noqa: BLE001
                LOG.warning("symmetric workspace unavailable (%r); collectives use NCCL", e)
                symm = None
        self._comm = C.Comm(device=self._device, symm=symm)
        self._mailbox = _Mailbox(self._comm.rank, self._comm.world)
        return self.barrier("start-worker", "handshake")

    def run(self, records: Iterable[Tuple[Any, Any]] = (), conf: Optional[Dict[str, Any]] = None,
            output_dir: Optional[str] = None) -> Context:
        """``CollectiveMapper.run`` (``:751-790``)."""
        ok = self.init_coll_comm_components()
        if not ok:
            raise IOError("Fail to do master barrier.")
        ctx = Context(conf, output_dir, self.get_self_id())
        reader = KeyValReader(records)
        try:
            self.setup(ctx)
            self.map_collective(reader, ctx)
            self.cleanup(ctx)
        finally:
            ctx.flush_output()
            self.log_resource_pools()
        return ctx

    # ---- identity (CollectiveMapper.java:317-365) -------------------------------------------
    @property
    def comm(self) -> C.Comm:
        assert self._comm is not None, "call run() or init_coll_comm_components() first"
        return self._comm

    @property
    def device(self) -> torch.device:
        return self.comm.device

    @property
    def symm(self):
        return self.comm.symm

    def get_self_id(self) -> int:
        return self.comm.rank

    def get_master_id(self) -> int:
        return 0

    def is_master(self) -> bool:
        return self.comm.rank == 0

    def get_num_workers(self) -> int:
        return self.comm.world

    def get_min_id(self) -> int:
        return 0

    def get_max_id(self) -> int:
        return self.comm.world - 1

    # ---- tracing ---------------------------------------------------------------------------
    @contextmanager
    def _range(self, context_name: str, operation_name: str):
        name = f"{context_name}/{operation_name}"
        pushed = False
        if self.device.type == "cuda":
            try:
                torch.cuda.nvtx.range_push(name)
                pushed = True
            except Exception:  # noqa: BLE001
                pushed = False
        t0 = time.perf_counter()
        try:
            yield
        finally:
            self._op_times[name] = self._op_times.get(name, 0.0) + (time.perf_counter() - t0)
            if pushed:
                torch.cuda.nvtx.range_pop()

    # ---- collectives (CollectiveMapper.java:374-663) ------------------------------------------
    def barrier(self, context_name: str, operation_name: str) -> bool:
        with self._range(context_name, operation_name):
            return C.barrier(self.comm, context_name, operation_name)

    def broadcast(self, context_name: str, operation_name: str, table: Table, bcast_worker_id: int,
                  use_mst_bcast: bool = False) -> bool:
        with self._range(context_name, operation_name):
            return C.broadcast(self.comm, table, bcast_worker_id, use_mst_bcast)

    def reduce(self, context_name: str, operation_name: str, table: Table, reduce_worker_id: int) -> bool:
        with self._range(context_name, operation_name):
            return C.reduce(self.comm, table, reduce_worker_id)

    def allgather(self, context_name: str, operation_name: str, table: Table) -> bool:
        with self._range(context_name, operation_name):
            return C.allgather(self.comm, table)

    def allreduce(self, context_name: str, operation_name: str, table: Table) -> bool:
        with self._range(context_name, operation_name):
            return C.allreduce(self.comm, table)

    def regroup(self, context_name: str, operation_name: str, table: Table,
                partitioner: Optional[Partitioner] = None) -> bool:
        with self._range(context_name, operation_name):
            return C.regroup(self.comm, table, partitioner)

    def aggregate(self, context_name: str, operation_name: str, table: Table,
                  partitioner: Optional[Partitioner], fn) -> bool:
        with self._range(context_name, operation_name):
            return C.aggregate(self.comm, table, partitioner, fn)

    def pull(self, context_name: str, operation_name: str, local_table: Table, global_table: Table,
             use_bcast: bool = False) -> bool:
        with self._range(context_name, operation_name):
            return C.pull(self.comm, local_table, global_table, use_bcast)

    def push(self, context_name: str, operation_name: str, local_table: Table, global_table: Table,
             partitioner: Optional[Partitioner] = None) -> bool:
        with self._range(context_name, operation_name):
            return C.push(self.comm, local_table, global_table, partitioner)

    def rotate(self, context_name: str, operation_name: str, global_table: Table,
               rotate_map: Optional[Dict[int, int]] = None) -> bool:
        with self._range(context_name, operation_name):
            return C.rotate(self.comm, global_table, rotate_map)

    def join(self, context_name: str, operation_name: str, dynamic_table: Table, static_table: Table) -> bool:
        with self._range(context_name, operation_name):
            return C.join(self.comm, dynamic_table, static_table)

    def group_by_key(self, context_name: str, operation_name: str, table: Table,
                     partitioner: Optional[Partitioner] = None, fn=None) -> bool:
        with self._range(context_name, operation_name):
            return C.group_by_key(self.comm, table, partitioner, fn)

    # ---- events (CollectiveMapper.java:623-663) ---------------------------------------------
    def get_event(self) -> Optional[Event]:
        return self._mailbox.poll()

    def wait_event(self, timeout_s: float = 1800.0) -> Optional[Event]:
        return self._mailbox.wait(timeout_s)

    def send_event(self, event: Event) -> bool:
        if event is None:
            return False
        return self._mailbox.send(event)

    # ---- housekeeping (CollectiveMapper.java:670-714) ---------------------------------------
    def free_memory(self) -> None:
        ResourcePool.get().clean()
        if self.device.type == "cuda":
            torch.cuda.empty_cache()

    def free_conn(self) -> None:
        """No pooled sockets to drop here (``ConnPool.clean`` in the reference)."""

    def log_mem_usage(self) -> Dict[str, float]:
        out: Dict[str, float] = {}
        if self.device.type == "cuda":
            st = torch.cuda.memory_stats(self.device)
            out["allocated_mb"] = st.get("allocated_bytes.all.current", 0) / 2 ** 20
            out["reserved_mb"] = st.get("reserved_bytes.all.current", 0) / 2 ** 20
            out["peak_allocated_mb"] = st.get("allocated_bytes.all.peak", 0) / 2 ** 20
        try:
            import psutil
            out["host_rss_mb"] = psutil.Process().memory_info().rss / 2 ** 20
        except Exception:  # noqa: BLE001
            pass
        LOG.info("[rank %d] memory: %s", self.get_self_id(), out)
        return out

    def log_gc_time(self) -> Dict[str, float]:
        """The JVM GC report has no analogue; per-op wall time of the collectives is reported instead."""
        LOG.info("[rank %d] collective time (s): %s", self.get_self_id(), self._op_times)
        return dict(self._op_times)

    def log_resource_pools(self) -> Dict[str, Dict[str, int]]:
        stats = ResourcePool.get().log()
        LOG.debug("[rank %d] pools: %s", self.get_self_id() if self._comm else -1, stats)
        return stats

    # camelCase aliases --------------------------------------------------------------------
    getSelfID, getMasterID, isMaster, getNumWorkers, getMinID, getMaxID = (
        get_self_id, get_master_id, is_master, get_num_workers, get_min_id, get_max_id)
    getEvent, waitEvent, sendEvent = get_event, wait_event, send_event
    freeMemory, freeConn, logMemUsage, logGCTime = free_memory, free_conn, log_mem_usage, log_gc_time


# ---------------------------------------------------------------------------------------
# process-group bootstrap helper used by the launchers
# ---------------------------------------------------------------------------------------
def init_distributed(backend: Optional[str] = None) -> Tuple[int, int, torch.device]:
    """Initialises torch.distributed from the torchrun environment (no-op for a single process).

    Returns ``(rank, world, device)``.  ``backend`` defaults to NCCL when a GPU is visible, gloo otherwise.
    """
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    use_cuda = torch.cuda.is_available() and os.environ.get("HARP_FORCE_CPU", "0") != "1"
    device = torch.device("cuda", local_rank % max(torch.cuda.device_count(), 1)) if use_cuda else torch.device("cpu")
    if use_cuda:
        torch.cuda.set_device(device)
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        be = backend or ("nccl" if use_cuda else "gloo")
        if be == "nccl":
            dist.init_process_group(be, device_id=device)
        else:
            dist.init_process_group(be)
    return rank, world, device
