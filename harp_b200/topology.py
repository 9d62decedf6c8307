"""Worker topology: ``Nodes`` / ``Workers`` / ``WorkerInfo``.

Reference: ``core/harp-collective/src/main/java/edu/iu/harp/worker/{Nodes,Workers,WorkerInfo}.java``.  Under
``torchrun`` the ranks already are the worker ids, so nothing in the collectives needs this; it is kept because
launch scripts and user code of the reference read a *nodes file* (``#rack`` lines followed by one host per
worker) and ask the resulting object for master / min / max / middle / next ids and rack membership
(``Workers.java:118-155``).  ``Workers.from_distributed()`` builds the same object from the process group.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, Iterator, List, Optional

DEFAULT_WORKER_PORT_BASE = 12800          # HC/io/Constant.java (DEFAULT_WORKER_POART_BASE)
UNKNOWN_WORKER_ID = -1


@dataclass(frozen=True)
class WorkerInfo:
    """``WorkerInfo.java``: id, host, port (= port base + id), rack."""
    id: int
    node: str
    port: int
    rack: int


class Nodes:
    """Rack id → hosts, in file order (``Nodes.java:89-118``)."""

    def __init__(self, racks: Dict[int, List[str]], rack_order: Optional[List[int]] = None):
        self._racks = {int(k): list(v) for k, v in racks.items()}
        self._order = list(rack_order) if rack_order is not None else list(self._racks.keys())

    @classmethod
    def from_file(cls, path: str) -> "Nodes":
        with open(path) as f:
            return cls.from_lines(f)

    @classmethod
    def from_lines(cls, lines) -> "Nodes":
        racks: Dict[int, List[str]] = {}
        order: List[int] = []
        cur: Optional[int] = None
        for line in lines:
            line = line.strip()
            if not line:
                continue
            if line.startswith("#"):
                cur = int(line[1:].strip() or 0)
                if cur not in racks:
                    racks[cur] = []
                    order.append(cur)
            else:
                if cur is None:                    # hosts before any rack line: rack 0
                    cur = 0
                    racks[cur] = []
                    order.append(cur)
                racks[cur].append(line)
        return cls(racks, order)

    def get_num_physical_nodes(self) -> int:
        return sum(len(v) for v in self._racks.values())

    def get_nodes(self) -> Dict[int, List[str]]:
        return self._racks

    def get_rack_list(self) -> List[int]:
        return self._order

    def get_node_list(self) -> List[str]:
        return [h for r in self._order for h in self._racks[r]]

    def sort_racks(self) -> None:
        """Racks with more nodes first (``Nodes.sortRacks``), stable for ties."""
        self._order.sort(key=lambda r: -len(self._racks[r]))

    getNumPhysicalNodes, getNodes, getRackList, getNodeList, sortRacks = (
        get_num_physical_nodes, get_nodes, get_rack_list, get_node_list, sort_racks)


class Workers(Nodes):
    """Worker ids are dealt in nodes-file order; the master is worker 0, ``middle = max // 2`` and the next worker
    of the last one is worker 0 (``Workers.java:118-155``)."""

    def __init__(self, nodes: Nodes, self_id: int = UNKNOWN_WORKER_ID, port_base: int = DEFAULT_WORKER_PORT_BASE):
        super().__init__(nodes.get_nodes(), nodes.get_rack_list())
        self._infos: Dict[int, WorkerInfo] = {}
        self._rack_workers: Dict[int, List[int]] = {}
        wid = -1
        for rack in self.get_rack_list():
            ids = self._rack_workers.setdefault(rack, [])
            for host in self.get_nodes()[rack]:
                wid += 1
                self._infos[wid] = WorkerInfo(wid, host, port_base + wid, rack)
                ids.append(wid)
        self._self = int(self_id)
        self._master, self._min, self._max = 0, 0, wid
        self._middle = wid // 2 if wid >= 0 else 0
        self._next = self._self + 1 if 0 <= self._self < self._max else 0

    @classmethod
    def from_file(cls, path: str, self_id: int = UNKNOWN_WORKER_ID) -> "Workers":   # type: ignore[override]
        return cls(Nodes.from_file(path), self_id)

    @classmethod
    def from_distributed(cls, group=None) -> "Workers":
        """One rack, one entry per rank of the process group (host names from the rendezvous are not needed)."""
        import torch.distributed as dist

        if dist.is_available() and dist.is_initialized():
            world, rank = dist.get_world_size(group), dist.get_rank(group)
        else:
            world, rank = 1, 0
        return cls(Nodes({0: ["127.0.0.1"] * world}), rank)

    def get_num_workers(self) -> int:
        return len(self._infos)

    def is_the_only_worker(self) -> bool:
        return len(self._infos) <= 1

    def get_master_id(self) -> int:
        return self._master

    def is_master(self) -> bool:
        return self._self == self._master

    def get_master_info(self) -> WorkerInfo:
        return self._infos[self._master]

    def get_self_id(self) -> int:
        return self._self

    def is_self_in_worker(self) -> bool:
        return 0 <= self._self <= self._max

    def get_self_info(self) -> Optional[WorkerInfo]:
        return self._infos.get(self._self)

    def is_max(self) -> bool:
        return self._self == self._max

    def get_min_id(self) -> int:
        return self._min

    def get_middle_id(self) -> int:
        return self._middle

    def get_max_id(self) -> int:
        return self._max

    def get_next_id(self) -> int:
        return self._next

    def get_next_info(self) -> WorkerInfo:
        return self._infos[self._next]

    def get_worker_info(self, worker_id: int) -> Optional[WorkerInfo]:
        return self._infos.get(int(worker_id))

    def get_worker_info_list(self) -> Iterator[WorkerInfo]:
        return iter(self._infos[i] for i in sorted(self._infos))

    def get_rack_workers(self) -> Dict[int, List[int]]:
        return self._rack_workers

    (getNumWorkers, isTheOnlyWorker, getMasterID, isMaster, getMasterInfo, getSelfID, isSelfInWorker, getSelfInfo,
     isMax, getMinID, getMiddleID, getMaxID, getNextID, getNextInfo, getWorkerInfo, getWorkerInfoList) = (
        get_num_workers, is_the_only_worker, get_master_id, is_master, get_master_info, get_self_id,
        is_self_in_worker, get_self_info, is_max, get_min_id, get_middle_id, get_max_id, get_next_id, get_next_info,
        get_worker_info, get_worker_info_list)
