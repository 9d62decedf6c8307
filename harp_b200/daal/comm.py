"""``HarpDAALComm`` (``core/harp-daal-interface/src/main/java/edu/iu/data_comm/HarpDAALComm.java``).

The reference ships Java-serialised DAAL objects through Harp: ``harpdaal_braodcast`` = barrier +
broadcast + barrier (``:78-116``), ``harpdaal_gather`` = ``reduce`` of a table whose partition ids are
the worker ids (``:203-246``), ``harpdaal_allgather`` = ``allreduce`` of such a table (``:296-337``).
The same three calls are kept (arbitrary picklable objects or tensors); new code should prefer
``linalg._allreduce_`` on fixed-shape partial results, which is what the algorithms here do.
"""
from __future__ import annotations

from typing import Any, List, Optional

import torch

from ..partition import FnCombiner, Partition, Table


class HarpDAALComm:
    def __init__(self, self_id: int, master_id: int, num_workers: int, mapper, daal_context: Any = None):
        self.self_id, self.master_id, self.num_workers, self.mapper = self_id, master_id, num_workers, mapper

    def harpdaal_broadcast(self, obj: Any, context_name: str, operation_name: str) -> Any:
        m = self.mapper
        m.barrier(context_name, operation_name + "-pre")
        t = Table(0, FnCombiner(lambda a, b: a))
        if m.get_self_id() == self.master_id:
            t.add_partition(Partition(self.master_id, obj))
        m.broadcast(context_name, operation_name, t, self.master_id, False)
        m.barrier(context_name, operation_name + "-post")
        return t.get_partition(self.master_id).get()

    # the reference really spells it this way
    harpdaal_braodcast = harpdaal_broadcast

    def harpdaal_gather(self, obj: Any, master_id: Optional[int], context_name: str, operation_name: str) -> Optional[List[Any]]:
        root = self.master_id if master_id is None else master_id
        m = self.mapper
        t = Table(0, FnCombiner(lambda a, b: a))
        t.add_partition(Partition(m.get_self_id(), obj))
        m.barrier(context_name, operation_name + "-pre")
        m.reduce(context_name, operation_name, t, root)
        m.barrier(context_name, operation_name + "-post")
        if m.get_self_id() != root:
            return None
        return [t.get_partition(i).get() for i in sorted(t.get_partition_ids())]

    def harpdaal_allgather(self, obj: Any, context_name: str, operation_name: str) -> List[Any]:
        m = self.mapper
        t = Table(0, FnCombiner(lambda a, b: a))
        t.add_partition(Partition(m.get_self_id(), obj))
        m.allreduce(context_name, operation_name, t)      # disjoint ids: behaves as allgather
        return [t.get_partition(i).get() for i in sorted(t.get_partition_ids())]
