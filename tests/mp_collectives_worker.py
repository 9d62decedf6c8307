"""torchrun worker (gloo, CPU): -verify style checks of every table collective against a plain
model of the Java semantics (union of ids, combine on collision, owner = pid % N, rotate maps).
Mirrors edu.iu.examples.* -verify (ml/java/.../examples/AllReduce.java:43-71, Rotate.java:40-78)."""
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
GPU = os.environ.get("HARP_MP_TESTS_GPU", "0") == "1"
if not GPU:
    os.environ["HARP_FORCE_CPU"] = "1"
import torch
import torch.distributed as dist

from harp_b200.mapper import CollectiveMapper, Event, EventType, init_distributed
from harp_b200.partition import (ArrayCombiner, DoubleArrPlus, FnCombiner, Operation, Partition, Partitioner, Table)
from harp_b200.keyval import Int2IntKVTable, Key2ValKVTable, TypeCombiner
from harp_b200.resource import DoubleArray, Writable, register_writable


@register_writable
class HarpString(Writable):
    def __init__(self, s=""):
        self.s = s
    def write(self, out):
        b = self.s.encode(); out.write(len(b).to_bytes(4, "little")); out.write(b)
    def read(self, inp):
        n = int.from_bytes(inp.read(4), "little"); self.s = inp.read(n).decode()
    def clear(self):
        self.s = ""


class Verify(CollectiveMapper):
    def map_collective(self, reader, ctx):
        N, me = self.get_num_workers(), self.get_self_id()
        T = lambda: Table(0, DoubleArrPlus())
        vec = lambda v, n=8: torch.full((n,), float(v), dtype=torch.float64)

        # ---- allreduce: same ids everywhere (dense path) ----
        t = T()
        for pid in range(4):
            t.add_partition(Partition(pid, vec(me + pid)))
        assert self.allreduce("v", "allreduce-1", t)
        for pid in range(4):
            exp = sum(r + pid for r in range(N))
            assert torch.equal(t.get_partition(pid).get(), vec(exp)), (pid, t.get_partition(pid).get())
        # repeated allreduce of ones == N^iter (examples/AllReduce.java verify mode)
        t = T(); t.add_partition(Partition(0, vec(1)))
        for it in range(3):
            self.allreduce("v", f"allreduce-ones-{it}", t)
        assert torch.equal(t.get_partition(0).get(), vec(N ** 3))
        # ---- allreduce with disjoint ids behaves like allgather ----
        t = T(); t.add_partition(Partition(100 + me, vec(me)))
        self.allreduce("v", "allreduce-disjoint", t)
        assert sorted(t.get_partition_ids()) == [100 + r for r in range(N)]
        # ---- allreduce MAX / MIN / MULTIPLY, unequal partition sizes (generic path) ----
        for op, fn in ((Operation.MAX, max), (Operation.MIN, min)):
            t = Table(0, ArrayCombiner(op))
            t.add_partition(Partition(0, vec(me, 5))); t.add_partition(Partition(1, vec(-me, 3)))
            self.allreduce("v", f"allreduce-{op.name}", t)
            assert torch.equal(t.get_partition(0).get(), vec(fn(range(N)), 5))
            assert torch.equal(t.get_partition(1).get(), vec(fn(-r for r in range(N)), 3))
        # ---- custom combiner: string concatenation (contrib/.../svm/HarpStringPlus.java) ----
        def concat(a, b):
            a.s = a.s + b.s
        t = Table(0, FnCombiner(concat)); t.add_partition(Partition(0, HarpString(f"w{me};")))
        self.allreduce("v", "allreduce-string", t)
        got = t.get_partition(0).get().s
        assert sorted(x for x in got.split(";") if x) == [f"w{r}" for r in range(N)], got
        # ---- broadcast ----
        t = T()
        if me == N - 1:
            for pid in range(3):
                t.add_partition(Partition(pid, vec(pid + 10)))
        self.broadcast("v", "bcast", t, N - 1, False)
        assert sorted(t.get_partition_ids()) == [0, 1, 2] and torch.equal(t.get_partition(2).get(), vec(12))
        # ---- reduce: root combines, others released ----
        t = T(); t.add_partition(Partition(0, vec(me + 1))); t.add_partition(Partition(me + 50, vec(7)))
        self.reduce("v", "reduce", t, 0)
        if me == 0:
            assert torch.equal(t.get_partition(0).get(), vec(sum(r + 1 for r in range(N))))
            assert sorted(t.get_partition_ids()) == [0] + [50 + r for r in range(N)]
        else:
            assert t.is_empty()
        # ---- allgather ----
        t = T(); t.add_partition(Partition(me, vec(me * 2)))
        self.allgather("v", "allgather", t)
        assert sorted(t.get_partition_ids()) == list(range(N))
        for r in range(N):
            assert torch.equal(t.get_partition(r).get(), vec(r * 2))
        # ---- regroup: pid -> pid % N, collisions combine ----
        t = T()
        for pid in range(2 * N):
            t.add_partition(Partition(pid, vec(1)))
        self.regroup("v", "regroup", t, Partitioner(N))
        assert sorted(t.get_partition_ids()) == [me, me + N], t.get_partition_ids()
        assert torch.equal(t.get_partition(me).get(), vec(N))
        # negative ids: owner = |pid % N| (Partitioner.java:36-43)
        t = T(); t.add_partition(Partition(-(me + 1), vec(1)))
        self.regroup("v", "regroup-neg", t, Partitioner(N))
        for pid in t.get_partition_ids():
            assert abs(pid) % N == me
        # ---- aggregate = regroup + fn + allgather ----
        t = T()
        for pid in range(N):
            t.add_partition(Partition(pid, vec(me + 1)))
        def halve(p):
            p.get().mul_(0.5)
        self.aggregate("v", "aggregate", t, Partitioner(N), halve)
        tot = sum(r + 1 for r in range(N))
        for pid in range(N):
            assert torch.equal(t.get_partition(pid).get(), vec(tot * 0.5))
        # ---- rotate: ring and explicit map ----
        t = T(); t.add_partition(Partition(me, vec(me)))
        self.rotate("v", "rotate-ring", t, None)
        assert t.get_partition_ids() == [(me - 1) % N]
        rmap = {r: (r + 2) % N for r in range(N)}
        self.rotate("v", "rotate-map", t, rmap)
        assert t.get_partition_ids() == [(me - 3) % N], t.get_partition_ids()
        # N ring rotations bring everything home (examples/Rotate.java verify)
        t = T(); t.add_partition(Partition(me, vec(me)))
        for j in range(N):
            self.rotate("v", f"rotate-home-{j}", t, None)
        assert t.get_partition_ids() == [me]
        # ---- push / pull ----
        local = T()
        for pid in range(N + 1):
            local.add_partition(Partition(pid, vec(1)))
        glob = T()
        self.push("v", "push", local, glob, Partitioner(N))
        assert local.get_num_partitions() == N + 1                      # push leaves local untouched
        for pid in glob.get_partition_ids():
            assert pid % N == me and torch.equal(glob.get_partition(pid).get(), vec(N))
        for p in local.get_partitions():
            p.get().zero_()
        self.pull("v", "pull", local, glob, True)
        for pid in range(N + 1):
            assert torch.equal(local.get_partition(pid).get(), vec(N)), pid
        # second push: ids already owned stay with their owner even under another partitioner
        class Shifted(Partitioner):
            def get_worker_id(self, pid):
                return (pid + 1) % self.num_workers
        self.push("v", "push2", local, glob, Shifted(N))
        for pid in glob.get_partition_ids():
            assert pid % N == me
        # ---- join ----
        static = T(); static.add_partition(Partition(me, vec(0)))
        dyn = T()
        for pid in range(N):
            dyn.add_partition(Partition(pid, vec(1)))
        self.join("v", "join", dyn, static)
        assert dyn.get_partition_ids() == [me] and torch.equal(dyn.get_partition(me).get(), vec(N))
        # ---- key-value tables + groupByKey (word count) ----
        kv = Int2IntKVTable(0, TypeCombiner("sum"))
        for key in range(10):
            kv.add_key_val(key, me + 1)
        self.allreduce("v", "kv-allreduce", kv)
        assert kv.get_val(3) == sum(r + 1 for r in range(N))
        wc = Key2ValKVTable(0, lambda a, b: a + b, num_partitions=8)
        for w in ("harp", "b200", f"only{me}"):
            wc.add_key_val(w, 1)
        self.group_by_key("v", "wordcount", wc, Partitioner(N))
        for pid in wc.get_partition_ids():
            assert pid % N == me
        allc = {}
        for d in self.comm.all_gather_object(dict(wc.items())):
            allc.update(d)
        assert allc["harp"] == N and allc["b200"] == N and allc[f"only{me}"] == 1
        # ---- events ----
        self.barrier("v", "pre-events")
        self.send_event(Event(EventType.MESSAGE_EVENT, "v", me, (me + 1) % N, f"hello from {me}"))
        ev = self.wait_event(30)
        assert ev is not None and ev.get_body() == f"hello from {(me - 1) % N}", ev
        self.send_event(Event(EventType.LOCAL_EVENT, "v", me, me, 42))
        assert self.wait_event(5).get_body() == 42
        self.barrier("v", "mid-events")
        if me == 0:
            self.send_event(Event(EventType.COLLECTIVE_EVENT, "v", 0, -1, "to-all"))
        else:
            assert self.wait_event(30).get_body() == "to-all"
        assert self.get_event() is None
        self.barrier("v", "done")
        # identity helpers
        assert self.get_master_id() == 0 and self.get_min_id() == 0 and self.get_max_id() == N - 1
        assert self.is_master() == (me == 0)


def main():
    rank, world, dev = init_distributed("nccl" if GPU else "gloo")
    Verify().run([])
    if world > 1:
        dist.barrier()
    if rank == 0:
        print("COLLECTIVES_OK", world, flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
