"""Unit tests of the L1 data model — mirror core/harp-collective/src/test/java/edu/iu/harp/
{combiner,partition,resource,keyval}/*Test.java."""
import io

import pytest
import torch

from harp_b200.keyval import (Int2IntKVTable, Int2LongKVTable, Key2ValKVTable, KVPartition, Long2DoubleKVTable,
                              Long2IntKVTable, TypeCombiner, segment_combine)
from harp_b200.partition import (ArrayCombiner, DoubleArrPlus, FnCombiner, LogSumExpCombiner, MappedPartitioner,
                                 NeverCombine, Operation, Partition, Partitioner, PartitionStatus,
                                 RandomPartitioner, Table, add_partitions_to_table)
from harp_b200.resource import (Array, ByteArray, DoubleArray, FloatArray, IntArray, LongArray, ResourcePool,
                                ShortArray, Writable, decode_payload, encode_payload, register_writable)


@pytest.mark.parametrize("dtype", [torch.uint8, torch.int16, torch.int32, torch.int64, torch.float32, torch.float64])
def test_array_combiner_ops(dtype):
    a = torch.tensor([1, 2, 3, 4], dtype=dtype)
    b = torch.tensor([4, 3, 2, 1], dtype=dtype)
    exp = {Operation.SUM: [5, 5, 5, 5], Operation.MULTIPLY: [4, 6, 6, 4], Operation.MAX: [4, 3, 3, 4],
           Operation.MIN: [1, 2, 2, 1]}
    for op, want in exp.items():
        cur = a.clone()
        assert ArrayCombiner(op).combine(cur, b) is PartitionStatus.COMBINED
        assert cur.tolist() == want
    if dtype != torch.uint8:
        cur = a.clone()
        ArrayCombiner(Operation.MINUS).combine(cur, b)
        assert cur.tolist() == [-3, -1, 1, 3]


def test_combiner_size_mismatch_fails():
    assert ArrayCombiner(Operation.SUM).combine(torch.zeros(3), torch.zeros(4)) is PartitionStatus.COMBINE_FAILED
    assert NeverCombine().combine(1, 2) is PartitionStatus.COMBINE_FAILED


def test_table_add_insert_combine_remove():
    t = Table(7, DoubleArrPlus())
    assert t.get_table_id() == 7 and t.is_empty()
    assert t.add_partition(None) is PartitionStatus.ADD_FAILED
    assert t.add_partition(Partition(3, torch.ones(4, dtype=torch.float64))) is PartitionStatus.ADDED
    assert t.add_partition(Partition(3, torch.full((4,), 2.0, dtype=torch.float64))) is PartitionStatus.COMBINED
    assert t.get_partition(3).get().tolist() == [3.0] * 4
    assert t.add_partition(Partition(3, torch.ones(5, dtype=torch.float64))) is PartitionStatus.COMBINE_FAILED
    assert t.get_num_partitions() == 1 and t.get_partition_ids() == [3]
    assert t.remove_partition(3).id() == 3 and t.is_empty()
    assert add_partitions_to_table(t, [Partition(1, torch.ones(2)), Partition(1, torch.ones(2))])
    assert t.get_partition(1).get().tolist() == [2.0, 2.0]
    t.release()
    assert t.is_empty()


def test_dense_table_views_share_slab():
    slab = torch.zeros(4, 6)
    t = Table.dense(0, DoubleArrPlus(), slab)
    assert t.is_dense() and t.get_num_partitions() == 4
    t.add_partition(Partition(2, torch.ones(6)))
    assert slab[2].tolist() == [1.0] * 6            # combine happened in place, inside the slab
    t.insert_partition(Partition(9, torch.ones(6)))
    assert not t.is_dense()


def test_fn_and_logsumexp_combiners():
    t = Table(0, FnCombiner(lambda a, b: a + b))
    t.add_partition(Partition(0, "ab")); t.add_partition(Partition(0, "cd"))
    assert t.get_partition(0).get() == "abcd"
    cur = torch.log(torch.tensor([1.0, 2.0]))
    LogSumExpCombiner().combine(cur, torch.log(torch.tensor([3.0, 4.0])))
    torch.testing.assert_close(cur.exp(), torch.tensor([4.0, 6.0]))


def test_partitioner_rules():
    p = Partitioner(4)
    assert [p.get_worker_id(i) for i in range(6)] == [0, 1, 2, 3, 0, 1]
    assert p.get_worker_id(-5) == 1 and p.get_worker_id(-4) == 0      # |pid % N| with Java remainder
    assert p.worker_ids(torch.tensor([-5, 7, 8])).tolist() == [1, 3, 0]
    assert MappedPartitioner(4, {10: 3}).get_worker_id(10) == 3
    assert MappedPartitioner(4, {10: 3}).get_worker_id(11) == 3
    r1, r2 = RandomPartitioner(8, 1234), RandomPartitioner(8, 1234)
    owners = [r1.get_worker_id(i) for i in range(1000)]
    assert owners == [r2.get_worker_id(i) for i in range(1000)]           # same seed -> same owners everywhere
    assert len(set(owners)) == 8 and min(owners.count(w) for w in range(8)) > 60


@pytest.mark.parametrize("cls,dtype,esz", [(ByteArray, torch.uint8, 1), (ShortArray, torch.int16, 2),
                                           (IntArray, torch.int32, 4), (FloatArray, torch.float32, 4),
                                           (LongArray, torch.int64, 8), (DoubleArray, torch.float64, 8)])
def test_arrays_and_pools(cls, dtype, esz):
    pool = ResourcePool.get().pool(dtype)
    pool.clean()
    a = cls.create(100, True)
    assert a.size() == 100 and a.start() == 0 and a.get().dtype == dtype and a.get().numel() == 100
    assert a.backing().numel() == 128                      # approximate -> next power of two
    assert a.get_num_encode_bytes() == 100 * esz + 5       # DoubleArray.java:44
    ptr = a.backing().data_ptr()
    a.release()
    assert pool.stats()["cached"] == 1
    b = cls.create(70, True)                               # same size class -> reused
    assert b.backing().data_ptr() == ptr
    b.free()
    assert pool.stats()["cached"] == 0 and pool.stats()["in_use"] == 0
    c = cls.create(100, False)
    assert c.backing().numel() == 100
    c.release(); ResourcePool.get().clean()
    assert pool.stats()["cached"] == 0


@register_writable
class _Pair(Writable):
    def __init__(self):
        self.a, self.b = 0, 0.0
    def write(self, out):
        import struct; out.write(struct.pack("<id", self.a, self.b))
    def read(self, inp):
        import struct; self.a, self.b = struct.unpack("<id", inp.read(12))
    def clear(self):
        self.a, self.b = 0, 0.0


def test_writable_roundtrip_and_pool():
    w = _Pair.create(); w.a, w.b = 7, 2.5
    assert w.get_num_write_bytes() == 12
    w2 = decode_payload(encode_payload(w))
    assert isinstance(w2, _Pair) and (w2.a, w2.b) == (7, 2.5)
    w.release()
    w3 = _Pair.create()
    assert w3 is w and w3.a == 0                             # pooled by class, cleared on release
    t = decode_payload(encode_payload(torch.arange(5)))
    assert t.tolist() == [0, 1, 2, 3, 4]
    arr = decode_payload(encode_payload(DoubleArray.wrap(torch.ones(3, dtype=torch.float64))))
    assert isinstance(arr, DoubleArray) and arr.get().tolist() == [1.0, 1.0, 1.0]
    log = ResourcePool.get().log()
    assert "writable" in log and "float64" in log


@pytest.mark.parametrize("cls", [Int2IntKVTable, Int2LongKVTable, Long2IntKVTable, Long2DoubleKVTable])
def test_kv_tables(cls):
    t = cls(0, TypeCombiner("sum"))
    for k in [5, 3, 5, 9, 3, 5]:
        t.add_key_val(k, 2)
    assert t.get_val(5) == 6 and t.get_val(3) == 4 and t.get_val(9) == 2 and t.get_val(77) == 0
    assert sorted(t.get_partition_ids()) == [3, 5, 9]          # partitionID = key (Int2IntKVTable.java:146-148)
    other = cls(0, TypeCombiner("sum")); other.add_key_val(5, 10); other.add_key_val(1, 1)
    for p in other.get_partitions():
        t.add_partition(p)
    assert t.get_val(5) == 16 and t.get_val(1) == 1
    part = t.get_partition(5).get()
    buf = io.BytesIO(); part.write(buf); buf.seek(0)
    q = KVPartition(); q.read(buf)
    assert dict(q.items()) == dict(part.items())
    t.add_batch(torch.tensor([5, 5, 2]), torch.tensor([1, 1, 1]))
    assert t.get_val(5) == 18 and t.get_val(2) == 1 and t.num_keys() == 5
    assert t.remove_val(2) == 1 and t.get_val(2) == 0


def test_segment_combine_and_key2val():
    k, v = segment_combine(torch.tensor([3, 1, 3, 2, 1]), torch.tensor([1.0, 2.0, 3.0, 4.0, 5.0]), "max")
    assert k.tolist() == [1, 2, 3] and v.tolist() == [5.0, 4.0, 3.0]
    wc = Key2ValKVTable(0, lambda a, b: a + b, num_partitions=4)
    for w in "a b a c a b".split():
        wc.add_key_val(w, 1)
    assert wc.get_val("a") == 3 and wc.get_val("b") == 2 and wc.get_val("zz") is None
    assert wc.remove_val("c") == 1 and dict(wc.items()) == {"a": 3, "b": 2}
