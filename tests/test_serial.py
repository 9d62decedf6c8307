"""Harp wire format (reference unit tests io/SerializerTest, DeserializerTest, DataUtilTest, resource/*ArrayTest:
round trips, encoded byte counts, golden big-endian bytes)."""
import io
import struct

import pytest
import torch

from harp_b200.io import serial as S
from harp_b200.partition import Partition, Table
from harp_b200.partition import ArrayCombiner, Operation
from harp_b200.resource import (ByteArray, DoubleArray, FloatArray, IntArray, LongArray, ShortArray, Writable,
                                register_writable)


def test_primitives_are_big_endian_and_round_trip():
    out = S.Serializer()
    out.write_boolean(True); out.write_byte(-3); out.write_short(-2); out.write_char("A"); out.write_int(0x01020304)
    out.write_long(-5); out.write_float(1.5); out.write_double(-2.25); out.write_chars("hé€")
    data = out.to_bytes()
    assert data[:1] == b"\x01" and data[1:2] == b"\xfd" and data[2:4] == b"\xff\xfe" and data[4:6] == b"\x00A"
    assert data[6:10] == b"\x01\x02\x03\x04"
    assert data[-10:] == struct.pack(">i", 3) + "hé€".encode("utf-16-be")           # int32 count + UTF-16BE units
    inp = S.Deserializer(data)
    assert (inp.read_boolean(), inp.read_byte(), inp.read_short(), inp.read_char(), inp.read_int()) == (True, -3, -2, "A", 0x01020304)
    assert (inp.read_long(), inp.read_float(), inp.read_double(), inp.read_chars()) == (-5, 1.5, -2.25, "hé€")
    assert inp.get_pos() == inp.get_length() == len(data)
    with pytest.raises(EOFError):
        inp.read_byte()


@pytest.mark.parametrize("cls,tag,width,vals", [
    (ByteArray, 1, 1, [1, 2, 250]), (ShortArray, 2, 2, [-1, 300]), (IntArray, 3, 4, [7, -8, 9]),
    (FloatArray, 4, 4, [1.5, -2.0]), (LongArray, 5, 8, [2 ** 40, -3]), (DoubleArray, 6, 8, [0.25, 1e100, -1.0])])
def test_typed_arrays_tag_length_elements(cls, tag, width, vals):
    arr = cls.wrap(torch.tensor(vals, dtype=cls.dtype))
    assert S.num_encode_bytes(arr) == len(vals) * width + 5                        # DoubleArray.java:44
    data = S.encode_trans_list([arr])
    assert data[0] == tag and struct.unpack(">i", data[1:5])[0] == len(vals) and len(data) == len(vals) * width + 5
    back = S.decode_simple_list(data)
    assert len(back) == 1 and type(back[0]) is cls and torch.equal(back[0].get(), arr.get())
    if cls is DoubleArray:
        assert data[5:13] == struct.pack(">d", 0.25)


@register_writable
class _Pair(Writable):
    def __init__(self):
        self.a, self.b = 0, ""

    def write(self, out):
        s = S.Serializer(out); s.write_int(self.a); s.write_chars(self.b)

    def read(self, inp):
        d = S.Deserializer(inp.read()); self.a, self.b = d.read_int(), d.read_chars()
        inp.seek(d.get_pos())

    def clear(self):
        self.a, self.b = 0, ""


def test_writable_tag_classname_body_and_mixed_lists():
    w = _Pair(); w.a, w.b = 42, "xy"
    data = S.encode_trans_list([w, IntArray.wrap(torch.tensor([1, 2], dtype=torch.int32))])
    assert data[0] == 7 and len(data) == S.num_trans_list_bytes([w, IntArray.wrap(torch.tensor([1, 2], dtype=torch.int32))])
    objs = S.decode_simple_list(data)
    assert isinstance(objs[0], _Pair) and (objs[0].a, objs[0].b) == (42, "xy")
    assert torch.equal(objs[1].get(), torch.tensor([1, 2], dtype=torch.int32))


def test_empty_list_is_one_zero_byte_and_partition_lists_carry_ids():
    assert S.encode_trans_list([]) == b"\x00" and S.decode_simple_list(b"\x00") == [] and S.decode_partition_list(b"\x00") == []
    t = Table(0, ArrayCombiner(Operation.SUM))
    t.add_partition(Partition(5, DoubleArray.wrap(torch.tensor([1.0, 2.0], dtype=torch.float64))))
    t.add_partition(Partition(-2, DoubleArray.wrap(torch.tensor([3.0], dtype=torch.float64))))
    n, body = S.encode_table(t)
    assert n == 2 and len(body) == (2 * 8 + 5 + 4) + (8 + 5 + 4)
    parts = {p.id(): p.get().get().tolist() for p in S.decode_partition_list(body)}
    assert parts == {5: [1.0, 2.0], -2: [3.0]}
    # payload first, id last (Partition.java:74-78)
    first_len = 2 * 8 + 5
    assert struct.unpack(">i", body[first_len:first_len + 4])[0] in (5, -2)


def test_data_message_head_layout_and_round_trip():
    t = Table(0, ArrayCombiner(Operation.SUM))
    t.add_partition(Partition(3, IntArray.wrap(torch.tensor([4, 5, 6], dtype=torch.int32))))
    t.add_partition(Partition(9, IntArray.wrap(torch.tensor([7], dtype=torch.int32))))
    d = S.Data.of_table("kmeans", "allreduce-1", 2, t)
    assert d.is_data() and d.is_operation_data() and d.is_partition_data() and d.partition_id == 2
    head, body = d.encode()
    assert len(head) == 9 + (4 + 2 * 6) + (4 + 2 * 11) + 4 and head[0] == S.PARTITION_LIST
    back = S.Data.decode(head, body)
    assert (back.context_name, back.operation_name, back.worker_id, back.partition_id) == ("kmeans", "allreduce-1", 2, 2)
    assert {p.id(): p.get().get().tolist() for p in back.body} == {3: [4, 5, 6], 9: [7]}
    ev = S.Data(S.SIMPLE_LIST, "ctx", 0, [DoubleArray.wrap(torch.tensor([1.0], dtype=torch.float64))])
    assert ev.is_data() and not ev.is_operation_data() and not ev.is_partition_data()
    h2, b2 = ev.encode()
    assert len(h2) == 9 + 4 + 2 * 3
    e2 = S.Data.decode(h2, b2)
    assert e2.operation_name is None and e2.body[0].get().tolist() == [1.0]
    assert not S.Data(S.SIMPLE_LIST, "ctx", S.UNKNOWN_WORKER_ID, []).is_data()
    with pytest.raises(ValueError):
        S.Data.decode(h2, b2 + b"\x00")
