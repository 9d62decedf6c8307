"""Harp's byte format: ``Serializer`` / ``Deserializer`` and the list encoders of ``DataUtil``.

Reference: ``core/harp-collective/src/main/java/edu/iu/harp/io/{Serializer,Deserializer,DataUtil,DataType}.java``,
``resource/{Array,DoubleArray,...,Writable}.java``, ``partition/Partition.java:60-78``.

Nothing in this build sends these bytes between workers (tensors travel as raw device memory, metadata as pickles
— see ``collective.Comm.exchange``); the module exists so that payloads can be exchanged with, or archived for, a
Java Harp peer, and so that the reference's (de)serialisation unit tests have a counterpart.  Format facts
(SURVEY §9.4): big-endian everywhere; a string is an int32 character count followed by UTF-16BE code units; a typed
array is a tag byte (1..6), an int32 length and the elements; a ``Writable`` is tag 7, its class name and its own
``write`` body; a partition is its payload followed by the int32 partition id; an empty list is a single 0 byte.
"""
from __future__ import annotations

import io
import struct
from typing import Any, List, Optional, Sequence, Tuple

import numpy as np
import torch

from ..partition import Partition
from ..resource import (Array, ByteArray, DoubleArray, FloatArray, IntArray, LongArray, ShortArray, Writable,
                        writable_class, writable_name)

UNKNOWN_DATA_TYPE, BYTE_ARRAY, SHORT_ARRAY, INT_ARRAY, FLOAT_ARRAY, LONG_ARRAY, DOUBLE_ARRAY, WRITABLE, SIMPLE_LIST, \
    PARTITION_LIST = range(10)

_ARRAY_BY_TAG = {BYTE_ARRAY: (ByteArray, ">i1"), SHORT_ARRAY: (ShortArray, ">i2"), INT_ARRAY: (IntArray, ">i4"),
                 FLOAT_ARRAY: (FloatArray, ">f4"), LONG_ARRAY: (LongArray, ">i8"), DOUBLE_ARRAY: (DoubleArray, ">f8")}
_TAG_BY_DTYPE = {torch.uint8: BYTE_ARRAY, torch.int8: BYTE_ARRAY, torch.int16: SHORT_ARRAY, torch.int32: INT_ARRAY,
                 torch.float32: FLOAT_ARRAY, torch.int64: LONG_ARRAY, torch.float64: DOUBLE_ARRAY}


class Serializer:
    """``java.io.DataOutput`` over a growing byte buffer (``Serializer.java``): every method appends big-endian."""

    def __init__(self, out: Optional[io.BytesIO] = None):
        self._out = out if out is not None else io.BytesIO()

    def get_pos(self) -> int:
        return self._out.tell()

    def to_bytes(self) -> bytes:
        return self._out.getvalue()

    def write(self, b: bytes) -> None:
        self._out.write(b)

    def write_boolean(self, v: bool) -> None:
        self._out.write(b"\x01" if v else b"\x00")

    def write_byte(self, v: int) -> None:
        self._out.write(struct.pack(">b", ((int(v) + 128) % 256) - 128))

    def write_short(self, v: int) -> None:
        self._out.write(struct.pack(">h", int(v)))

    def write_char(self, v) -> None:
        self._out.write(struct.pack(">H", ord(v) if isinstance(v, str) else int(v)))

    def write_int(self, v: int) -> None:
        self._out.write(struct.pack(">i", int(v)))

    def write_long(self, v: int) -> None:
        self._out.write(struct.pack(">q", int(v)))

    def write_float(self, v: float) -> None:
        self._out.write(struct.pack(">f", float(v)))

    def write_double(self, v: float) -> None:
        self._out.write(struct.pack(">d", float(v)))

    def write_chars(self, s: str) -> None:
        """``Serializer.writeChars`` (``:223-241``): int32 number of UTF-16 code units, then the units."""
        units = s.encode("utf-16-be")
        self.write_int(len(units) // 2)
        self._out.write(units)

    write_utf = write_chars
    (getPos, writeBoolean, writeByte, writeShort, writeChar, writeInt, writeLong, writeFloat, writeDouble, writeChars,
     writeUTF) = (get_pos, write_boolean, write_byte, write_short, write_char, write_int, write_long, write_float,
                  write_double, write_chars, write_chars)


class Deserializer:
    """``java.io.DataInput`` over bytes (``Deserializer.java``); reading past the end raises ``EOFError``."""

    def __init__(self, data: bytes, start: int = 0, length: Optional[int] = None):
        self._data = memoryview(data)
        self._pos = int(start)
        self._end = len(data) if length is None else int(start) + int(length)

    def get_pos(self) -> int:
        return self._pos

    def get_length(self) -> int:
        return self._end

    def _take(self, n: int) -> memoryview:
        if self._pos + n > self._end:
            raise EOFError(f"need {n} bytes at {self._pos}, buffer ends at {self._end}")
        v = self._data[self._pos:self._pos + n]
        self._pos += n
        return v

    def read(self, n: int) -> bytes:
        return bytes(self._take(n))

    def read_boolean(self) -> bool:
        return self._take(1)[0] != 0

    def read_byte(self) -> int:
        return struct.unpack(">b", self._take(1))[0]

    def read_short(self) -> int:
        return struct.unpack(">h", self._take(2))[0]

    def read_char(self) -> str:
        return chr(struct.unpack(">H", self._take(2))[0])

    def read_int(self) -> int:
        return struct.unpack(">i", self._take(4))[0]

    def read_long(self) -> int:
        return struct.unpack(">q", self._take(8))[0]

    def read_float(self) -> float:
        return struct.unpack(">f", self._take(4))[0]

    def read_double(self) -> float:
        return struct.unpack(">d", self._take(8))[0]

    def read_chars(self) -> str:
        n = self.read_int()
        return bytes(self._take(2 * n)).decode("utf-16-be")

    read_utf = read_chars
    (getPos, getLength, readBoolean, readByte, readShort, readChar, readInt, readLong, readFloat, readDouble,
     readChars, readUTF) = (get_pos, get_length, read_boolean, read_byte, read_short, read_char, read_int, read_long,
                            read_float, read_double, read_chars, read_chars)


# --------------------------------------------------------------------------------------------------------------
# Simple objects: typed arrays and writables
# --------------------------------------------------------------------------------------------------------------
def _utf16_len(s: str) -> int:
    """``String.length()`` of Java: UTF-16 code units (characters outside the BMP count twice)."""
    return len(s.encode("utf-16-be")) // 2


def _tensor_of(obj: Any) -> Optional[torch.Tensor]:
    if isinstance(obj, Array):
        return obj.get()
    return obj if isinstance(obj, torch.Tensor) else None


def num_encode_bytes(obj: Any) -> int:
    """``getNumEnocdeBytes``: arrays ``size * width + 5`` (``DoubleArray.java:44``), writables tag + name + body,
    partitions 4 more (``Partition.java:60-66``)."""
    if isinstance(obj, Partition):
        return 4 + num_encode_bytes(obj.get())
    t = _tensor_of(obj)
    if t is not None:
        return t.numel() * t.element_size() + 5
    if isinstance(obj, Writable):
        return 1 + 4 + 2 * _utf16_len(writable_name(obj)) + obj.get_num_write_bytes()
    raise TypeError(f"not a Harp Simple object: {type(obj)!r}")


def encode_simple(obj: Any, out: Serializer) -> None:
    if isinstance(obj, Partition):
        encode_simple(obj.get(), out)
        out.write_int(obj.id())
        return
    t = _tensor_of(obj)
    if t is not None:
        tag = _TAG_BY_DTYPE.get(t.dtype)
        if tag is None or t.dim() != 1:
            raise TypeError(f"no Harp array type for a {t.dim()}-d {t.dtype} tensor")
        out.write_byte(tag)
        out.write_int(t.numel())
        arr = t.detach().cpu().contiguous().numpy()
        if arr.dtype == np.uint8:
            arr = arr.view(np.int8)
        out.write(arr.astype(_ARRAY_BY_TAG[tag][1], copy=False).tobytes())
        return
    if isinstance(obj, Writable):
        out.write_byte(WRITABLE)
        out.write_chars(writable_name(obj))
        body = io.BytesIO()
        obj.write(body)
        out.write(body.getvalue())
        return
    raise TypeError(f"not a Harp Simple object: {type(obj)!r}")


def decode_simple(inp: Deserializer, tag: Optional[int] = None) -> Any:
    """One typed array (returned as the matching ``*Array``) or one registered ``Writable``."""
    if tag is None:
        tag = inp.read_byte()
    if tag in _ARRAY_BY_TAG:
        cls, fmt = _ARRAY_BY_TAG[tag]
        n = inp.read_int()
        width = np.dtype(fmt).itemsize
        arr = np.frombuffer(inp.read(n * width), dtype=fmt).astype(np.dtype(fmt).newbyteorder("="))
        t = torch.from_numpy(arr.copy())
        if cls is ByteArray and t.dtype != ByteArray.dtype:
            t = t.view(ByteArray.dtype)
        return cls.wrap(t)
    if tag == WRITABLE:
        cls = writable_class(inp.read_chars())
        obj = cls.create()
        # Writable.read consumes exactly what write produced; hand it a stream positioned at the body
        stream = io.BytesIO(bytes(inp._data[inp._pos:inp._end]))
        obj.read(stream)
        inp._pos += stream.tell()
        return obj
    raise ValueError(f"unknown Harp data type tag {tag}")


# --------------------------------------------------------------------------------------------------------------
# Lists (DataUtil.encodeTransList / decodeSimpleList / decodePartitionList)
# --------------------------------------------------------------------------------------------------------------
def num_trans_list_bytes(objs: Sequence[Any]) -> int:
    return sum(num_encode_bytes(o) for o in objs) or 1


def encode_trans_list(objs: Sequence[Any]) -> bytes:
    """A list of Simple objects or of Partitions; the empty list is the single byte 0 (``DataUtil.java:452-463``)."""
    out = Serializer()
    if not objs:
        out.write_byte(UNKNOWN_DATA_TYPE)
    else:
        for o in objs:
            encode_simple(o, out)
    data = out.to_bytes()
    assert len(data) == num_trans_list_bytes(objs)
    return data


def decode_simple_list(data: bytes) -> List[Any]:
    inp = Deserializer(data)
    objs: List[Any] = []
    while inp.get_pos() < inp.get_length():
        tag = inp.read_byte()
        if tag == UNKNOWN_DATA_TYPE:
            break
        objs.append(decode_simple(inp, tag))
    return objs


def decode_partition_list(data: bytes) -> List[Partition]:
    inp = Deserializer(data)
    parts: List[Partition] = []
    while inp.get_pos() < inp.get_length():
        tag = inp.read_byte()
        if tag == UNKNOWN_DATA_TYPE:
            break
        payload = decode_simple(inp, tag)
        parts.append(Partition(inp.read_int(), payload))
    return parts


def encode_table(table) -> Tuple[int, bytes]:
    """``(number of partitions, body bytes)`` of a table — what a ``PARTITION_LIST`` message carries (its head's
    partition-id field holds the count)."""
    parts = table.get_partitions()
    return len(parts), encode_trans_list(parts)


# --------------------------------------------------------------------------------------------------------------
# Data: one message = head + body  (io/Data.java:31-133, head layout :504-569 / :391-451)
# --------------------------------------------------------------------------------------------------------------
UNKNOWN_WORKER_ID = -1
UNKNOWN_PARTITION_ID = -2147483648          # Constant.UNKNOWN_PARTITION_ID (Integer.MIN_VALUE)


class Data:
    """``bodyType | contextName | workerID | bodySize [| operationName [| partitionID]]`` followed by the body
    (a ``SIMPLE_LIST`` or a ``PARTITION_LIST``).  For partition lists the partition-id field of the head carries the
    NUMBER of partitions in the message (``PartitionUtil.java:62-70``)."""

    def __init__(self, body_type: int, context_name: str, worker_id: int, body: Sequence[Any],
                 operation_name: Optional[str] = None, partition_id: int = UNKNOWN_PARTITION_ID):
        self.body_type = int(body_type)
        self.context_name = context_name
        self.worker_id = int(worker_id)
        self.body = list(body)
        self.operation_name = operation_name
        self.partition_id = int(partition_id)

    @classmethod
    def of_table(cls, context_name: str, operation_name: str, worker_id: int, table) -> "Data":
        parts = table.get_partitions()
        return cls(PARTITION_LIST, context_name, worker_id, parts, operation_name, len(parts))

    # ``Data.isData / isOperationData / isPartitionData`` (``Data.java:264-300``)
    def is_data(self) -> bool:
        return (self.body_type in (SIMPLE_LIST, PARTITION_LIST) and self.context_name is not None
                and self.worker_id != UNKNOWN_WORKER_ID and self.body is not None)

    def is_operation_data(self) -> bool:
        return self.is_data() and self.operation_name is not None

    def is_partition_data(self) -> bool:
        return self.is_operation_data() and self.partition_id != UNKNOWN_PARTITION_ID

    def encode_body(self) -> bytes:
        return encode_trans_list(self.body)

    def encode_head(self, body_size: Optional[int] = None) -> bytes:
        if not self.is_data():
            raise ValueError("not a complete Data object")
        out = Serializer()
        out.write_byte(self.body_type)
        out.write_chars(self.context_name)
        out.write_int(self.worker_id)
        out.write_int(num_trans_list_bytes(self.body) if body_size is None else body_size)
        if self.is_operation_data():
            out.write_chars(self.operation_name)
            if self.is_partition_data():
                out.write_int(self.partition_id)
        head = out.to_bytes()
        expect = 9 + 4 + 2 * _utf16_len(self.context_name)
        if self.is_operation_data():
            expect += 4 + 2 * _utf16_len(self.operation_name) + (4 if self.is_partition_data() else 0)
        assert len(head) == expect                              # Data.java:509-519
        return head

    def encode(self) -> Tuple[bytes, bytes]:
        body = self.encode_body()
        return self.encode_head(len(body)), body

    @classmethod
    def decode(cls, head: bytes, body: bytes) -> "Data":
        inp = Deserializer(head)
        body_type, ctx, wid, size = inp.read_byte(), inp.read_chars(), inp.read_int(), inp.read_int()
        op = inp.read_chars() if inp.get_pos() < inp.get_length() else None
        pid = inp.read_int() if inp.get_pos() < inp.get_length() else UNKNOWN_PARTITION_ID
        if size != len(body):
            raise ValueError(f"head announces {size} body bytes, got {len(body)}")
        if body_type == PARTITION_LIST:
            objs: List[Any] = decode_partition_list(body)
        elif body_type == SIMPLE_LIST:
            objs = decode_simple_list(body)
        else:
            raise ValueError(f"unknown body type {body_type}")
        return cls(body_type, ctx, wid, objs, op, pid)

    isData, isOperationData, isPartitionData, encodeHead, encodeBody = (
        is_data, is_operation_data, is_partition_data, encode_head, encode_body)
