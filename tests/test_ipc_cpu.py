"""Host half of the IPC framing (no GPU): the hand-built flatbuffer metadata of libbgpu (bgpu_ipc_schema_message,
bgpu_ipc_record_batch_metadata) must be read by Arrow's own reader (pyarrow = Arrow C++, whose flatbuffer verifier
rejects malformed tables) exactly as written."""
import io

import numpy as np
import pyarrow as pa
import pytest


def test_schema_message_round_trips_every_supported_type(bb):
    from ballista_b200 import ipc

    schema = pa.schema([pa.field("a", pa.int64(), False), pa.field("s", pa.string()), pa.field("d", pa.decimal128(15, 2)),
                        pa.field("big", pa.decimal128(38, 10)), pa.field("dt", pa.date32()), pa.field("dt64", pa.date64()),
                        pa.field("f", pa.float64()), pa.field("f32", pa.float32()), pa.field("b", pa.bool_()),
                        pa.field("u8", pa.uint8()), pa.field("i8", pa.int8()), pa.field("u16", pa.uint16()),
                        pa.field("i16", pa.int16()), pa.field("u32", pa.uint32()), pa.field("i32", pa.int32(), False),
                        pa.field("u64", pa.uint64()), pa.field("bin", pa.binary()),
                        pa.field("ts", pa.timestamp("us", tz="UTC")), pa.field("tsn", pa.timestamp("ns")),
                        pa.field("a rather long column name, with punctuation: ü", pa.int64())])
    msg = ipc.schema_message(schema)
    assert msg[:4] == b"\xff\xff\xff\xff" and len(msg) % 8 == 0
    assert pa.ipc.open_stream(io.BytesIO(msg + ipc.EOS)).schema.equals(schema)
    assert pa.ipc.open_stream(io.BytesIO(ipc.schema_message(pa.schema([])) + ipc.EOS)).schema.equals(pa.schema([]))


def test_dictionary_fields_carry_their_encoding(bb):
    """Dictionary<Int32, Utf8 | Binary | Int64> fields: value type + DictionaryEncoding{id = position among the dictionary
    fields, Int32 indices}; Arrow's reader accepts the message and reports the same schema. Other index types are refused."""
    from ballista_b200 import ipc

    schema = pa.schema([("a", pa.int64()), ("id1", pa.dictionary(pa.int32(), pa.string())), ("x", pa.float64()),
                        pa.field("id3", pa.dictionary(pa.int32(), pa.binary()), False), ("n", pa.dictionary(pa.int32(), pa.int64()))])
    msg = ipc.schema_message(schema)
    assert pa.ipc.open_stream(io.BytesIO(msg + ipc.EOS)).schema.equals(schema)
    with pytest.raises(bb.BgpuError):
        ipc.schema_message(pa.schema([("d", pa.dictionary(pa.int8(), pa.string()))]))


def test_record_batch_metadata_frames_a_hand_laid_body(bb):
    from ballista_b200 import ipc

    schema = pa.schema([pa.field("a", pa.int64()), pa.field("s", pa.string()), pa.field("b", pa.bool_())])
    body = bytearray()

    def add(b):
        off = len(body)
        body.extend(b)
        body.extend(b"\0" * (-len(b) % 8))
        return off, len(b)
    buffers = [add(b""), add(np.array([1, -2, 3], dtype=np.int64).tobytes()),          # a: no nulls
               add(bytes([0b101])), add(np.array([0, 1, 1, 3], dtype=np.int32).tobytes()), add(b"xyz"),  # s: ["x", None, "yz"]
               add(bytes([0b011])), add(bytes([0b001]))]                                 # b: [True, False, None]
    md = ipc.record_batch_metadata(3, [(3, 0), (3, 1), (3, 1)], buffers, len(body))
    assert len(md) % 8 == 0
    two = ipc.schema_message(schema) + (md + bytes(body)) * 2 + ipc.EOS  # two batches in one stream
    t = pa.ipc.open_stream(io.BytesIO(two)).read_all()
    assert t.to_pydict() == {"a": [1, -2, 3] * 2, "s": ["x", None, "yz"] * 2, "b": [True, False, None] * 2}
    # metadata size depends on the schema only (the encoder reserves the hole before it knows the values)
    assert len(ipc.record_batch_metadata(10**12, [(10**12, 7)] * 3, [(2**40, 2**41)] * 7, 2**50)) == len(md)
