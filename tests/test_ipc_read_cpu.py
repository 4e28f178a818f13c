"""Host half of the IPC reader (no GPU): the stream / flatbuffer walker must read schemas written by Arrow's own writer
(pyarrow = Arrow C++), and the LZ4 frame walker + block decoder the device kernel is built from (run here on the host with
one lane through the library's test hook) must invert Arrow's own LZ4_FRAME codec."""
import io

import numpy as np
import pyarrow as pa
import pytest

from helpers import independent_frame, reference_style_lz4_stream


def _stream(schema, batches=(), compression=None):
    sink = io.BytesIO()
    with pa.ipc.new_stream(sink, schema, options=pa.ipc.IpcWriteOptions(compression=compression)) as w:
        for b in batches:
            w.write_batch(b)
    return sink.getvalue()


def test_reader_reads_arrow_written_schemas(bb):
    from ballista_b200 import ipc

    schema = pa.schema([pa.field("a", pa.int64(), False), pa.field("s", pa.string()), pa.field("d", pa.decimal128(15, 2)),
                        pa.field("big", pa.decimal128(38, 10)), pa.field("dt", pa.date32()), pa.field("dt64", pa.date64()),
                        pa.field("f", pa.float64()), pa.field("f32", pa.float32()), pa.field("b", pa.bool_()),
                        pa.field("u8", pa.uint8()), pa.field("i8", pa.int8()), pa.field("u16", pa.uint16()),
                        pa.field("i16", pa.int16()), pa.field("u32", pa.uint32()), pa.field("i32", pa.int32(), False),
                        pa.field("u64", pa.uint64()), pa.field("bin", pa.binary()),
                        pa.field("ts", pa.timestamp("us", tz="UTC")), pa.field("tsn", pa.timestamp("ns")),
                        pa.field("t32", pa.time32("ms")), pa.field("t64", pa.time64("ns")),
                        pa.field("a rather long column name, with punctuation: ü", pa.int64())])
    host = ipc._HostContext()
    r = ipc.StreamReader(host, _stream(schema))
    assert r.schema.equals(schema)
    # our own schema message reads back too
    r2 = ipc.StreamReader(host, ipc.schema_message(schema) + ipc.EOS)
    assert r2.schema.equals(schema)
    assert ipc.StreamReader(host, _stream(pa.schema([]))).schema.equals(pa.schema([]))


def test_reader_rejects_malformed_streams(bb):
    from ballista_b200 import ipc

    host = ipc._HostContext()
    with pytest.raises(bb.BgpuError, match="empty stream"):
        ipc.StreamReader(host, b"")
    with pytest.raises(bb.BgpuError, match="beyond the stream"):
        ipc.StreamReader(host, b"\xff\xff\xff\xff\x00\x10\x00\x00abc")
    schema = pa.schema([("a", pa.int64())])
    batch = pa.RecordBatch.from_arrays([pa.array([1, 2, 3], type=pa.int64())], schema=schema)
    data = _stream(schema, [batch])
    first = 8 + int.from_bytes(data[4:8], "little")
    with pytest.raises(bb.BgpuError, match="Unable to read IPC message as schema"):
        ipc.StreamReader(host, data[first:])  # starts with a RecordBatch message
    # Dictionary<Int32, T> fields are read (round 2); other index types are not
    dsch = pa.schema([("d", pa.dictionary(pa.int32(), pa.string())), ("n", pa.dictionary(pa.int32(), pa.int64()))])
    assert ipc.StreamReader(host, _stream(dsch)).schema.equals(dsch)
    with pytest.raises(bb.BgpuError) as e:
        ipc.StreamReader(host, _stream(pa.schema([("d", pa.dictionary(pa.int16(), pa.string()))])))
    assert "dictionary" in str(e.value)
    with pytest.raises(bb.BgpuError):
        ipc.StreamReader(host, _stream(pa.schema([("l", pa.list_(pa.int64()))])))


def test_reader_next_fails_loudly_without_a_device(bb):
    import torch

    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from ballista_b200 import ipc

    schema = pa.schema([("a", pa.int64())])
    batch = pa.RecordBatch.from_arrays([pa.array([1, 2, 3], type=pa.int64())], schema=schema)
    r = ipc.StreamReader(ipc._HostContext(), _stream(schema, [batch]))
    with pytest.raises(bb.BgpuError, match="no CPU fallback"):
        r.maybe_next()


def _cases():
    rng = np.random.default_rng(11)
    yield "empty", b""
    yield "one byte", b"x"
    yield "short text", b"hello hello hello hello hello"
    yield "run (overlapping match, offset 1)", b"a" * 100_000
    yield "period 3", b"abc" * 70_000
    yield "incompressible 1 KB", rng.integers(0, 256, 1000, dtype=np.uint8).tobytes()
    yield "incompressible 300 KB (stored blocks)", rng.integers(0, 256, 300_000, dtype=np.uint8).tobytes()
    yield "int64 column 2 MB", rng.integers(0, 1000, 250_000).astype(np.int64).tobytes()
    yield "decimal128 column 5 MB (several blocks)", np.stack([rng.integers(0, 10_000_000, 320_000), np.zeros(320_000, dtype=np.int64)], axis=1).astype(np.int64).tobytes()
    yield "mixed", (b"".join(bytes([i % 251]) * (i % 40) for i in range(20_000)) + rng.integers(0, 4, 100_000, dtype=np.uint8).tobytes())
    yield "long literal run then long match", rng.integers(0, 256, 5000, dtype=np.uint8).tobytes() * 3


@pytest.mark.parametrize("name,data", list(_cases()), ids=[n for n, _ in _cases()])
def test_lz4_frame_decoder_inverts_arrows_codec(bb, name, data):
    from ballista_b200 import ipc

    frame = pa.compress(data, codec="lz4", asbytes=True)
    assert ipc.lz4_frame_decode_host(frame, len(data)) == data


def test_lz4_frame_decoder_rejects_corruption(bb):
    from ballista_b200 import ipc

    data = b"abcdefgh" * 1000
    frame = bytearray(pa.compress(data, codec="lz4", asbytes=True))
    with pytest.raises(bb.BgpuError, match="bad magic"):
        ipc.lz4_frame_decode_host(b"\0\0\0\0" + bytes(frame[4:]), len(data))
    with pytest.raises(bb.BgpuError):
        ipc.lz4_frame_decode_host(bytes(frame), len(data) + 1)  # wrong uncompressed length
    with pytest.raises(bb.BgpuError):
        ipc.lz4_frame_decode_host(bytes(frame[: len(frame) // 2]), len(data))  # truncated


@pytest.mark.parametrize("block_checksum,content_size", [(False, False), (True, True)])
def test_lz4_block_independent_frames(bb, block_checksum, content_size):
    from ballista_b200 import ipc

    rng = np.random.default_rng(5)
    data = (rng.integers(0, 50, 200_000).astype(np.int64).tobytes() + rng.integers(0, 256, 70_000, dtype=np.uint8).tobytes()
            + b"xyz" * 33_333)
    frame = independent_frame(data, block_checksum=block_checksum, content_size=content_size)
    assert ipc.lz4_frame_decode_host(frame, len(data)) == data


def test_reference_style_lz4_stream_is_valid_arrow_ipc(bb):
    """The hand-framed stream the GPU reader tests use (buffers as block-independent LZ4 frames behind their int64 length
    prefix, BodyCompression in libbgpu's RecordBatch metadata) is read by Arrow's own reader exactly as written."""
    from helpers import random_batch

    rng = np.random.default_rng(8)
    batches = [random_batch(rng, n, with_dict=False) for n in (5, 0, 70_000)]
    data = reference_style_lz4_stream(batches)
    got = list(pa.ipc.open_stream(io.BytesIO(data)))
    assert len(got) == 3
    for g, w in zip(got, batches):
        assert g.equals(w)
