"""Interoperability with the reference's own LZ4 code, EXECUTED (tests/lz4_flex_ffi.py: lz4_flex-0.11.5 as compiled into a
Rust library of this image; the reference's shuffle files are compressed and read by that crate through arrow-ipc,
/root/reference/ballista/core/src/execution_plans/shuffle_writer.rs:365-376, shuffle_reader.rs:771-785):

  * frames whose blocks were compressed by lz4_flex's block compressor, framed the way its FrameEncoder frames them,
    are decoded exactly by libbgpu's LZ4 frame walker + block decoder (the kernel's logic, run on the host by the test
    hook) - K8 reads what the reference writes;
  * the stand-ins the rest of the suite uses for "a file the reference wrote" (helpers.independent_frame,
    helpers.reference_style_lz4_stream) are accepted by lz4_flex's FrameDecoder and give back the data - the stand-ins are
    what they claim to be.
Skipped on a machine without that library."""
import io

import numpy as np
import pyarrow as pa
import pytest

import lz4_flex_ffi as F
from helpers import _xxh32_short, independent_frame, random_batch, record_batch_message_buffers, reference_style_lz4_stream

pytestmark = pytest.mark.skipif(not F.available(), reason="lz4_flex (wandb/bin/librust_parquet_ffi.so) not in this image")


def payloads():
    rng = np.random.default_rng(4)
    text = (b"1996-03-13|DELIVER IN PERSON|TRUCK|" * 4000)
    return {"empty": b"", "one": b"x", "short": b"hello hello hello hello", "zeros": bytes(300_000),
            "low_entropy": bytes(rng.integers(0, 4, 200_000, dtype=np.uint8)), "text": text,
            "random": bytes(rng.integers(0, 256, 70_000, dtype=np.uint8)),
            "int64_runs": (np.arange(40_000, dtype=np.int64) // 50).tobytes()}


def lz4_flex_frame(data, block=65536):
    """FrameEncoder's default layout (block-independent 64 KB blocks, no checksums but the header's) around blocks
    compressed by lz4_flex itself; a block that does not shrink is stored."""
    out = bytearray(b"\x04\x22\x4d\x18" + bytes([0x60, 0x40]))
    out += bytes([(_xxh32_short(bytes(out[4:])) >> 8) & 0xFF])
    for lo in range(0, len(data), block):
        piece = data[lo:lo + block]
        comp = F.block_compress(piece)
        assert F.block_decompress(comp, len(piece)) == piece
        if len(comp) >= len(piece):
            out += (len(piece) | 0x80000000).to_bytes(4, "little") + piece
        else:
            out += len(comp).to_bytes(4, "little") + comp
    return bytes(out + b"\0\0\0\0")


@pytest.mark.parametrize("name", list(payloads()))
def test_our_decoder_reads_blocks_compressed_by_lz4_flex(bb, name):
    from ballista_b200 import ipc

    data = payloads()[name]
    frame = lz4_flex_frame(data)
    assert F.frame_decompress(frame) == data            # the frame is one lz4_flex itself accepts
    assert ipc.lz4_frame_decode_host(frame, len(data)) == data


@pytest.mark.parametrize("name", list(payloads()))
def test_stand_in_frames_are_valid_for_lz4_flex(name):
    data = payloads()[name]
    assert F.frame_decompress(independent_frame(data)) == data
    assert F.frame_decompress(independent_frame(data, content_size=True)) == data
    assert F.frame_decompress(pa.compress(data, codec="lz4", asbytes=True)) == data  # Arrow C++'s block-linked frames too


def test_reference_style_stream_buffers_decode_with_lz4_flex():
    """Every compressed buffer of the stand-in shuffle file decodes with lz4_flex to the bytes Arrow's reader gives."""
    rng = np.random.default_rng(9)
    batch = random_batch(rng, 50_000, with_dict=False)
    stream = reference_style_lz4_stream([batch])
    assert pa.ipc.open_stream(io.BytesIO(stream)).read_all().to_batches()[0].equals(batch)
    # walk the stream: schema message, then the record batch message
    reader = pa.ipc.MessageReader.open_stream(io.BytesIO(stream))
    reader.read_next_message()
    msg = reader.read_next_message()
    compressed, bufs = record_batch_message_buffers(msg.serialize().to_pybytes())
    assert compressed
    seen = 0
    for b in bufs:
        if not b:
            continue
        raw_len = int.from_bytes(b[:8], "little", signed=True)
        if raw_len == -1:
            continue
        got = F.frame_decompress(b[8:])
        assert len(got) == raw_len
        seen += 1
    assert seen >= 5


def test_message_walk_on_arrow_written_messages():
    """The same walk the GPU test does over the device's messages, here over Arrow C++'s: every compressed buffer of an
    LZ4_FRAME RecordBatch message decodes with lz4_flex to the buffer of the uncompressed message."""
    rng = np.random.default_rng(12)
    batch = random_batch(rng, 20_000, with_dict=False)

    def message(compression):
        sink = io.BytesIO()
        with pa.ipc.new_stream(sink, batch.schema, options=pa.ipc.IpcWriteOptions(compression=compression)) as w:
            w.write_batch(batch)
        reader = pa.ipc.MessageReader.open_stream(io.BytesIO(sink.getvalue()))
        reader.read_next_message()
        return reader.read_next_message().serialize().to_pybytes()
    compressed, cbufs = record_batch_message_buffers(message("lz4"))
    plain, rbufs = record_batch_message_buffers(message(None))
    assert compressed and not plain and len(cbufs) == len(rbufs)
    frames = 0
    for cb, rb in zip(cbufs, rbufs):
        if not cb:
            continue
        raw_len = int.from_bytes(cb[:8], "little", signed=True)
        if raw_len == -1:
            assert cb[8:8 + len(rb)] == rb[:len(cb) - 8]
            continue
        got = F.frame_decompress(cb[8:])
        assert len(got) == raw_len and got == rb[:raw_len]
        frames += 1
    assert frames >= 5


def test_stream_with_lz4_flex_blocks_is_valid_arrow_ipc():
    """The stand-in shuffle file with its blocks compressed by lz4_flex itself (what the GPU reader test feeds the device
    decoder) is a stream Arrow's own reader accepts and reads back exactly."""
    rng = np.random.default_rng(21)
    batches = [random_batch(rng, n, with_dict=False) for n in (1, 3000, 40_000)]
    stream = reference_style_lz4_stream(batches, block_compress=F.block_compress)
    got = pa.ipc.open_stream(io.BytesIO(stream)).read_all().to_batches()
    assert len(got) == 3 and all(g.equals(b) for g, b in zip(got, batches))
