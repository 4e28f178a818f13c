"""The reference's reader can read what the device writes: every LZ4 frame that lz4_compress_kernel + the host layout of
bgpu_ipc_encode_partitions(compression = LZ4_FRAME) put into a RecordBatch message is decoded by the EXECUTED lz4_flex
FrameDecoder (tests/lz4_flex_ffi.py - the crate arrow-ipc's reader calls in the reference build,
/root/reference/ballista/core/src/execution_plans/shuffle_reader.rs:771-785) to exactly the bytes of the same buffer in the
uncompressed message. Skipped where that library is missing."""
import numpy as np
import pyarrow as pa
import pytest

import lz4_flex_ffi as F
from helpers import random_batch, record_batch_message_buffers, reference_style_lz4_stream

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not F.available(), reason="lz4_flex (wandb/bin/librust_parquet_ffi.so) not in this image")]


@pytest.mark.parametrize("n_rows", [1, 1000, 300_000])
def test_device_lz4_frames_decode_with_lz4_flex(bb, ctx, n_rows):
    from ballista_b200 import ipc

    rng = np.random.default_rng(n_rows + 1)
    batch = random_batch(rng, n_rows, with_dict=False)
    low = pa.RecordBatch.from_arrays([pa.array(np.arange(n_rows, dtype=np.int64) // 50), pa.array(np.zeros(n_rows, dtype=np.int64)),
                                      pa.array(rng.integers(0, 3, n_rows).astype(np.int32))], names=["run", "zero", "small"])
    frames = stored = 0
    for b in (batch, low):
        dev = bb.DeviceBatch.upload(ctx, b)
        part, offs = bb.BatchPartitioner.try_new(bb.Partitioning.Hash([0], 4), ctx).partition_device(dev)
        enc = ipc.encode_partitions(ctx, part, offs, compression="lz4")
        raw = ipc.encode_partitions(ctx, part, offs)
        for p in range(4):
            if offs[p + 1] == offs[p]:
                continue
            compressed, cbufs = record_batch_message_buffers(enc.message(p))
            plain, rbufs = record_batch_message_buffers(raw.message(p))
            assert compressed and not plain and len(cbufs) == len(rbufs)
            for cb, rb in zip(cbufs, rbufs):
                if not cb:
                    assert not rb
                    continue
                raw_len = int.from_bytes(cb[:8], "little", signed=True)
                if raw_len == -1:          # stored: compression did not pay
                    assert cb[8:] == rb
                    stored += 1
                    continue
                assert raw_len == len(rb)
                assert F.frame_decompress(cb[8:]) == rb
                frames += 1
        enc.close()
        raw.close()
    assert frames > 0 or n_rows == 1   # buffers of a one-row batch are too small for a frame to pay: all stored


@pytest.mark.parametrize("sizes", [[1], [7, 0, 1000], [200_000, 33_333]])
def test_device_reader_reads_files_with_lz4_flex_blocks(bb, ctx, sizes):
    """The other direction: a shuffle file framed as the reference's writer frames it, its blocks compressed by lz4_flex
    itself, read by libbgpu's stream reader (frames walked on the host, blocks decoded by the kernel), bit-exact."""
    from ballista_b200 import ipc

    rng = np.random.default_rng(sum(sizes))
    batches = [random_batch(rng, n, with_dict=False) for n in sizes]
    stream = reference_style_lz4_stream(batches, block_compress=F.block_compress)
    got = [b.download() for b in ipc.StreamReader(ctx, stream)]
    assert [g.num_rows for g in got] == sizes
    for g, w in zip(got, batches):
        assert g.schema.equals(w.schema)
        for a, b, name in zip(g.columns, w.columns, w.schema.names):
            assert a.null_count == b.null_count and a.equals(b), name
