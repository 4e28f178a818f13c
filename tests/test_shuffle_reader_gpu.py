"""Consumer side of the shuffle on the device: libbgpu's IPC stream reader (LZ4_FRAME blocks
decoded by a kernel), bgpu_batch_concat, and the GpuShuffleReaderExec / CoalesceBatchesExec / CoalesceTasksExec mirrors.
Parity is bit-exact against Arrow's own reader / concat on the same bytes, and - for the two-stage job - against the oracle.
The reference's own `test_read_local_shuffle`
(/root/reference/ballista/core/src/execution_plans/shuffle_reader.rs:1006-1048) is restated."""
import io
import os
import sys

import numpy as np
import pyarrow as pa
import pytest

from helpers import expr_to_tree, lineitem_batch, random_batch

pytestmark = pytest.mark.gpu


def _stream(schema, batches, compression):
    sink = io.BytesIO()
    with pa.ipc.new_stream(sink, schema, options=pa.ipc.IpcWriteOptions(compression=compression)) as w:
        for b in batches:
            w.write_batch(b)
    return sink.getvalue()


def _same(got, want):
    """Bit-exact equality of two RecordBatches, null counts included."""
    assert got.schema.equals(want.schema), (got.schema, want.schema)
    assert got.num_rows == want.num_rows
    for g, w, name in zip(got.columns, want.columns, want.schema.names):
        assert g.null_count == w.null_count, name
        assert g.equals(w), name


@pytest.mark.parametrize("compression", [None, "lz4"])
@pytest.mark.parametrize("sizes", [[1], [7, 0, 1000], [8192, 8192, 3], [300_000]])
def test_stream_reader_matches_arrows_reader(bb, ctx, compression, sizes):
    from ballista_b200 import ipc

    rng = np.random.default_rng(len(sizes) * 1000 + sum(sizes))
    batches = [random_batch(rng, n, null_frac=0.1 if i % 2 == 0 else 0.0, with_dict=False) for i, n in enumerate(sizes)]
    schema = batches[0].schema
    data = _stream(schema, batches, compression)
    want = list(pa.ipc.open_stream(io.BytesIO(data)))
    r = ipc.StreamReader(ctx, data)
    assert r.schema.equals(schema)
    got = [b.download() for b in r]
    assert len(got) == len(want) and r.num_rows == sum(sizes)
    for g, w in zip(got, want):
        _same(g, w)
    assert r.maybe_next() is None  # finished stays finished


def test_stream_reader_low_entropy_columns_many_blocks(bb, ctx):
    """2 M rows of highly compressible columns: every buffer is a multi-block LZ4 frame with long overlapping matches."""
    from ballista_b200 import ipc

    n = 2_000_000
    rng = np.random.default_rng(3)
    batch = pa.RecordBatch.from_arrays(
        [pa.array(np.repeat(np.arange(n // 1000, dtype=np.int64), 1000)), pa.array(np.zeros(n, dtype=np.int64)),
         pa.array(rng.integers(0, 3, n).astype(np.int32)), pa.array(np.tile(np.array(["A", "N", "R", "F"]), n // 4))],
        names=["run", "zero", "small", "flag"])
    data = _stream(batch.schema, [batch], "lz4")
    assert len(data) < batch.nbytes // 4
    got = [b.download() for b in ipc.StreamReader(ctx, data)]
    assert len(got) == 1
    _same(got[0], batch)


def test_stream_reader_reads_our_own_writer(bb, ctx):
    from ballista_b200 import ipc

    rng = np.random.default_rng(9)
    batch = random_batch(rng, 5000, with_dict=False)
    dev = bb.DeviceBatch.upload(ctx, batch)
    part, offs = bb.BatchPartitioner.try_new(bb.Partitioning.Hash([0], 4), ctx).partition_device(dev)
    enc = ipc.encode_partitions(ctx, part, offs)
    whole = part.download()
    for p in range(4):
        data = ipc.schema_message(batch.schema, ctx) + bytes(enc.message(p)) + ipc.EOS
        got = [b.download() for b in ipc.StreamReader(ctx, data)]
        assert len(got) == 1
        _same(got[0], pa.RecordBatch.from_pylist(whole.slice(offs[p], offs[p + 1] - offs[p]).to_pylist(), schema=batch.schema))


def test_stream_reader_rejects_a_corrupt_lz4_block(bb, ctx):
    from ballista_b200 import ipc

    batch = pa.RecordBatch.from_arrays([pa.array(np.arange(100_000, dtype=np.int64) // 7)], names=["a"])
    data = bytearray(_stream(batch.schema, [batch], "lz4"))
    # damage the middle of the compressed body (not the framing)
    for i in range(len(data) // 2, len(data) // 2 + 64):
        data[i] ^= 0xFF
    r = ipc.StreamReader(ctx, bytes(data))
    try:
        got = r.maybe_next().download()
    except bb.BgpuError as e:
        assert "LZ4" in str(e)
    else:  # the damage may decode to well-formed but different bytes: then it must at least differ
        assert not got.equals(batch)


@pytest.mark.parametrize("damage", ["decreasing", "negative", "beyond_data"])
def test_stream_reader_rejects_corrupt_utf8_offsets(bb, ctx, damage):
    """A shuffle file from another executor whose utf8 offsets are not what Arrow promises (arrow-rs's reader validates
    them): every kernel downstream would index the data buffer with them, so the reader refuses the batch."""
    from ballista_b200 import ipc

    strings = [f"s{i:05d}" for i in range(3000)]
    batch = pa.RecordBatch.from_arrays([pa.array(np.arange(3000, dtype=np.int64)), pa.array(strings)], names=["a", "s"])
    data = bytearray(_stream(batch.schema, [batch], None))
    offs = np.arange(3001, dtype=np.int32) * 6  # every string is 6 bytes: find the offsets buffer in the body
    at = bytes(data).find(offs.tobytes())
    assert at > 0
    word = {"decreasing": (at + 4 * 1500, 10), "negative": (at, -8), "beyond_data": (at + 4 * 3000, 6 * 3000 + 4096)}[damage]
    data[word[0]: word[0] + 4] = np.int32(word[1]).tobytes()
    r = ipc.StreamReader(ctx, bytes(data))
    with pytest.raises(bb.BgpuError) as e:
        r.maybe_next()
    assert e.value.code == 1 and "offsets" in str(e.value)  # BGPU_EINVAL
    # the undamaged stream still reads
    ok = ipc.StreamReader(ctx, _stream(batch.schema, [batch], None)).maybe_next().download()
    assert ok.equals(batch)


@pytest.mark.parametrize("sizes", [[5], [3, 5], [8, 0, 9, 1], [1000, 1, 63, 64, 65], [40_000, 70_001, 13]])
def test_concat_matches_arrow(bb, ctx, sizes):
    rng = np.random.default_rng(sum(sizes))
    # every other input has no nulls at all (no validity buffer): the result still needs one
    batches = [random_batch(rng, n, null_frac=0.2 if i % 2 == 0 else 0.0, with_dict=False) for i, n in enumerate(sizes)]
    want = pa.Table.from_batches(batches).combine_chunks().to_batches()[0]
    got = bb.DeviceBatch.concat(ctx, [bb.DeviceBatch.upload(ctx, b) for b in batches]).download()
    _same(got, want)


def test_concat_of_partition_slices_round_trips(bb, ctx):
    """Concatenating the partitions of a stable hash repartition = the partitioned batch itself (sliced inputs:
    utf8 offsets that do not start at 0 come from the IPC reader)."""
    from ballista_b200 import ipc

    rng = np.random.default_rng(21)
    batch = random_batch(rng, 20_000, with_dict=False)
    dev = bb.DeviceBatch.upload(ctx, batch)
    part, offs = bb.BatchPartitioner.try_new(bb.Partitioning.Hash([0, 9], 8), ctx).partition_device(dev)
    enc = ipc.encode_partitions(ctx, part, offs)
    pieces = []
    for p in range(8):
        data = ipc.schema_message(batch.schema, ctx) + bytes(enc.message(p)) + ipc.EOS
        pieces.extend(ipc.StreamReader(ctx, data))
    got = bb.DeviceBatch.concat(ctx, pieces).download()
    _same(got, pa.RecordBatch.from_pylist(part.download().to_pylist(), schema=batch.schema))


def test_concat_rejects_mismatched_schemas(bb, ctx):
    a = bb.DeviceBatch.upload(ctx, pa.RecordBatch.from_arrays([pa.array([1, 2], type=pa.int64())], names=["a"]))
    b = bb.DeviceBatch.upload(ctx, pa.RecordBatch.from_arrays([pa.array([1, 2], type=pa.int32())], names=["a"]))
    with pytest.raises(bb.BgpuError):
        bb.DeviceBatch.concat(ctx, [a, b])


def test_coalesce_batches_semantics(bb, ctx):
    rng = np.random.default_rng(2)
    sizes = [10, 20, 0, 100, 5, 300, 40, 7]
    host = [random_batch(rng, n, with_dict=False, with_strings=False) for n in sizes]
    out = list(bb.CoalesceBatchesExec([bb.DeviceBatch.upload(ctx, b) for b in host], ctx, target_batch_size=100))
    # 10+20+100 -> 130; 5+300 -> 305 (the buffer is not empty, so the big batch is buffered too); 40+7 flushed at the end
    assert [b.num_rows for b in out] == [130, 305, 47]
    want = pa.Table.from_batches(host).combine_chunks().to_batches()[0]
    _same(pa.Table.from_batches([b.download() for b in out]).combine_chunks().to_batches()[0], want)
    big = bb.DeviceBatch.upload(ctx, host[5])
    assert list(bb.CoalesceBatchesExec([big], ctx, target_batch_size=100))[0] is big  # passes through untouched


def test_coalesce_tasks_fan_in(bb):
    parts = {0: ["a0", "a1", "a2"], 1: ["b0"], 2: [], 3: ["d0", "d1"]}
    one = bb.CoalesceTasksExec(lambda p: parts[p], [0])
    assert list(one.execute()) == parts[0]
    many = bb.CoalesceTasksExec(lambda p: parts[p], [0, 1, 2, 3])
    got = list(many.execute())
    assert sorted(got) == sorted(sum(parts.values(), [])) and got[:3] == ["a0", "b0", "d0"]
    assert many.output_partitioning().partition_count() == 1


@pytest.mark.parametrize("ipc_mode", ["device", "host"])
def test_read_local_shuffle(bb, ctx, tmp_path, ipc_mode):
    """shuffle_reader.rs:1006-1048: write one partition with two batches, read the file back: 2 batches, each equal."""
    schema = pa.schema([pa.field("a", pa.uint32(), True), pa.field("b", pa.string(), True)])
    batch = pa.RecordBatch.from_arrays([pa.array([1, 2], type=pa.uint32()), pa.array(["hello", "world"])], schema=schema)
    writer = bb.GpuShuffleWriterExec.try_new("local_file", 1, [0], [batch, batch], str(tmp_path), bb.Partitioning.Hash([0], 1), ctx,
                                             input_schema=schema, ipc=ipc_mode)
    meta = writer.execute(0)[0]
    path = meta.column(1)[0].as_py()
    loc = bb.PartitionLocation([0], "local_file", 1, 0, path)
    reader = bb.GpuShuffleReaderExec([[loc]], schema, ctx)
    result = [b.download() for b in reader.execute(0)]
    assert len(result) == 2
    for b in result:
        assert b.equals(batch)
    assert reader.metrics["num_rows"] == 4 and reader.metrics["bytes_read"] == os.path.getsize(path)


def test_fetch_failed_mapping(bb, ctx, tmp_path):
    """A missing or unreadable partition file surfaces as FetchFailed carrying the map stage and partitions
    (shuffle_reader.rs:751-759, test_fetch_partitions_error_mapping :942-993)."""
    meta = bb.ExecutorMetadata(id="exec-7")
    loc = bb.PartitionLocation([3, 4], "job", 2, 0, str(tmp_path / "missing.arrow"), meta)
    reader = bb.GpuShuffleReaderExec([[loc]], pa.schema([("a", pa.int64())]), ctx)
    with pytest.raises(bb.FetchFailed) as e:
        list(reader.execute(0))
    assert e.value.executor_id == "exec-7" and e.value.map_stage_id == 2 and e.value.map_partitions == [3, 4]
    assert "Failed to open partition file" in e.value.message
    bad = tmp_path / "bad.arrow"
    bad.write_bytes(b"not an arrow stream at all")
    reader = bb.GpuShuffleReaderExec([[bb.PartitionLocation([0], "job", 2, 0, str(bad), meta)]], pa.schema([("a", pa.int64())]), ctx)
    with pytest.raises(bb.FetchFailed) as e:
        list(reader.execute(0))
    assert "Failed to new arrow StreamReader" in e.value.message
    # locations whose statistics say "no rows" are never fetched
    empty = bb.PartitionLocation([0], "job", 2, 0, str(tmp_path / "missing.arrow"), meta, bb.PartitionStats(0, 0, 0))
    assert list(bb.GpuShuffleReaderExec([[empty]], pa.schema([("a", pa.int64())]), ctx).execute(0)) == []


@pytest.mark.parametrize("ipc_mode", ["device", "host"])
def test_two_stage_q1_job_through_shuffle_files(bb, ctx, oracle, tmp_path, ipc_mode):
    """The whole path as the scheduler plans it (planner.rs:376-427), data on the device between the operators:
    stage 1 = CoalesceTasksExec(3 input partitions) -> Partial aggregate -> GpuShuffleWriterExec(Hash(keys), 4);
    stage 2, per reduce partition = GpuShuffleReaderExec -> CoalesceBatchesExec -> FinalPartitioned aggregate.
    Two map tasks write the files; the union of the reduce outputs must equal the oracle's Partial -> Final."""
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import bench

    rng = np.random.default_rng(17)
    schema = bench.lineitem_schema()
    gi, aggs, filt = bench.q1_plan()
    n_out = 4
    inputs = {t: {p: [lineitem_batch(rng, int(n)) for n in rng.integers(1, 4000, 3)] for p in range(3)} for t in range(2)}
    locations = [[] for _ in range(n_out)]
    state_schema = None
    for task in range(2):
        fan_in = bb.CoalesceTasksExec(lambda p, t=task: (bb.DeviceBatch.upload(ctx, b) for b in inputs[t][p]), [0, 1, 2], schema)
        agg = bb.GpuAggregateExec(ctx, bb.AggregateMode.Partial, gi, aggs, schema, filter=filt)
        for b in fan_in.execute():
            agg.update(b)
        state = agg.emit()
        state_schema = state.download().schema
        writer = bb.GpuShuffleWriterExec.try_new("q1", 1, [3 * task, 3 * task + 1, 3 * task + 2],
                                                 [state if ipc_mode == "device" else state.download()], str(tmp_path),
                                                 bb.Partitioning.Hash(list(range(len(gi))), n_out), ctx,
                                                 input_schema=state_schema, ipc=ipc_mode)
        for w in writer.execute_shuffle_write():
            locations[w.output_partition].append(bb.PartitionLocation.from_shuffle_write("q1", 1, w))
    reader = bb.GpuShuffleReaderExec(locations, state_schema, ctx)
    finals = []
    for p in range(n_out):
        fin = bb.GpuAggregateExec(ctx, bb.AggregateMode.FinalPartitioned, gi, aggs, schema)
        seen = 0
        for b in bb.CoalesceBatchesExec(reader.execute(p), ctx):
            fin.update(b)
            seen += b.num_rows
        if seen:
            finals.append(fin.emit().download())
    got = pa.Table.from_batches(finals)

    o_aggs = [(f, expr_to_tree(a), nm) for f, a, nm in aggs]
    states = []
    for task in range(2):
        op = oracle.GroupedAggregate("partial", gi, o_aggs, filter=expr_to_tree(filt))
        for p in range(3):
            for b in inputs[task][p]:
                op.update(b)
        states.append(op.emit())
    arg_types = [None if f == "count" else states[0].schema.field(f"{nm}[sum]").type for f, _, nm in aggs]
    arg_types = [None if t is None else pa.decimal128(t.precision - 10, t.scale) for t in arg_types]
    of = oracle.GroupedAggregate("final", list(range(len(gi))), [(f, 0, nm) for f, _, nm in aggs], arg_types=arg_types)
    for s in states:
        of.update(s)
    want = pa.Table.from_batches([of.emit()])
    keys = [(n_, "ascending") for n_ in got.schema.names[: len(gi)]]
    assert got.sort_by(keys).equals(want.sort_by(keys))


@pytest.mark.parametrize("sizes", [[1], [7, 0, 1000], [200_000, 33_333]])
def test_stream_reader_reads_reference_style_lz4_files(bb, ctx, sizes):
    """Streams framed as the reference's writer frames them - block-independent 64 KB LZ4 blocks (lz4_flex) behind the int64
    length prefix, stored buffers with -1: every block is one decode chain, decoded in shared memory by its own warp."""
    from helpers import reference_style_lz4_stream
    from ballista_b200 import ipc

    rng = np.random.default_rng(77 + sum(sizes))
    batches = [random_batch(rng, n, null_frac=0.1 if i % 2 == 0 else 0.0, with_dict=False) for i, n in enumerate(sizes)]
    data = reference_style_lz4_stream(batches)
    got = [b.download() for b in ipc.StreamReader(ctx, data)]
    assert len(got) == len(batches)
    for g, w in zip(got, batches):
        _same(g, w)


def test_edge_cases_empty_inputs(bb, ctx):
    """Empty streams, empty batches and empty partitions through the reader, the concat and the LZ4 encoder."""
    from ballista_b200 import ipc

    rng = np.random.default_rng(0)
    proto = random_batch(rng, 10, with_dict=False)
    empty = proto.slice(0, 0)
    # a stream with a schema and no batch: nothing to iterate, the schema is there
    r = ipc.StreamReader(ctx, _stream(proto.schema, [], None))
    assert r.schema.equals(proto.schema) and list(r) == []
    # zero-row batches (compressed and not) come back as zero-row batches
    for comp in (None, "lz4"):
        got = [b.download() for b in ipc.StreamReader(ctx, _stream(proto.schema, [empty, proto, empty], comp))]
        assert [g.num_rows for g in got] == [0, 10, 0]
        _same(got[1], proto)
    # concat of nothing but empty batches, and of empty + one
    e = bb.DeviceBatch.upload(ctx, empty)
    out = bb.DeviceBatch.concat(ctx, [e, e]).download()
    assert out.num_rows == 0 and out.schema.equals(proto.schema)
    _same(bb.DeviceBatch.concat(ctx, [e, bb.DeviceBatch.upload(ctx, proto), e]).download(), proto)
    # every partition empty: no message at all, with and without compression
    part, offs = bb.BatchPartitioner.try_new(bb.Partitioning.Hash([0], 4), ctx).partition_device(e)
    for comp in (None, "lz4"):
        enc = ipc.encode_partitions(ctx, part, offs, compression=comp)
        assert enc.total_bytes == 0 and all(len(enc.message(p)) == 0 for p in range(4))
    # one row: three of four partitions are empty
    one = bb.DeviceBatch.upload(ctx, proto.slice(0, 1))
    part, offs = bb.BatchPartitioner.try_new(bb.Partitioning.Hash([0], 4), ctx).partition_device(one)
    enc = ipc.encode_partitions(ctx, part, offs, compression="lz4")
    assert sum(len(enc.message(p)) > 0 for p in range(4)) == 1
