"""Dictionary<Int32, Utf8> columns through everything that moves rows of several batches together (SURVEY.md §8 a5, a11, a12;
db-benchmark's id1..id3 are dictionary-encoded, /root/reference/benchmarks/db-benchmark/groupby-datafusion.py:56): concat,
the exchange, shuffle files written on the device (DictionaryBatch + RecordBatch messages, read back by Arrow's reader) and
files Arrow wrote (read by the device reader). Rows are compared by VALUE: dictionaries are re-built on the way."""
import io

import numpy as np
import pyarrow as pa
import pytest

from helpers import decode_dictionaries, random_batch

pytestmark = pytest.mark.gpu


def _dict_batch(rng, n, card=50, null_frac=0.1):
    vals = pa.array([f"id{int(i):03d}" for i in rng.integers(1, card + 1, n)], type=pa.string(),
                    mask=rng.random(n) < null_frac if null_frac else None).dictionary_encode()
    wide = pa.array([("w%06d" % i) * 3 for i in rng.integers(0, 7, n)], type=pa.string()).dictionary_encode()
    return pa.RecordBatch.from_arrays([pa.array(rng.integers(0, 9, n), type=pa.int64()), vals,
                                       pa.array(rng.random(n)), wide], names=["k", "id", "f", "wide"])


def _same_rows(got, want):
    assert got.schema.equals(want.schema), (got.schema, want.schema)
    assert decode_dictionaries(got).equals(decode_dictionaries(want))


@pytest.mark.parametrize("sizes", [[5], [300, 0, 7, 4000], [70_000, 3]])
def test_concat_of_batches_with_their_own_dictionaries(bb, ctx, sizes):
    rng = np.random.default_rng(len(sizes))
    batches = [_dict_batch(rng, n, card=10 + 7 * i) for i, n in enumerate(sizes)]
    got = bb.DeviceBatch.concat(ctx, [bb.DeviceBatch.upload(ctx, b) for b in batches]).download()
    _same_rows(got, pa.Table.from_batches(batches))


def test_exchange_and_small_exchange_carry_dictionary_columns(bb, ctx, oracle):
    rng = np.random.default_rng(4)
    ex = bb.Exchange(ctx, 0, 1)
    for rows, n in ((20_000, 6), (100, 16)):
        b = _dict_batch(rng, rows)
        part, offs = bb.BatchPartitioner.try_new(bb.Partitioning.Hash([1, 0], n), ctx).partition_device(bb.DeviceBatch.upload(ctx, b))
        got = ex.all_to_all(part, offs, n).download()
        want = pa.Table.from_batches([pb for _, pb in oracle.partition(b, [1, 0], n)[0]])
        _same_rows(got, want)
        if rows <= 256:
            small = ex.small_exchange(bb.DeviceBatch.upload(ctx, b), [1, 0], n)
            assert small is not None
            _same_rows(pa.Table.from_batches([small.download()]).sort_by([("f", "ascending")]), want.sort_by([("f", "ascending")]))
    ex.close()


@pytest.mark.parametrize("compression", [None, "lz4"])
def test_device_written_dictionary_messages_are_read_by_arrow_and_by_us(bb, ctx, oracle, compression):
    from ballista_b200 import ipc

    rng = np.random.default_rng(9)
    b = _dict_batch(rng, 30_000, null_frac=0.2)
    n = 5
    part, offs = bb.BatchPartitioner.try_new(bb.Partitioning.Hash([1], n), ctx).partition_device(bb.DeviceBatch.upload(ctx, b))
    enc = ipc.encode_partitions(ctx, part, offs, compression)
    schema_msg = ipc.schema_message(b.schema)
    want = dict(oracle.partition(b, [1], n)[0])
    for p in range(n):
        msg = bytes(enc.message(p))
        if p not in want:
            assert msg == b""
            continue
        stream = schema_msg + msg + msg + ipc.EOS  # twice: the second dictionary batch replaces the first
        got = pa.ipc.open_stream(io.BytesIO(stream)).read_all()
        _same_rows(got, pa.Table.from_batches([want[p], want[p]]))
        ours = [x.download() for x in ipc.StreamReader(ctx, stream)]
        assert len(ours) == 2
        _same_rows(pa.Table.from_batches(ours), pa.Table.from_batches([want[p], want[p]]))
    enc.close()


@pytest.mark.parametrize("compression", [None, "lz4"])
def test_reader_takes_dictionary_streams_written_by_arrow(bb, ctx, compression):
    """What the reference's writer puts into a shuffle file for a Partial state with dictionary keys: a dictionary batch
    per dictionary, replaced when a later batch brings another dictionary."""
    from ballista_b200 import ipc

    rng = np.random.default_rng(12)
    batches = [_dict_batch(rng, n, card=c) for n, c in ((1000, 5), (1, 3), (5000, 40))]
    sink = io.BytesIO()
    with pa.ipc.new_stream(sink, batches[0].schema, options=pa.ipc.IpcWriteOptions(compression=compression)) as w:
        for b in batches:
            w.write_batch(b)
    got = [x.download() for x in ipc.StreamReader(ctx, sink.getvalue())]
    assert [g.num_rows for g in got] == [b.num_rows for b in batches]
    for g, b in zip(got, batches):
        _same_rows(g, b)


def test_reader_rejects_an_index_beyond_its_dictionary(bb, ctx):
    from ballista_b200 import ipc

    idx = pa.array(np.arange(2000, dtype=np.int32) % 7)
    arr = pa.DictionaryArray.from_arrays(idx, pa.array([f"v{i}" for i in range(7)]))
    b = pa.RecordBatch.from_arrays([arr], names=["d"])
    sink = io.BytesIO()
    with pa.ipc.new_stream(sink, b.schema) as w:
        w.write_batch(b)
    data = bytearray(sink.getvalue())
    at = bytes(data).find(idx.buffers()[1].to_pybytes()[:400])
    assert at > 0
    data[at + 40: at + 44] = np.int32(7).tobytes()  # one index past the end of the dictionary
    with pytest.raises(bb.BgpuError, match="dictionary index"):
        ipc.StreamReader(ctx, bytes(data)).maybe_next()


def test_partial_state_with_dictionary_key_through_a_shuffle_file_and_final(bb, ctx, oracle, tmp_path):
    """db-benchmark q1's shape end to end: Partial (dictionary key) -> device-framed partition files -> device reader ->
    FinalPartitioned, against the oracle's Partial -> Final."""
    from ballista_b200.shuffle_writer import GpuShuffleWriterExec

    rng = np.random.default_rng(21)
    b = random_batch(rng, 40_000)
    gi = [b.schema.names.index("d")]
    spec = [("sum", b.schema.names.index("small"), "s"), ("avg", b.schema.names.index("f"), "a"), ("count", None, "n")]
    op = oracle.GroupedAggregate("partial", gi, spec)
    op.update(b)
    of = oracle.GroupedAggregate("final", [0], spec, arg_types=[pa.int64(), pa.float64(), None])
    of.update(op.emit())
    want = of.emit()
    gp = bb.GpuAggregateExec(ctx, bb.AggregateMode.Partial, gi, spec, b.schema)
    gp.update(b)
    state = gp.emit()
    w = GpuShuffleWriterExec.try_new("job", 1, [0], [state], str(tmp_path), bb.Partitioning.Hash([0], 4), ctx)
    locs = w.execute_shuffle_write()
    gf = bb.GpuAggregateExec(ctx, bb.AggregateMode.FinalPartitioned, gi, spec, b.schema)
    rows = 0
    for l in locs:
        arrow_rows = pa.ipc.open_stream(l.path).read_all()   # Arrow reads the device-framed file
        rows += arrow_rows.num_rows
        for batch in ipc_batches(bb, ctx, l.path):
            gf.update(batch)
    assert rows == state.num_rows
    got = gf.emit().download()
    from test_aggregate_gpu import assert_same_groups
    assert_same_groups(got, want, 1)


def ipc_batches(bb, ctx, path):
    from ballista_b200 import ipc

    with open(path, "rb") as f:
        data = f.read()
    return list(ipc.StreamReader(ctx, data))


@pytest.mark.parametrize("distinct", [3_000, 180_000])
def test_many_dictionaries_with_many_values_fold_into_one(bb, ctx, oracle, distinct):
    """Every batch brings its OWN dictionary (duplicates inside, values shared between batches - the shape of states that
    arrive from several ranks): the canonical dictionary the aggregate emits holds every value once, whatever growth of
    the value table and its byte arena (1 MiB at first) the batches forced; groups and sums are the oracle's."""
    rng = np.random.default_rng(distinct)
    batches = []
    for j in range(3):
        n = 120_000
        ids = rng.integers(0, distinct, n)
        dict_vals = pa.array([f"id{int(i):010d}" for i in ids], type=pa.string())   # an identity dictionary with duplicates
        key = pa.DictionaryArray.from_arrays(pa.array(np.arange(n, dtype=np.int32)), dict_vals)
        batches.append(pa.RecordBatch.from_arrays([key, pa.array(rng.integers(-9, 9, n), type=pa.int64())], names=["id3", "v"]))
    spec = [("sum", 1, "s"), ("count", None, "n")]
    want = oracle.GroupedAggregate("partial", [0], spec)
    agg = bb.GpuAggregateExec(ctx, bb.AggregateMode.Partial, [0], spec, batches[0].schema)
    for round_ in range(2):      # a second stream through the same operator (reset) starts from an empty dictionary again
        for b in batches:
            agg.update(b)
            if round_ == 0:
                want.update(b)
        got = agg.emit().download()
        if round_ == 0:
            expect = want.emit()
        from test_aggregate_gpu import assert_same_groups
        assert_same_groups(got, expect, 1)
        d = got.column(0).dictionary
        assert len(d) == len(set(d.to_pylist())) == got.num_rows   # every value once, none invented
        agg.reset()
