"""GPU parity of the streamed input (bgpu_agg_push_host: host batches of the reference's 8192 rows -> pinned wire chunks ->
widen on the device -> one launch per chunk without a host round trip) against the oracle, bit-exact for integers and
Decimal128, 1e-9 relative for Float64 (the tolerance of tests/test_aggregate_gpu.py)."""
import decimal
import threading

import numpy as np
import pyarrow as pa
import pytest

from helpers import expr_to_tree, lineitem_batch, q1_plan, random_batch
from test_aggregate_gpu import Q1_AGGS, assert_same_groups

pytestmark = pytest.mark.gpu


def _spec(schema, groups, aggs):
    gi = [schema.names.index(g) for g in groups]
    return gi, [(f, (None if c is None else schema.names.index(c)), nm) for f, c, nm in aggs]


def _oracle_partial(oracle, batches, gi, spec, **kw):
    op = oracle.GroupedAggregate("partial", gi, spec, **kw)
    for b in batches:
        op.update(b)
    return op.emit()


@pytest.mark.parametrize("groups", [["flag"], ["flag", "small"], ["k32"], ["date", "b"], ["i16", "u8"], ["dec"]])
def test_pushed_batches_match_oracle(bb, ctx, oracle, groups, monkeypatch):
    """Every column kind through the wire form (nulls, strings of mixed lengths, bools, narrow and wide integers,
    Decimal128, floats), small chunks so that several launches and both device buffer sets are used."""
    monkeypatch.setenv("BGPU_INGEST_CHUNK_ROWS", "20000")
    rng = np.random.default_rng(23)
    whole = random_batch(rng, 90_000, with_dict=False)
    batches = [whole.slice(i, 8192) for i in range(0, whole.num_rows, 8192)]
    gi, spec = _spec(whole.schema, groups, Q1_AGGS)
    agg = bb.GpuAggregateExec(ctx, bb.AggregateMode.Partial, gi, spec, whole.schema)
    for b in batches:
        agg.push(b)
    got = agg.emit().download()
    st = agg.ingest_stats
    assert st["rows"] == whole.num_rows and st["batches"] == len(batches) and st["chunks"] >= 4
    assert_same_groups(got, _oracle_partial(oracle, batches, gi, spec), len(gi))


def test_thousand_tiny_batches_with_growth_and_rung_change(bb, ctx, oracle, monkeypatch):
    """>= 1000 batches; the number of groups grows mid-stream past the private rung's capacity (launches that do not commit
    are replayed), past the shared rungs and into the two-pass path with a table growth."""
    monkeypatch.setenv("BGPU_INGEST_CHUNK_ROWS", "16384")
    rng = np.random.default_rng(5)
    n_batches, rows = 1200, 512
    n = n_batches * rows
    phase = np.arange(n) // (n // 6)
    card = np.array([3, 40, 900, 20_000, 150_000, 150_000])[np.minimum(phase, 5)]
    k = (rng.integers(0, 1 << 40, n) % card).astype(np.int64)
    whole = pa.RecordBatch.from_arrays(
        [pa.array(k), pa.array(rng.integers(-1000, 1000, n), type=pa.int64(), mask=rng.random(n) < 0.05),
         pa.array(rng.random(n)), pa.array([decimal.Decimal(int(x)).scaleb(-2) for x in rng.integers(0, 10**9, n)], type=pa.decimal128(15, 2))],
        names=["k", "v", "f", "d"])
    batches = [whole.slice(i, rows) for i in range(0, n, rows)]
    gi, spec = _spec(whole.schema, ["k"], [("sum", "v", "s"), ("count", "v", "c"), ("avg", "f", "a"), ("sum", "d", "sd"),
                                           ("avg", "d", "ad"), ("count", None, "n")])
    agg = bb.GpuAggregateExec(ctx, bb.AggregateMode.Partial, gi, spec, whole.schema)
    paths = set()
    for i, b in enumerate(batches):
        agg.push(b)
        if i % 100 == 99:
            paths.add(agg.last_path)
    got = agg.emit().download()
    st = agg.ingest_stats
    assert st["batches"] == n_batches and st["rows"] == n
    assert st["replays"] >= 1, st           # capacity / rung changes were met by launches nobody had looked at
    assert st["async_launches"] >= 5, st    # and most launches ran without a host round trip
    assert len(paths) >= 2, paths
    assert_same_groups(got, _oracle_partial(oracle, [whole], gi, spec), 1)


def test_q1_pushed_from_four_threads_matches_oracle(bb, ctx, oracle, monkeypatch):
    """TPC-H Q1's Partial aggregate (filter + decimal expressions fused) fed 8192-row batches from four threads at once:
    bit-exact against the oracle, and the wire is several times narrower than the Arrow buffers."""
    monkeypatch.setenv("BGPU_INGEST_CHUNK_ROWS", "100000")
    rng = np.random.default_rng(11)
    li = lineitem_batch(rng, 400_000)
    batches = [li.slice(i, 8192) for i in range(0, li.num_rows, 8192)]
    gi, aggs, filt = q1_plan()
    agg = bb.GpuAggregateExec(ctx, bb.AggregateMode.Partial, gi, aggs, li.schema, filter=filt)
    errors = []

    def feed(part):
        try:
            for b in part:
                agg.push(b)
        except Exception as e:  # noqa: BLE001
            errors.append(e)

    threads = [threading.Thread(target=feed, args=(batches[t::4],), daemon=True) for t in range(4)]
    for t in threads:
        t.start()
    for t in threads:
        t.join(timeout=120)
    assert not errors and not any(t.is_alive() for t in threads), errors
    got = agg.emit().download()
    st = agg.ingest_stats
    assert st["rows"] == li.num_rows
    assert st["wire_bytes"] * 4 < st["arrow_bytes"], st
    o_aggs = [(f, expr_to_tree(a), nm) for f, a, nm in aggs]
    want = _oracle_partial(oracle, [li], gi, o_aggs, filter=expr_to_tree(filt))
    assert_same_groups(got, want, len(gi))


def test_push_edge_cases(bb, ctx, oracle):
    """Empty batches, a batch larger than a segment, sliced batches (array offsets), a constant column, all-null values."""
    rng = np.random.default_rng(2)
    n = 150_000
    big = pa.RecordBatch.from_arrays(
        [pa.array(rng.integers(0, 5, n), type=pa.int32()), pa.array(np.full(n, 7), type=pa.int64()),
         pa.array([None] * n, type=pa.int64()), pa.array(rng.integers(-2**62, 2**62, n), type=pa.int64()),
         pa.array(rng.integers(0, 2, n).astype(bool)), pa.array(["x" * int(v) for v in rng.integers(0, 4, n)])],
        names=["k", "c", "nul", "w", "flag", "s"])
    batches = [big.slice(0, 0), big.slice(3, 70_001), big.slice(70_004, 5), big.slice(70_009, 0), big.slice(70_009)]
    for groups in (["k"], ["flag", "s"]):
        gi, spec = _spec(big.schema, groups, [("sum", "c", "sc"), ("sum", "nul", "sn"), ("count", "nul", "cn"), ("sum", "w", "sw"),
                                              ("avg", "c", "ac"), ("count", "w", "cw"), ("count", None, "n")])
        agg = bb.GpuAggregateExec(ctx, bb.AggregateMode.Partial, gi, spec, big.schema)
        for b in batches:
            agg.push(b)
        agg.flush()
        agg.push(big.slice(1, 2))  # pushing goes on after a flush
        got = agg.emit().download()
        assert_same_groups(got, _oracle_partial(oracle, batches + [big.slice(1, 2)], gi, spec), len(gi))


def test_push_and_update_agree(bb, ctx):
    """The streamed input and bgpu_batch_upload + bgpu_agg_update give the same state, bit for bit (integers)."""
    rng = np.random.default_rng(8)
    li = lineitem_batch(rng, 100_000)
    gi, aggs, filt = q1_plan()
    a1 = bb.GpuAggregateExec(ctx, bb.AggregateMode.Partial, gi, aggs, li.schema, filter=filt)
    a2 = bb.GpuAggregateExec(ctx, bb.AggregateMode.Partial, gi, aggs, li.schema, filter=filt)
    for i in range(0, li.num_rows, 8192):
        a1.push(li.slice(i, 8192))
        a2.update(li.slice(i, 8192))
    t1 = pa.Table.from_batches([a1.emit().download()])
    t2 = pa.Table.from_batches([a2.emit().download()])
    keys = [(nm, "ascending") for nm in t1.schema.names[: len(gi)]]
    assert t1.sort_by(keys).equals(t2.sort_by(keys))
