"""GPU parity: GpuAggregateExec (Partial, FinalPartitioned) against the oracle's GroupedHashAggregateStream
restatement. Integer / Decimal128 sums, counts and decimal averages are bit-exact; Float64 within 1e-9 relative
(the north star's tolerance — the reference's own float sum order is not deterministic)."""
import decimal

import numpy as np
import pyarrow as pa
import pyarrow.compute as pc
import pytest

from helpers import random_batch

pytestmark = pytest.mark.gpu
RTOL = 1e-9


def _sorted(rb, n_keys):
    t = pa.Table.from_batches([rb])
    cols = []
    for i in range(n_keys):
        c = t.column(i)
        if pa.types.is_dictionary(c.type):
            c = c.cast(c.type.value_type)
        cols.append(c)
    key_tbl = pa.table({f"k{i}": c for i, c in enumerate(cols)})
    idx = pc.sort_indices(key_tbl, sort_keys=[(f"k{i}", "ascending") for i in range(n_keys)], null_placement="at_end")
    return t.take(idx), key_tbl.take(idx)


def assert_same_groups(got, want, n_keys):
    assert got.schema.names == want.schema.names
    assert got.num_rows == want.num_rows
    g, gk = _sorted(got, n_keys)
    w, wk = _sorted(want, n_keys)
    assert gk.equals(wk), "group keys differ"
    for i in range(n_keys, g.num_columns):
        a, b = g.column(i).combine_chunks(), w.column(i).combine_chunks()
        name = g.schema.names[i]
        assert a.type == b.type, (name, a.type, b.type)
        if pa.types.is_floating(a.type):
            assert a.is_null().equals(b.is_null()), name
            x, y = a.fill_null(0).to_numpy(), b.fill_null(0).to_numpy()
            assert np.allclose(x, y, rtol=RTOL, atol=0), name
        else:
            assert a.equals(b), f"column {name} differs"


def run_both(bb, ctx, oracle, batches, group_names, aggs, n_partials=2):
    schema = batches[0].schema
    gi = [schema.names.index(g) for g in group_names]
    spec = [(f, (None if c is None else schema.names.index(c)), nm) for f, c, nm in aggs]
    # oracle: partial per input partition, then final
    OP = [oracle.GroupedAggregate("partial", gi, spec) for _ in range(n_partials)]
    GP = [bb.GpuAggregateExec(ctx, bb.AggregateMode.Partial, gi, spec, schema) for _ in range(n_partials)]
    for i, b in enumerate(batches):
        OP[i % n_partials].update(b)
        GP[i % n_partials].update(b)
    ostates = [p.emit() for p in OP]
    gstates = [p.emit().download() for p in GP]
    for o, g_ in zip(ostates, gstates):
        assert_same_groups(g_, o, len(gi))
    arg_types = [schema.field(c).type if c is not None else None for _, c, _ in spec]
    OF = oracle.GroupedAggregate("final", list(range(len(gi))), spec, arg_types=arg_types)
    GF = bb.GpuAggregateExec(ctx, bb.AggregateMode.FinalPartitioned, gi, spec, schema)
    for o, g_ in zip(ostates, gstates):
        OF.update(o)
        GF.update(g_)
    of, gf = OF.emit(), GF.emit().download()
    assert_same_groups(gf, of, len(gi))
    return gf


Q1_AGGS = [("sum", "dec", "sum_qty"), ("avg", "dec", "avg_qty"), ("avg", "f", "avg_f"), ("count", None, "count_order"),
           ("sum", "k32", "sum_i"), ("avg", "k32", "avg_i"), ("count", "k32", "cnt_i"), ("sum", "f", "sum_f"),
           ("sum", "u8", "sum_u")]


@pytest.mark.parametrize("groups", [["flag"], ["flag", "small"], ["small"], ["d"], ["d", "flag"], ["k32"], ["date", "b"],
                                    ["i16", "u8"], ["k64"], ["dec"]])
def test_partial_final_match_oracle(bb, ctx, oracle, groups):
    rng = np.random.default_rng(17)
    whole = random_batch(rng, 40_000)
    batches = [whole.slice(i, 8192) for i in range(0, whole.num_rows, 8192)]
    run_both(bb, ctx, oracle, batches, groups, Q1_AGGS, n_partials=2)


def test_no_nulls_then_nulls(bb, ctx, oracle):
    rng = np.random.default_rng(3)
    a = random_batch(rng, 5000, null_frac=0)
    b = random_batch(rng, 5000, null_frac=0.3)
    run_both(bb, ctx, oracle, [a, b, a.slice(0, 100)], ["flag", "small"], Q1_AGGS, n_partials=1)


def test_all_null_group_gives_null_sum(bb, ctx, oracle):
    k = pa.array([1, 1, 2, 2, 3], type=pa.int64())
    v = pa.array([None, None, 5, None, 7], type=pa.int64())
    b = pa.RecordBatch.from_arrays([k, v], names=["k", "v"])
    out = run_both(bb, ctx, oracle, [b], ["k"], [("sum", "v", "s"), ("avg", "v", "a"), ("count", "v", "c")], 1)
    t = pa.Table.from_batches([out]).sort_by("k")
    assert t["s"].to_pylist() == [None, 5, 7] and t["c"].to_pylist() == [0, 1, 1] and t["a"].to_pylist() == [None, 5.0, 7.0]


def test_high_cardinality_grows_table(bb, ctx, oracle):
    rng = np.random.default_rng(9)
    n = 400_000
    b = pa.RecordBatch.from_arrays([pa.array(rng.integers(0, 300_000, n), type=pa.int64()),
                                    pa.array(rng.integers(-100, 100, n), type=pa.int64()),
                                    pa.array(rng.random(n))], names=["k", "v", "f"])
    run_both(bb, ctx, oracle, [b.slice(0, 250_000), b.slice(250_000)], ["k"],
             [("sum", "v", "s"), ("count", None, "c"), ("avg", "f", "a")], 2)


@pytest.mark.parametrize("groups", [["k64"], ["d", "k64"], ["k64", "dec"]])
def test_radix_prepass_gives_the_same_groups(bb, ctx, oracle, groups, monkeypatch):
    """The radix pre-pass of the two-pass path (rows moved into table-bucket order before intern / accumulate; opt-in, see
    profiles/r2_orderkey.md), forced on for a small input: nullable arguments, a dictionary key, a key wider than 16 bytes (interned ids) and several batches
    into the same table - bit-exact against the oracle, and the path reports itself."""
    monkeypatch.setenv("BGPU_RADIX", "1")
    monkeypatch.setenv("BGPU_RADIX_MIN_MB", "0")
    monkeypatch.setenv("BGPU_RADIX_MIN_ROWS", "1000")
    monkeypatch.setenv("BGPU_RADIX_BUCKETS", "64")
    rng = np.random.default_rng(77)
    whole = random_batch(rng, 120_000)
    k = pa.array(rng.integers(0, 90_000, whole.num_rows), type=pa.int64())
    whole = whole.set_column(whole.schema.names.index("k64"), "k64", k)
    batches = [whole.slice(0, 70_000), whole.slice(70_000)]
    gi = [whole.schema.names.index(c) for c in groups]
    spec = [(op, whole.schema.names.index(c) if c else None, name) for op, c, name in Q1_AGGS]
    gp = bb.GpuAggregateExec(ctx, bb.AggregateMode.Partial, gi, spec, whole.schema)
    want = oracle.GroupedAggregate("partial", gi, spec)
    for b in batches:
        gp.update(b)
        assert gp.last_path == "global_radix", gp.last_path
        want.update(b)
    assert_same_groups(gp.emit().download(), want.emit(), len(gi))


def test_int64_sum_wraps(bb, ctx, oracle):
    b = pa.RecordBatch.from_arrays([pa.array([0, 0, 0, 1], type=pa.int64()),
                                    pa.array([2**62, 2**62, 2**62, -5], type=pa.int64())], names=["k", "v"])
    out = run_both(bb, ctx, oracle, [b], ["k"], [("sum", "v", "s")], 1)
    assert sorted(out.column(1).to_pylist()) == [-(2**62), -5]


def test_long_string_key_switches_to_interned_keys(bb, ctx, oracle):
    """A string key beyond its inline field in the FIRST batch of a stream: the key layout switches to interned ids
    (GroupValuesRows has no width limit) and the batch runs again; the result is the oracle's."""
    b = pa.RecordBatch.from_arrays([pa.array(["short", "this string is far too long to pack", "short", None, "",
                                              "this string is far too long to pack"]), pa.array([1, 2, 3, 4, 5, 6])],
                                   names=["k", "v"])
    out = run_both(bb, ctx, oracle, [b, b.slice(1, 3)], ["k"], [("sum", "v", "s"), ("count", None, "n")], 1)
    assert out.num_rows == 4


def test_long_string_key_after_committed_batches_is_an_error(bb, ctx):
    """Later in a stream the table is keyed inline already: the error stands (and names the budget)."""
    short = pa.RecordBatch.from_arrays([pa.array(["a", "bc"]), pa.array([1, 2])], names=["k", "v"])
    long = pa.RecordBatch.from_arrays([pa.array(["this string is far too long to pack"]), pa.array([3])], names=["k", "v"])
    agg = bb.GpuAggregateExec(ctx, bb.AggregateMode.Partial, [0], [("sum", 1, "s")], short.schema)
    agg.update(short)
    with pytest.raises(bb.BgpuError, match="packed-key budget"):
        agg.update(long)


@pytest.mark.parametrize("groups", [["s"], ["s", "flag", "k64"], ["k64", "dec"], ["k64", "small", "date", "s"], ["dec", "s", "b"]])
def test_group_keys_wider_than_16_bytes(bb, ctx, oracle, groups):
    """Keys that do not fit the 16-byte packed key - strings up to 51 bytes, Int64 + Decimal128, four columns - go through
    the interner: bit-exact against the oracle through Partial (two partitions, several batches each) and Final."""
    rng = np.random.default_rng(41)
    whole = random_batch(rng, 60_000, with_dict=False)
    batches = [whole.slice(i, 8192) for i in range(0, whole.num_rows, 8192)]
    run_both(bb, ctx, oracle, batches, groups, Q1_AGGS, n_partials=2)


def test_wide_keys_many_distinct_values_grow_the_interner(bb, ctx, oracle):
    rng = np.random.default_rng(43)
    n = 300_000
    keys = [f"customer#{int(i):09d}-{'x' * int(i % 7)}" for i in rng.integers(0, 200_000, n)]
    b = pa.RecordBatch.from_arrays([pa.array(keys), pa.array(rng.integers(-5, 5, n), type=pa.int64())], names=["k", "v"])
    run_both(bb, ctx, oracle, [b.slice(0, 100_000), b.slice(100_000)], ["k"], [("sum", "v", "s"), ("count", None, "n")], 2)


def test_empty_input(bb, ctx):
    b = random_batch(np.random.default_rng(0), 0, with_dict=False)
    agg = bb.GpuAggregateExec(ctx, bb.AggregateMode.Partial, [b.schema.names.index("small")],
                              [("sum", b.schema.names.index("dec"), "s")], b.schema)
    agg.update(b)
    out = agg.emit().download()
    assert out.num_rows == 0 and out.schema.names == ["small", "s[sum]"]


def test_wide_decimal_sums_few_groups(bb, ctx, oracle):
    """i128 sums whose per-row high-word contribution is large, small-negative and zero, on the thread-private
    single-pass kernel (the private high words are 32-bit with a 64-bit spill word per group)."""
    rng = np.random.default_rng(23)
    n = 50_000
    mag = rng.choice([1, 10**9, 10**19, 10**25, 10**30], n)
    vals = [int(m) * int(x) for m, x in zip(mag, rng.integers(-9999, 9999, n))]
    b = pa.RecordBatch.from_arrays(
        [pa.array(rng.integers(0, 3, n), type=pa.int64()),
         pa.array([decimal.Decimal(v) for v in vals], type=pa.decimal128(38, 0)),
         pa.array([decimal.Decimal(int(x)) for x in rng.integers(-5, 5, n)], type=pa.decimal128(10, 0))],
        names=["k", "wide", "narrow"])
    spec = [("sum", "wide", "sw"), ("sum", "narrow", "sn"), ("count", None, "c"), ("avg", "narrow", "an")]
    out = run_both(bb, ctx, oracle, [b.slice(0, 30_000), b.slice(30_000)], ["k"], spec, 1)
    got = dict(zip(out.column(0).to_pylist(), out.column(1).to_pylist()))
    for k in range(3):
        assert got[k] == decimal.Decimal(sum(v for v, kk in zip(vals, b.column(0).to_pylist()) if kk == k))


def zipf_keys(rng, n, domain):
    """Zipf(1)-like keys over 1..domain: floor((domain + 1)^u) (tools/microbench_config5.py)."""
    return np.clip(np.floor(np.exp(rng.random(n) * np.log(domain + 1.0))), 1, domain).astype(np.int64)


@pytest.mark.parametrize("nulls", [False, True])
def test_skewed_keys_global_tier_hot_group_cache(bb, ctx, oracle, nulls):
    """BASELINE config 5 shape: Zipf-skewed Int64 keys over a domain large enough for the two-pass path's global-atomic
    tier, where each CTA keeps hot groups in a shared-memory cache that is folded into the global records at the end:
    sums / counts / decimal averages must stay bit-exact (Float64 within 1e-9), null tracking included."""
    rng = np.random.default_rng(31 + nulls)
    n = 300_000
    batches = []
    for _ in range(2):
        k = zipf_keys(rng, n, 60_000)
        mask = (rng.random(n) < 0.3) if nulls else None
        dec = [decimal.Decimal(int(v)) / 100 for v in rng.integers(-10**9, 10**9, n)]
        batches.append(pa.RecordBatch.from_arrays(
            [pa.array(k), pa.array(rng.integers(1, 16, n), type=pa.int64(), mask=mask),
             pa.array(np.round(rng.random(n) * 100, 6), mask=mask), pa.array(dec, type=pa.decimal128(15, 2), mask=mask)],
            names=["id6", "v2", "v3", "d"]))
    aggs = [("sum", "v2", "s2"), ("sum", "v3", "s3"), ("sum", "d", "sd"), ("avg", "d", "ad"), ("count", "v2", "c2"), ("count", None, "c")]
    gp = bb.GpuAggregateExec(ctx, bb.AggregateMode.Partial, [0], [(f, None if c is None else batches[0].schema.names.index(c), nm)
                                                                 for f, c, nm in aggs], batches[0].schema)
    gp.update(batches[0])
    assert gp.last_path == "global", gp.last_path  # the tier this test is about
    out = run_both(bb, ctx, oracle, batches, ["id6"], aggs, n_partials=1)
    hot = pa.Table.from_batches([out]).filter(pc.equal(pa.Table.from_batches([out]).column("id6"), 1))
    assert hot.num_rows == 1 and hot.column("c")[0].as_py() > 0.04 * 2 * n  # the hottest group really is hot


@pytest.mark.parametrize("domain,nulls,zipf", [(100, False, True), (100, True, True), (100, True, False), (700, True, True)])
def test_shared_rung_warp_private_copies(bb, ctx, oracle, domain, nulls, zipf):
    """db-benchmark q4 shape (k = 1e2, and 700 groups so that the capacity has to grow): the multi-CTA shared rung with one
    accumulator copy per warp - equal group ids inside a warp take turns, everything else is a plain read-modify-write.
    Zipf keys put the same group in a third of a warp's lanes. Bit-exact sums / counts, Float64 within 1e-9."""
    rng = np.random.default_rng(domain + 2 * nulls + zipf)
    n = 400_000
    batches = []
    for _ in range(2):
        k = zipf_keys(rng, n, domain) if zipf else rng.integers(1, domain + 1, n)
        mask = (rng.random(n) < 0.3) if nulls else None
        dec = [decimal.Decimal(int(v)) / 100 for v in rng.integers(-10**9, 10**9, n)]
        batches.append(pa.RecordBatch.from_arrays(
            [pa.array(k, type=pa.int64()), pa.array(rng.integers(1, 6, n), type=pa.int64(), mask=mask),
             pa.array(rng.integers(1, 16, n), type=pa.int64()), pa.array(np.round(rng.random(n) * 100, 6), mask=mask),
             pa.array(dec, type=pa.decimal128(15, 2), mask=mask)],
            names=["id4", "v1", "v2", "v3", "d"]))
    aggs = [("avg", "v1", "a1"), ("avg", "v2", "a2"), ("avg", "v3", "a3"), ("sum", "d", "sd"), ("count", "v1", "c1"), ("count", None, "c")]
    gp = bb.GpuAggregateExec(ctx, bb.AggregateMode.Partial, [0], [(f, None if c is None else batches[0].schema.names.index(c), nm)
                                                                 for f, c, nm in aggs], batches[0].schema)
    gp.update(batches[0])
    # (700 groups of this many accumulators no longer fit two CTAs per SM: the one-CTA shared rung takes them)
    assert gp.last_path == ("fused_shared2" if domain <= 128 else "fused_shared1"), gp.last_path
    run_both(bb, ctx, oracle, batches, ["id4"], aggs, n_partials=1)


def test_partial_that_does_not_reduce_passes_rows_through(bb, ctx, oracle):
    """A first batch of mostly distinct keys: the Partial aggregate emits its input rows as state rows (key, sum = the
    argument, count = 1; what DataFusion later called skip_partial_aggregation) and the FinalPartitioned side does the
    only aggregation. The final result equals the oracle's Partial -> Final, bit for bit; with BGPU_NO_PASSTHROUGH the
    same input takes the two-pass path and gives the same result."""
    import decimal
    import os

    rng = np.random.default_rng(31)
    n = (1 << 20) + 77
    b = pa.RecordBatch.from_arrays(
        [pa.array(rng.integers(0, 3_000_000, n), type=pa.int64()),
         pa.array([decimal.Decimal(int(x)).scaleb(-2) for x in rng.integers(100, 5000, n)], type=pa.decimal128(15, 2)),
         pa.array(rng.integers(-10**6, 10**6, n), type=pa.int64()), pa.array(rng.random(n))],
        names=["k", "d", "v", "f"])
    aggs = [("sum", "d", "sd"), ("avg", "d", "ad"), ("sum", "v", "sv"), ("count", "v", "cv"), ("avg", "f", "af"), ("count", None, "n")]
    gi, spec = [0], [(f, (None if c is None else b.schema.names.index(c)), nm) for f, c, nm in aggs]
    op = oracle.GroupedAggregate("partial", gi, spec)
    op.update(b)
    of = oracle.GroupedAggregate("final", [0], spec, arg_types=[b.schema.field(c).type if c is not None else None for _, c, _ in spec])
    of.update(op.emit())
    want = of.emit()
    for skip in (False, True):
        if skip:
            os.environ["BGPU_NO_PASSTHROUGH"] = "1"
        try:
            gp = bb.GpuAggregateExec(ctx, bb.AggregateMode.Partial, gi, spec, b.schema)
            gp.update(bb.DeviceBatch.upload(ctx, b))
            assert gp.last_path == ("global" if skip else "passthrough")
            state = gp.emit()
            assert (state.num_rows == n) == (not skip)
            gf = bb.GpuAggregateExec(ctx, bb.AggregateMode.FinalPartitioned, gi, spec, b.schema)
            gf.update(state)
            assert_same_groups(gf.emit().download(), want, 1)
        finally:
            os.environ.pop("BGPU_NO_PASSTHROUGH", None)
