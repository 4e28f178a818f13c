"""GPU parity for aggregate arguments given as expressions and the fused filter (TPC-H Q1 shape): the product
evaluates them per row inside the aggregate kernels; the oracle evaluates whole columns generically with Python
integers (arrow-arith decimal rules). Everything here is integer/decimal arithmetic: bit-exact."""
import decimal

import numpy as np
import pyarrow as pa
import pytest

from helpers import expr_to_tree, lineitem_batch, q1_plan
from test_aggregate_gpu import assert_same_groups

pytestmark = pytest.mark.gpu


def run_q1(bb, ctx, oracle, batches, n_partials=2, expect_path=None):
    schema = batches[0].schema
    gi, aggs, filt = q1_plan()
    o_aggs = [(f, expr_to_tree(a), n) for f, a, n in aggs]
    OP = [oracle.GroupedAggregate("partial", gi, o_aggs, filter=expr_to_tree(filt)) for _ in range(n_partials)]
    GP = [bb.GpuAggregateExec(ctx, bb.AggregateMode.Partial, gi, aggs, schema, filter=filt) for _ in range(n_partials)]
    for i, b in enumerate(batches):
        OP[i % n_partials].update(b)
        GP[i % n_partials].update(b)
        if expect_path:
            assert GP[i % n_partials].last_path.startswith(expect_path), GP[i % n_partials].last_path
    ostates = [p.emit() for p in OP]
    gstates = [p.emit().download() for p in GP]
    for o, g_ in zip(ostates, gstates):
        assert_same_groups(g_, o, len(gi))
    arg_types = [ostates[0].schema.field(f"{n}[sum]").type if f != "count" else None for f, _, n in aggs]
    # the Final aggregate sees state columns; argument types are those the Partial evaluated
    arg_types = [None if t is None else pa.decimal128(t.precision - 10, t.scale) for t in arg_types]
    OF = oracle.GroupedAggregate("final", list(range(len(gi))), [(f, 0, n) for f, _, n in aggs], arg_types=arg_types)
    GF = bb.GpuAggregateExec(ctx, bb.AggregateMode.FinalPartitioned, gi, aggs, schema)
    for o, g_ in zip(ostates, gstates):
        OF.update(o)
        GF.update(g_)
    of, gf = OF.emit(), GF.emit().download()
    assert_same_groups(gf, of, len(gi))
    return gf


def test_q1_matches_oracle(bb, ctx, oracle):
    rng = np.random.default_rng(42)
    whole = lineitem_batch(rng, 30_000)
    out = run_q1(bb, ctx, oracle, [whole.slice(i, 8192) for i in range(0, whole.num_rows, 8192)], expect_path="fused")
    assert out.num_rows == 4
    assert out.schema.field("sum_disc_price").type == pa.decimal128(38, 4)
    assert out.schema.field("sum_charge").type == pa.decimal128(38, 6)
    assert out.schema.field("avg_disc").type == pa.decimal128(19, 6)
    assert out.schema.field("sum_qty").type == pa.decimal128(25, 2)


def test_q1_with_nulls(bb, ctx, oracle):
    rng = np.random.default_rng(1)
    a = lineitem_batch(rng, 10_000, null_frac=0.0)
    b = lineitem_batch(rng, 10_000, null_frac=0.2)
    run_q1(bb, ctx, oracle, [a, b, a.slice(0, 1000)], n_partials=1)


def test_q1_one_large_batch(bb, ctx, oracle):
    rng = np.random.default_rng(5)
    run_q1(bb, ctx, oracle, [lineitem_batch(rng, 200_000)], n_partials=1, expect_path="fused_private")


def test_filter_drops_everything(bb, ctx):
    from ballista_b200.expr import col, lit

    b = lineitem_batch(np.random.default_rng(0), 1000)
    agg = bb.GpuAggregateExec(ctx, bb.AggregateMode.Partial, [4], [("count", None, "c")], b.schema,
                              filter=col(6) < lit(0, 3))
    agg.update(b)
    assert agg.emit().download().num_rows == 0


def test_decimal_multiply_overflow_is_an_error(bb, ctx):
    from ballista_b200.expr import col

    big = decimal.Decimal(10) ** 30
    b = pa.RecordBatch.from_arrays([pa.array([1, 1], type=pa.int64()),
                                    pa.array([big, big], type=pa.decimal128(38, 0))], names=["k", "v"])
    agg = bb.GpuAggregateExec(ctx, bb.AggregateMode.Partial, [0], [("sum", col(1) * col(1), "s")], b.schema)
    with pytest.raises(bb.BgpuError) as e:
        agg.update(b)
        agg.emit()
    assert e.value.code == 5  # BGPU_EOVERFLOW


def test_unsupported_expression_shape(bb, ctx):
    from ballista_b200.expr import col

    b = lineitem_batch(np.random.default_rng(0), 10)
    with pytest.raises(bb.BgpuError) as e:
        bb.GpuAggregateExec(ctx, bb.AggregateMode.Partial, [4], [("sum", (col(0) + col(1)) + col(2), "s")], b.schema)
    assert e.value.code == 2  # BGPU_EUNSUPPORTED, no silent fallback


@pytest.mark.parametrize("k,expect", [(4, "fused_private"), (100, "fused_shared2"), (2000, "fused_shared1"), (200_000, "global")])
def test_ladder_by_cardinality(bb, ctx, oracle, k, expect):
    """db-benchmark q1 shape (sum(v1) by id4) at growing cardinality walks the ladder; results stay exact."""
    rng = np.random.default_rng(k)
    n = 300_000
    b = pa.RecordBatch.from_arrays([pa.array(rng.integers(1, k + 1, n), type=pa.int64()),
                                    pa.array(rng.integers(1, 6, n), type=pa.int64()),
                                    pa.array(np.round(rng.random(n) * 100, 6))], names=["id4", "v1", "v3"])
    for aggs, check_path in (([("sum", 1, "v1")], True), ([("sum", 1, "v1"), ("avg", 2, "v3"), ("count", None, "c")], False)):
        O = oracle.GroupedAggregate("partial", [0], aggs)
        G = bb.GpuAggregateExec(ctx, bb.AggregateMode.Partial, [0], aggs, b.schema)
        for piece in (b.slice(0, 200_000), b.slice(200_000)):
            O.update(piece)
            G.update(piece)
        if check_path:
            assert G.last_path.startswith(expect), G.last_path
        assert_same_groups(G.emit().download(), O.emit(), 1)


def test_int64_and_float_expressions(bb, ctx, oracle):
    from ballista_b200.expr import col, lit

    rng = np.random.default_rng(8)
    n = 20_000
    b = pa.RecordBatch.from_arrays([pa.array(rng.integers(0, 5, n), type=pa.int64()),
                                    pa.array(rng.integers(-2**62, 2**62, n), type=pa.int64()),
                                    pa.array(rng.integers(-100, 100, n), type=pa.int64(), mask=rng.random(n) < 0.1),
                                    pa.array(rng.random(n))], names=["k", "a", "b", "f"])
    aggs = [("sum", col(1) * col(2), "wrapping_product"), ("sum", col(1) + col(2), "s"), ("count", col(1) - col(2), "c"),
            ("sum", col(3) * (lit(1.0) - col(3)), "fprod"), ("avg", col(3) * lit(2.0), "favg")]
    O = oracle.GroupedAggregate("partial", [0], [(f, expr_to_tree(a), nm) for f, a, nm in aggs])
    G = bb.GpuAggregateExec(ctx, bb.AggregateMode.Partial, [0], aggs, b.schema)
    O.update(b)
    G.update(b)
    assert_same_groups(G.emit().download(), O.emit(), 1)
