"""The aggregate VALUES the reference's own test-suite pins - SUM(id) = 28, AVG(id) = 3.5, COUNT(id) = 8 over
testdata/alltypes_plain.parquet (/root/reference/ballista/client/src/context.rs:784-815; fixture
tests/golden/alltypes_plain_aggregates.json) - through the CUDA path: one GpuAggregateExec(Partial) per input partition ->
hash repartition of the states -> FinalPartitioned, with a constant group key and with the table's own columns as keys."""
import pyarrow as pa
import pytest

from helpers import alltypes_plain

pytestmark = pytest.mark.gpu

AGGS = [("sum", 1, "SUM(test.id)"), ("avg", 1, "AVG(test.id)"), ("count", 1, "COUNT(test.id)")]


def partial_final(bb, ctx, batch, key):
    partitioner = bb.BatchPartitioner.try_new(bb.Partitioning.Hash([0], 4), ctx)
    fin = bb.GpuAggregateExec(ctx, bb.AggregateMode.FinalPartitioned, [key], AGGS, batch.schema)
    for part in (batch.slice(0, 4), batch.slice(4)):
        agg = bb.GpuAggregateExec(ctx, bb.AggregateMode.Partial, [key], AGGS, batch.schema)
        agg.update(part)
        state, _ = partitioner.partition_device(agg.emit())
        fin.update(state)
    return fin.emit().download()


def test_reference_asserted_aggregates(bb, ctx):
    batch, want = alltypes_plain()
    out = partial_final(bb, ctx, batch, 0)
    assert out.num_rows == 1
    row = out.to_pylist()[0]
    assert row["SUM(test.id)"] == want["SUM(id)"] == 28 and out.schema.field("SUM(test.id)").type == pa.int64()
    assert row["AVG(test.id)"] == want["AVG(id)"] == 3.5 and out.schema.field("AVG(test.id)").type == pa.float64()
    assert row["COUNT(test.id)"] == want["COUNT(id)"] == 8 and out.schema.field("COUNT(test.id)").type == pa.int64()


@pytest.mark.parametrize("key", ["string_col", "bigint_col", "bool_col"])
def test_grouped_results_add_up_to_the_asserted_totals(bb, ctx, key):
    batch, want = alltypes_plain()
    rows = partial_final(bb, ctx, batch, batch.schema.names.index(key)).to_pylist()
    assert len(rows) == 2
    assert sum(r["SUM(test.id)"] for r in rows) == want["SUM(id)"]
    assert sum(r["COUNT(test.id)"] for r in rows) == want["COUNT(id)"]
    assert sorted(r["AVG(test.id)"] for r in rows) == [3.0, 4.0]   # ids {0,2,4,6} and {1,3,5,7}
