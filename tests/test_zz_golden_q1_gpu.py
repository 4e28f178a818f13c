"""The CUDA path against the golden answer on the reference's own fixture (tests/golden/q1_lineitem_testdata.json: Q1 over
/root/reference/ballista/scheduler/testdata/lineitem/partition{0,1}.tbl, expected values by plain integer arithmetic):
one GpuAggregateExec(Partial) per input partition -> hash repartition of the states -> FinalPartitioned, bit-exact."""
import pyarrow as pa
import pytest

from helpers import q1_plan, q1_result_as_unscaled, q1_testdata

pytestmark = pytest.mark.gpu


def test_gpu_q1_on_reference_testdata(bb, ctx):
    batches, expected = q1_testdata()
    gi, aggs, filt = q1_plan()
    schema = batches[0].schema
    partitioner = bb.BatchPartitioner.try_new(bb.Partitioning.Hash(list(range(len(gi))), 16), ctx)
    fin = bb.GpuAggregateExec(ctx, bb.AggregateMode.FinalPartitioned, gi, aggs, schema)
    for b in batches:
        part = bb.GpuAggregateExec(ctx, bb.AggregateMode.Partial, gi, aggs, schema, filter=filt)
        part.update(b)
        state, _ = partitioner.partition_device(part.emit())
        fin.update(state)
    got = fin.emit().download()
    assert q1_result_as_unscaled(got) == expected
    types = {f.name: f.type for f in got.schema}
    assert types["sum_qty"] == pa.decimal128(25, 2) and types["avg_qty"] == pa.decimal128(19, 6) and types["count_order"] == pa.int64()
