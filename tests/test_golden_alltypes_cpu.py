"""The only aggregate VALUES the reference's own test-suite pins - SUM(id) = 28, AVG(id) = 3.5, COUNT(id) = 8 over
testdata/alltypes_plain.parquet (/root/reference/ballista/client/src/context.rs:784-815) - reproduced by the oracle as
Partial (two input partitions) -> Final, with a constant group key and with the table's own columns as keys."""
import pyarrow as pa
import pytest

from helpers import alltypes_plain

AGGS = [("sum", 1, "SUM(test.id)"), ("avg", 1, "AVG(test.id)"), ("count", 1, "COUNT(test.id)")]


def partial_final(oracle, batch, key):
    states = []
    for part in (batch.slice(0, 4), batch.slice(4)):
        op = oracle.GroupedAggregate("partial", [key], AGGS)
        op.update(part)
        states.append(op.emit())
    fin = oracle.GroupedAggregate("final", [0], [(f, 0, n) for f, _, n in AGGS], arg_types=[pa.int64()] * 3)
    for s in states:
        fin.update(s)
    return fin.emit()


def test_reference_asserted_aggregates(oracle):
    batch, want = alltypes_plain()
    out = partial_final(oracle, batch, 0)
    assert out.num_rows == 1
    row = out.to_pylist()[0]
    assert row["SUM(test.id)"] == want["SUM(id)"] == 28 and out.schema.field("SUM(test.id)").type == pa.int64()
    assert row["AVG(test.id)"] == want["AVG(id)"] == 3.5 and out.schema.field("AVG(test.id)").type == pa.float64()
    assert row["COUNT(test.id)"] == want["COUNT(id)"] == 8 and out.schema.field("COUNT(test.id)").type == pa.int64()


@pytest.mark.parametrize("key", ["bool_col", "string_col", "bigint_col"])
def test_grouped_results_add_up_to_the_asserted_totals(oracle, key):
    batch, want = alltypes_plain()
    out = partial_final(oracle, batch, batch.schema.names.index(key))
    rows = out.to_pylist()
    assert len(rows) == 2
    assert sum(r["SUM(test.id)"] for r in rows) == want["SUM(id)"]
    assert sum(r["COUNT(test.id)"] for r in rows) == want["COUNT(id)"]
    assert sum(r["AVG(test.id)"] * r["COUNT(test.id)"] for r in rows) / want["COUNT(id)"] == want["AVG(id)"]
    assert sorted(r["AVG(test.id)"] for r in rows) == [3.0, 4.0]   # ids {0,2,4,6} and {1,3,5,7}
