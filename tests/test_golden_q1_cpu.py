"""The oracle against a golden answer on the reference's own fixture: TPC-H Q1 over
/root/reference/ballista/scheduler/testdata/lineitem/partition{0,1}.tbl (the two-partition `lineitem` table of the
reference's planner tests), expected values from tests/golden/make_q1_testdata.py (plain Python integers, independent of
the oracle). One Partial aggregate per input partition, then FinalPartitioned - the plan of planner.rs:376-427."""
import pyarrow as pa

from helpers import expr_to_tree, q1_plan, q1_result_as_unscaled, q1_testdata


def test_oracle_q1_on_reference_testdata(oracle):
    batches, expected = q1_testdata()
    assert [b.num_rows for b in batches] == [10, 10]
    gi, aggs, filt = q1_plan()
    o_aggs = [(f, expr_to_tree(a), nm) for f, a, nm in aggs]
    states = []
    for b in batches:
        op = oracle.GroupedAggregate("partial", gi, o_aggs, filter=expr_to_tree(filt))
        op.update(b)
        states.append(op.emit())
    arg_types = [None if f == "count" else states[0].schema.field(f"{nm}[sum]").type for f, _, nm in aggs]
    arg_types = [None if t is None else pa.decimal128(t.precision - 10, t.scale) for t in arg_types]
    fin = oracle.GroupedAggregate("final", list(range(len(gi))), [(f, 0, nm) for f, _, nm in aggs], arg_types=arg_types)
    for s in states:
        fin.update(s)
    got = fin.emit()
    # result types as DataFusion derives them: sum(decimal(15,2)) -> (25,2); products add scales; avg -> scale + 4
    types = {f.name: f.type for f in got.schema}
    assert types["sum_qty"] == pa.decimal128(25, 2) and types["sum_disc_price"].scale == 4 and types["sum_charge"].scale == 6
    assert types["avg_qty"] == pa.decimal128(19, 6) and types["count_order"] == pa.int64()
    assert q1_result_as_unscaled(got) == expected
