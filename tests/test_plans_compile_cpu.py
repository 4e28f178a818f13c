"""Every plan shape of the BASELINE configs generates kernels NVRTC accepts for sm_100a - checked without a device through
bgpu_agg_debug_compile (the same planner + code generator + prelude the product runs; only the module load and the launch
are missing): TPC-H Q1 (config 1/4a: fused filter + decimal expressions), the db-benchmark group-bys q1-q5 (config 2/5:
dictionary and Int64 keys, /root/reference/benchmarks/db-benchmark/groupby-datafusion.py:73-141) and the l_orderkey
group-by (config 4b), in Partial and FinalPartitioned mode, with and without validity bitmaps."""
import os
import sys

import pyarrow as pa
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DICT = pa.dictionary(pa.int32(), pa.string())
G1 = pa.schema([("id1", DICT), ("id2", DICT), ("id3", DICT), ("id4", pa.int64()), ("id5", pa.int64()), ("id6", pa.int64()),
                ("v1", pa.int64()), ("v2", pa.int64()), ("v3", pa.float64())])
D15 = pa.decimal128(15, 2)
ORDERKEY = pa.schema([("l_orderkey", pa.int64()), ("l_quantity", D15), ("l_extendedprice", D15)])
PLANS = {
    "g1_q1": (G1, [0], [("sum", 6, "v1")]),
    "g1_q2": (G1, [0, 1], [("sum", 6, "v1")]),
    "g1_q3": (G1, [2], [("sum", 6, "v1"), ("avg", 8, "v3")]),
    "g1_q4": (G1, [3], [("avg", 6, "v1"), ("avg", 7, "v2"), ("avg", 8, "v3")]),
    "g1_q5": (G1, [5], [("sum", 6, "v1"), ("sum", 7, "v2"), ("sum", 8, "v3")]),
    "orderkey": (ORDERKEY, [0], [("sum", 1, "q"), ("sum", 2, "p"), ("count", None, "c")]),
}


@pytest.mark.parametrize("name", list(PLANS))
def test_baseline_plan_shapes_compile(bb, name):
    from ballista_b200.aggregate import debug_compile

    schema, group_by, aggs = PLANS[name]
    # flags: bit 0 = every column carries validity, bit 1 = dictionary keys need a remap, bit 2 = seen tracking
    debug_compile(bb.AggregateMode.Partial, group_by, aggs, schema, flags=0)
    debug_compile(bb.AggregateMode.FinalPartitioned, group_by, aggs, schema, flags=7 if name.startswith("g1_q") else 5)


def test_q1_plan_compiles_in_both_modes_and_geometries(bb):
    sys.path.insert(0, ROOT)
    import bench
    from ballista_b200.aggregate import debug_compile

    gi, aggs, filt = bench.q1_plan()
    schema = bench.lineitem_schema()
    debug_compile(bb.AggregateMode.Partial, gi, aggs, schema, filter=filt, threads=640, rows_per_thread=1, min_ctas=1)
    debug_compile(bb.AggregateMode.Partial, gi, aggs, schema, filter=filt, flags=1)
    debug_compile(bb.AggregateMode.FinalPartitioned, gi, aggs, schema)
