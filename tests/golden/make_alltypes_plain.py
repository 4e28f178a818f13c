"""Writes tests/golden/alltypes_plain_aggregates.json: the columns of the reference's own test table
/root/reference/ballista/client/testdata/alltypes_plain.parquet (8 rows) that its aggregate test reads, together with the
answers that test asserts (/root/reference/ballista/client/src/context.rs:736-816, `test_aggregate_func`):

    select SUM("id") from test   -> 28      (:784-793)
    select AVG("id") from test   -> 3.5     (:795-804)
    select COUNT("id") from test -> 8       (:806-815)

These are the only aggregate VALUES the reference's test-suite pins. The queries have no GROUP BY (DataFusion runs them
through AggregateStream); the parity tests reproduce the arithmetic of the Partial -> Final merge with a constant group key
and with the table's own `bool_col` / `string_col` as keys (whose per-group results must add up to the pinned totals).
`id` is Int32 in the file; the planner coerces SUM's argument to Int64 and AVG's to Float64, which is what the fixture
consumers do. Run in the build container only (reads /root/reference): python tests/golden/make_alltypes_plain.py
"""
import json
import os

import pyarrow.parquet as pq

HERE = os.path.dirname(os.path.abspath(__file__))
t = pq.read_table("/root/reference/ballista/client/testdata/alltypes_plain.parquet")
doc = {
    "source": "ballista/client/testdata/alltypes_plain.parquet",
    "columns": {
        "id": t.column("id").to_pylist(),
        "bool_col": t.column("bool_col").to_pylist(),
        "bigint_col": t.column("bigint_col").to_pylist(),
        "double_col": t.column("double_col").to_pylist(),
        "string_col": [b.decode() for b in t.column("string_col").to_pylist()],
    },
    "reference_asserts": {"test": "ballista/client/src/context.rs:736-816 test_aggregate_func",
                          "SUM(id)": 28, "AVG(id)": 3.5, "COUNT(id)": 8},
}
with open(os.path.join(HERE, "alltypes_plain_aggregates.json"), "w") as f:
    json.dump(doc, f, indent=1)
print(doc)
