"""Golden fixture: TPC-H Q1 over the reference's own lineitem test data
(/root/reference/ballista/scheduler/testdata/lineitem/partition0.tbl and partition1.tbl, 10 rows each - the files the
reference's planner tests register as the two-partition `lineitem` table, ballista/scheduler/src/test_utils.rs), with
the expected answer computed HERE by plain Python integer arithmetic, independent of oracle/ and of the CUDA path.

Semantics restated from DataFusion 34: l_quantity, l_extendedprice, l_discount, l_tax are
Decimal128(15,2); `a * b` of two decimals multiplies the unscaled values and adds the scales; `1 - d` / `1 + t` with the
Int64 literal coerced to the decimal's scale; sum(decimal(p,s)) -> decimal(min(38,p+10), s); avg(decimal(p,s)) ->
decimal(min(38,p+4), s+4) = (sum * 10^4) / count truncated toward zero; count -> Int64. Q1's filter:
l_shipdate <= date '1998-12-01' - interval '90' day = 1998-09-02 (benchmarks/queries/q1.sql).

Run in the build container (needs /root/reference); writes tests/golden/q1_lineitem_testdata.json.
"""
import datetime
import json
import os

REF = "/root/reference/ballista/scheduler/testdata/lineitem"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "q1_lineitem_testdata.json")
EPOCH = datetime.date(1970, 1, 1)
CUTOFF = (datetime.date(1998, 9, 2) - EPOCH).days


def cents(s):
    whole, _, frac = s.partition(".")
    frac = (frac + "00")[:2]
    sign = -1 if whole.startswith("-") else 1
    return sign * (abs(int(whole)) * 100 + int(frac))


def trunc_div(a, b):
    q = abs(a) // b
    return q if a >= 0 else -q


partitions = []
for name in ("partition0.tbl", "partition1.tbl"):
    rows = []
    for line in open(os.path.join(REF, name)):
        f = line.rstrip("\n").split("|")
        rows.append({"l_quantity": cents(f[4]), "l_extendedprice": cents(f[5]), "l_discount": cents(f[6]), "l_tax": cents(f[7]),
                     "l_returnflag": f[8], "l_linestatus": f[9],
                     "l_shipdate": (datetime.date.fromisoformat(f[10]) - EPOCH).days})
    partitions.append(rows)

groups = {}
for rows in partitions:
    for r in rows:
        if r["l_shipdate"] > CUTOFF:
            continue
        g = groups.setdefault((r["l_returnflag"], r["l_linestatus"]),
                              {"sum_qty": 0, "sum_base_price": 0, "sum_disc_price": 0, "sum_charge": 0, "sum_disc": 0, "count_order": 0})
        disc_price = r["l_extendedprice"] * (100 - r["l_discount"])        # scale 4
        g["sum_qty"] += r["l_quantity"]
        g["sum_base_price"] += r["l_extendedprice"]
        g["sum_disc_price"] += disc_price
        g["sum_charge"] += disc_price * (100 + r["l_tax"])                 # scale 6
        g["sum_disc"] += r["l_discount"]
        g["count_order"] += 1
expected = []
for (rf, ls), g in sorted(groups.items()):
    n = g["count_order"]
    expected.append({"l_returnflag": rf, "l_linestatus": ls,
                     "sum_qty": g["sum_qty"], "sum_base_price": g["sum_base_price"],      # unscaled, scale 2
                     "sum_disc_price": g["sum_disc_price"],                                # unscaled, scale 4
                     "sum_charge": g["sum_charge"],                                        # unscaled, scale 6
                     "avg_qty": trunc_div(g["sum_qty"] * 10**4, n),                        # unscaled, scale 6
                     "avg_price": trunc_div(g["sum_base_price"] * 10**4, n),
                     "avg_disc": trunc_div(g["sum_disc"] * 10**4, n),
                     "count_order": n})
json.dump({"source": [REF + "/partition0.tbl", REF + "/partition1.tbl"], "units": "decimals as unscaled integers (scale 2 in the input)",
           "partitions": partitions, "expected": expected}, open(OUT, "w"), indent=1)
print(f"wrote {OUT}: {sum(len(p) for p in partitions)} rows, {len(expected)} groups")
