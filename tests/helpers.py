"""Shared builders for parity tests."""
import decimal
import json
import os

import numpy as np
import pyarrow as pa

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def golden_hash_cases():
    """[(name, pyarrow arrays (key columns), expected u64 hashes)] from tests/golden/ahash_vectors.json."""
    v = json.load(open(os.path.join(GOLDEN, "ahash_vectors.json")))
    ctx = decimal.Context(prec=60)
    cases = []
    for name, t in (("i64", pa.int64()), ("i32", pa.int32()), ("u32", pa.uint32()), ("i16", pa.int16()),
                    ("i8", pa.int8()), ("f64", pa.float64()), ("utf8", pa.string())):
        cases.append((name, [pa.array([x for x, _ in v[name]], type=t)], [h for _, h in v[name]]))
    cases.append(("bool", [pa.array([bool(x) for x, _ in v["bool"]])], [h for _, h in v["bool"]]))
    cases.append(("binary", [pa.array([bytes.fromhex(x) for x, _ in v["binary"]], type=pa.binary())],
                  [h for _, h in v["binary"]]))
    cases.append(("decimal128", [pa.array([ctx.create_decimal(x) for x, _ in v["decimal128"]],
                                          type=pa.decimal128(38, 0))], [h for _, h in v["decimal128"]]))
    cases.append(("i64_utf8", [pa.array([a for a, _, _ in v["i64_utf8"]], type=pa.int64()),
                               pa.array([s for _, s, _ in v["i64_utf8"]])], [h for _, _, h in v["i64_utf8"]]))
    # dictionary-encoded strings hash like plain strings
    cases.append(("dict_utf8", [pa.array([x for x, _ in v["utf8"]]).dictionary_encode()], [h for _, h in v["utf8"]]))
    return cases


def random_batch(rng, n, null_frac=0.1, with_strings=True, with_dict=True):
    """A batch covering every column type on the path, with nulls."""
    def mask():
        return rng.random(n) < null_frac if null_frac else None
    cols = {
        "k64": pa.array(rng.integers(-2**62, 2**62, n), type=pa.int64(), mask=mask()),
        "k32": pa.array(rng.integers(-1000, 1000, n), type=pa.int32(), mask=mask()),
        "small": pa.array(rng.integers(0, 7, n), type=pa.int64()),
        "f": pa.array(rng.random(n), mask=mask()),
        "u8": pa.array(rng.integers(0, 255, n), type=pa.uint8(), mask=mask()),
        "i16": pa.array(rng.integers(-3000, 3000, n), type=pa.int16()),
        "date": pa.array(rng.integers(8000, 11000, n).astype("int32"), type=pa.date32()),
        "dec": pa.array([decimal.Decimal(int(x)) / 100 for x in rng.integers(-10**12, 10**12, n)],
                        type=pa.decimal128(15, 2), mask=mask()),
        "b": pa.array(rng.random(n) < 0.5, mask=mask()),
    }
    if with_strings:
        words = ["", "a", "bc", "def", "ghij", "klmnopqr", "stuvwxyz0", "0123456789abcdef", "0123456789abcdefg" * 3]
        cols["s"] = pa.array([words[i] for i in rng.integers(0, len(words), n)], type=pa.string(), mask=mask())
        cols["flag"] = pa.array(rng.choice(["A", "N", "R"], n).tolist(), type=pa.string())
    if with_dict:
        cols["d"] = pa.array([f"id{int(i):03d}" for i in rng.integers(1, 100, n)], type=pa.string(), mask=mask()).dictionary_encode()
    return pa.RecordBatch.from_arrays(list(cols.values()), names=list(cols.keys()))


def expr_to_tree(e):
    """ballista_b200.expr tree -> the oracle's nested-tuple form."""
    from ballista_b200 import expr as X

    if e is None or isinstance(e, int):
        return e
    if isinstance(e, X.Column):
        return ("col", e.index)
    if isinstance(e, X.Literal):
        typ = {X.T_INT64: "int64", X.T_FLOAT64: "float64", X.T_DECIMAL128: "decimal128", X.T_DATE32: "date32"}[e.type]
        return ("lit", e.value, typ, e.precision, e.scale)
    op = {X.X_ADD: "+", X.X_SUB: "-", X.X_MUL: "*", X.X_LT: "<", X.X_LE: "<=", X.X_GT: ">", X.X_GE: ">=", X.X_EQ: "=",
          X.X_NE: "!=", X.X_AND: "and"}[e.op]
    return (op, expr_to_tree(e.left), expr_to_tree(e.right))


def lineitem_batch(rng, n, null_frac=0.0):
    """Synthetic TPC-H lineitem slice with the reference's schema (/root/reference/benchmarks/src/bin/tpch.rs:923-940):
    Decimal128(15,2) measures, Utf8 flags, Date32 ship date; distributions per SURVEY.md §8(d)."""
    def dec(lo, hi):
        vals = rng.integers(lo, hi + 1, n)
        mask = rng.random(n) < null_frac if null_frac else None
        return pa.array([decimal.Decimal(int(v)).scaleb(-2) for v in vals], type=pa.decimal128(15, 2), mask=mask)
    ship = rng.integers(8036, 10561 + 1, n).astype("int32")  # 1992-01-02 .. 1998-12-01
    receipt = ship + rng.integers(1, 31, n)
    cutoff = 9298  # 1995-06-17
    rf = np.where(receipt <= cutoff, np.where(rng.random(n) < 0.5, "R", "A"), "N")
    ls = np.where(ship > cutoff, "O", "F")
    return pa.RecordBatch.from_arrays(
        [dec(100, 5000), dec(90000, 10495000), dec(0, 10), dec(0, 8), pa.array(rf.tolist(), type=pa.string()),
         pa.array(ls.tolist(), type=pa.string()), pa.array(ship, type=pa.date32())],
        names=["l_quantity", "l_extendedprice", "l_discount", "l_tax", "l_returnflag", "l_linestatus", "l_shipdate"])


def q1_plan():
    """TPC-H Q1's Partial aggregate over lineitem_batch columns (/root/reference/benchmarks/queries/q1.sql):
    returns (group_by, aggr_expr, filter) in ballista_b200.expr form."""
    from ballista_b200.expr import col, lit

    qty, price, disc, tax = col(0), col(1), col(2), col(3)
    disc_price = price * (lit(1) - disc)
    aggs = [("sum", qty, "sum_qty"), ("sum", price, "sum_base_price"), ("sum", disc_price, "sum_disc_price"),
            ("sum", disc_price * (lit(1) + tax), "sum_charge"), ("avg", qty, "avg_qty"), ("avg", price, "avg_price"),
            ("avg", disc, "avg_disc"), ("count", None, "count_order")]
    filt = col(6) <= lit(10471, 3)  # l_shipdate <= date '1998-12-01' - interval '90' day = 1998-09-02
    return [4, 5], aggs, filt
