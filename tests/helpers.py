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
