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


def _xxh32_short(data, seed=0):
    """XXH32 of fewer than 16 bytes (the LZ4 frame descriptor): the header checksum byte is (xxh32 >> 8) & 0xff."""
    P1, P2, P3, P4, P5, M = 2654435761, 2246822519, 3266489917, 668265263, 374761393, 0xFFFFFFFF
    assert len(data) < 16

    def rotl(x, r):
        return ((x << r) | (x >> (32 - r))) & M
    h = (seed + P5 + len(data)) & M
    i = 0
    while i + 4 <= len(data):
        h = (h + int.from_bytes(data[i:i + 4], "little") * P3) & M
        h = (rotl(h, 17) * P4) & M
        i += 4
    while i < len(data):
        h = (h + data[i] * P5) & M
        h = (rotl(h, 11) * P1) & M
        i += 1
    h ^= h >> 15
    h = (h * P2) & M
    h ^= h >> 13
    h = (h * P3) & M
    h ^= h >> 16
    return h


def independent_frame(data, block=65536, block_checksum=False, content_size=False, block_compress=None):
    """An LZ4 frame with block-independent 64 KB blocks, as lz4_flex's FrameEncoder (arrow-ipc's codec in the reference
    build) writes by default; built from Arrow's raw LZ4 block codec. The header checksum is real; block checksums are placeholders."""
    flg = 0x40 | 0x20 | (0x10 if block_checksum else 0) | (0x08 if content_size else 0)
    out = bytearray(b"\x04\x22\x4d\x18" + bytes([flg, 0x40]))
    if content_size:
        out += len(data).to_bytes(8, "little")
    out += bytes([(_xxh32_short(bytes(out[4:])) >> 8) & 0xFF])  # header checksum
    for lo in range(0, len(data), block):
        piece = data[lo:lo + block]
        # blocks from Arrow's raw LZ4 codec (liblz4), or from `block_compress` (tests/lz4_flex_ffi.py: lz4_flex itself)
        comp = block_compress(piece) if block_compress else pa.compress(piece, codec="lz4_raw", asbytes=True)
        if len(comp) >= len(piece):
            out += (len(piece) | 0x80000000).to_bytes(4, "little") + piece
        else:
            out += len(comp).to_bytes(4, "little") + comp
        if block_checksum:
            out += b"\0\0\0\0"
    out += b"\0\0\0\0"
    return bytes(out)


def reference_style_lz4_stream(batches, block=65536, block_compress=None):
    """An Arrow IPC stream of `batches` (no dictionary columns) framed the way the reference's writer frames a shuffle file
    (arrow-ipc StreamWriter with LZ4_FRAME, /root/reference/ballista/core/src/execution_plans/shuffle_writer.rs:365-376):
    every buffer = int64 uncompressed length + an LZ4 frame of block-independent 64 KB blocks (lz4_flex's FrameEncoder
    default), or length -1 + the raw bytes when compression does not pay; buffers padded to 64 bytes. Metadata comes from
    libbgpu's own flatbuffer builder (ballista_b200.ipc)."""
    from ballista_b200 import ipc

    def raw_buffers(col):
        n = len(col)
        bufs = col.buffers()
        out = []
        validity = bufs[0]
        if validity is not None and col.null_count:
            assert col.offset == 0
            out.append(validity.to_pybytes()[: (n + 7) // 8])
        else:
            out.append(b"")
        t = col.type
        if pa.types.is_string(t) or pa.types.is_binary(t):
            offs = np.frombuffer(bufs[1], dtype=np.int32, count=n + 1) if n else np.zeros(1, np.int32)
            out.append(offs.tobytes())
            out.append(bufs[2].to_pybytes()[: int(offs[-1])] if bufs[2] is not None else b"")
        elif pa.types.is_boolean(t):
            out.append(bufs[1].to_pybytes()[: (n + 7) // 8])
        else:
            out.append(bufs[1].to_pybytes()[: n * t.byte_width])
        return out

    schema = batches[0].schema
    stream = bytearray(ipc.schema_message(schema))
    for b in batches:
        b = pa.RecordBatch.from_arrays([pa.concat_arrays([c]) for c in b.columns], schema=schema)  # offset 0, compact
        body, buffers, nodes = bytearray(), [], []
        for col in b.columns:
            nodes.append((len(col), col.null_count))
            for raw in raw_buffers(col):
                if not raw:
                    buffers.append((len(body), 0))
                    continue
                frame = independent_frame(raw, block, block_compress=block_compress)
                enc = (len(raw).to_bytes(8, "little", signed=True) + frame) if len(frame) < len(raw) else \
                    ((-1).to_bytes(8, "little", signed=True) + raw)
                buffers.append((len(body), len(enc)))
                body += enc
                body += b"\0" * (-len(body) % 64)
        stream += ipc.record_batch_metadata(b.num_rows, nodes, buffers, len(body), compression=0)
        stream += body
    return bytes(stream + ipc.EOS)


def q1_testdata():
    """The reference's own two-partition lineitem test data as Q1 input batches + the expected answer
    (tests/golden/q1_lineitem_testdata.json, written by tests/golden/make_q1_testdata.py with plain integer arithmetic)."""
    d = json.load(open(os.path.join(GOLDEN, "q1_lineitem_testdata.json")))
    dec = pa.decimal128(15, 2)
    batches = []
    for rows in d["partitions"]:
        cols = [pa.array([decimal.Decimal(r[k]).scaleb(-2) for r in rows], type=dec)
                for k in ("l_quantity", "l_extendedprice", "l_discount", "l_tax")]
        cols += [pa.array([r[k] for r in rows], type=pa.string()) for k in ("l_returnflag", "l_linestatus")]
        cols.append(pa.array(np.array([r["l_shipdate"] for r in rows], dtype="int32"), type=pa.date32()))
        batches.append(pa.RecordBatch.from_arrays(cols, names=["l_quantity", "l_extendedprice", "l_discount", "l_tax", "l_returnflag",
                                                               "l_linestatus", "l_shipdate"]))
    return batches, d["expected"]


def q1_result_as_unscaled(table):
    """A Q1 result (pyarrow Table / RecordBatch) as the golden file's rows: decimals as unscaled integers, sorted by key."""
    out = []
    for row in pa.Table.from_batches([table]).to_pylist() if isinstance(table, pa.RecordBatch) else table.to_pylist():
        r = {}
        for k, v in row.items():
            r[k] = int(v.scaleb(-v.as_tuple().exponent)) if isinstance(v, decimal.Decimal) else v
        out.append(r)
    return sorted(out, key=lambda r: (r["l_returnflag"], r["l_linestatus"]))


def record_batch_message_buffers(msg):
    """The body buffers of ONE encapsulated RecordBatch message (bytes: continuation marker, metadata length, flatbuffer
    metadata, body): returns (compressed?, [bytes of each buffer as it lies in the body]). A small flatbuffer walk over
    Message{header_type@1, header@2, bodyLength@3} -> RecordBatch{length@0, nodes@1, buffers@2, compression@3}."""
    import struct

    msg = bytes(msg)
    assert msg[:4] == b"\xff\xff\xff\xff"
    meta_len = struct.unpack_from("<i", msg, 4)[0]
    meta, body = msg[8:8 + meta_len], msg[8 + meta_len:]

    def table(pos):
        vt = pos - struct.unpack_from("<i", meta, pos)[0]
        vsize = struct.unpack_from("<H", meta, vt)[0]

        def field(i):
            slot = 4 + 2 * i
            if slot + 2 > vsize:
                return 0
            off = struct.unpack_from("<H", meta, vt + slot)[0]
            return pos + off if off else 0
        return field

    def indirect(at):
        return at + struct.unpack_from("<I", meta, at)[0]
    message = table(indirect(0))
    assert meta[message(1)] == 3, "not a RecordBatch message"
    rb = table(indirect(message(2)))
    vec = indirect(rb(2))
    n = struct.unpack_from("<I", meta, vec)[0]
    bufs = [struct.unpack_from("<qq", meta, vec + 4 + 16 * i) for i in range(n)]
    return bool(rb(3)), [body[o:o + ln] for o, ln in bufs]


def alltypes_plain():
    """The reference's alltypes_plain test table as the aggregate input (tests/golden/alltypes_plain_aggregates.json) and
    the answers the reference's own test asserts. Columns: g (constant Int64 key), id (Int64: the planner's coercion of the
    Int32 column), bool_col, bigint_col, double_col, string_col."""
    d = json.load(open(os.path.join(GOLDEN, "alltypes_plain_aggregates.json")))
    c = d["columns"]
    n = len(c["id"])
    batch = pa.RecordBatch.from_arrays(
        [pa.array([0] * n, type=pa.int64()), pa.array(c["id"], type=pa.int64()), pa.array(c["bool_col"]),
         pa.array(c["bigint_col"], type=pa.int64()), pa.array(c["double_col"], type=pa.float64()), pa.array(c["string_col"])],
        names=["g", "id", "bool_col", "bigint_col", "double_col", "string_col"])
    return batch, d["reference_asserts"]


def decode_dictionaries(t):
    """A Table / RecordBatch with every dictionary column cast to its value type (logical comparison: two encodings of the
    same rows differ in dictionaries and indices)."""
    import pyarrow as pa

    tbl = pa.Table.from_batches([t]) if isinstance(t, pa.RecordBatch) else t
    cols, fields = [], []
    for f, c in zip(tbl.schema, tbl.columns):
        if pa.types.is_dictionary(f.type):
            c = c.cast(f.type.value_type)
            f = pa.field(f.name, f.type.value_type, f.nullable)
        cols.append(c)
        fields.append(f)
    return pa.Table.from_arrays(cols, schema=pa.schema(fields))
