"""ORACLE (test infrastructure only — never imported by the product path).

ctypes + pyarrow glue over oracle/_build/liboracle.so, the CPU restatement of the reference's hot path
(DataFusion v34 create_hashes / BatchPartitioner / take / GroupedHashAggregateStream; SURVEY.md
Appendix A; reference call sites /root/reference/ballista/core/src/execution_plans/shuffle_writer.rs:336-346
and /root/reference/ballista/scheduler/src/planner.rs:376-427).  Parity status (what is pinned by execution and by the reference's fixtures, what is not): see oracle.h.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this.
"""
import ctypes as C
import os
import subprocess

import numpy as np
import pyarrow as pa

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "liboracle.so")

ORC_FIXED, ORC_BOOL, ORC_UTF8, ORC_BINARY, ORC_DICT_I32 = range(5)


class OrcCol(C.Structure):
    pass


OrcCol._fields_ = [
    ("type", C.c_int32),
    ("width", C.c_int32),
    ("values", C.c_void_p),
    ("offsets", C.c_void_p),
    ("validity", C.c_void_p),
    ("offset", C.c_int64),
    ("dict", C.POINTER(OrcCol)),
    ("dict_len", C.c_int64),
]


def build(force=False):
    """Compile the oracle with gcc (building the checker is not using it)."""
    srcs = [os.path.join(_HERE, f) for f in ("oracle.cpp", "pipeline.cpp", "oracle.h", "ahash_fallback.h")]
    if not force and os.path.exists(_SO) and all(os.path.getmtime(_SO) >= os.path.getmtime(s) for s in srcs):
        return _SO
    subprocess.check_call(["make", "-s", "-C", _HERE])
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_SO)
        _lib.orc_take_utf8_offsets.restype = C.c_int64
        _lib.orc_groups_new.restype = C.c_void_p
        _lib.orc_groups_intern.restype = C.c_int64
        _lib.orc_avg_decimal_evaluate.restype = C.c_int64
        _lib.orc_ahash_keyed_bytes.restype = C.c_uint64
        _lib.orc_ahash_keyed_u64s.restype = C.c_uint64
    return _lib


def _addr(buf):
    return buf.address if buf is not None else None


def _fixed_width(t):
    if pa.types.is_decimal128(t):
        return 16
    if pa.types.is_boolean(t):
        return None
    try:
        return t.bit_width // 8
    except ValueError:
        return None


def to_col(arr, keep):
    """pyarrow Array -> OrcCol. `keep` collects objects that must outlive the struct."""
    if isinstance(arr, pa.ChunkedArray):
        arr = arr.combine_chunks()
    keep.append(arr)
    c = OrcCol()
    t = arr.type
    bufs = arr.buffers()
    c.offset = arr.offset
    c.validity = _addr(bufs[0]) if arr.null_count else None
    c.dict = None
    if pa.types.is_dictionary(t):
        assert t.index_type == pa.int32(), "oracle handles Dictionary<Int32,_> only"
        c.type, c.width = ORC_DICT_I32, 4
        c.values = _addr(bufs[1])
        d = to_col(arr.dictionary, keep)
        keep.append(d)
        c.dict = C.pointer(d)
        c.dict_len = len(arr.dictionary)
    elif pa.types.is_string(t) or pa.types.is_binary(t):
        c.type = ORC_UTF8 if pa.types.is_string(t) else ORC_BINARY
        c.offsets = _addr(bufs[1])
        c.values = _addr(bufs[2]) if bufs[2] is not None else None
    elif pa.types.is_boolean(t):
        c.type = ORC_BOOL
        c.values = _addr(bufs[1])
    else:
        w = _fixed_width(t)
        assert w in (1, 2, 4, 8, 16), f"unsupported type {t}"
        c.type, c.width = ORC_FIXED, w
        c.values = _addr(bufs[1])
    return c


def _cols(arrays, keep):
    cs = (OrcCol * len(arrays))(*[to_col(a, keep) for a in arrays])
    keep.append(cs)
    return cs


def create_hashes(arrays):
    """datafusion-common hash_utils.rs create_hashes with RandomState::with_seeds(0,0,0,0)."""
    n = len(arrays[0])
    out = np.zeros(n, dtype=np.uint64)
    keep = []
    rc = lib().orc_create_hashes(_cols(arrays, keep), len(arrays), C.c_int64(n), out.ctypes.data_as(C.c_void_p))
    assert rc == 0
    return out


def ahash_keyed_bytes(keys, data):
    """RandomState{k0..k3}.hash_one(&[u8]) of the restated fallback hasher."""
    ks = (C.c_uint64 * 4)(*keys)
    return int(lib().orc_ahash_keyed_bytes(ks, bytes(data), C.c_uint64(len(data))))


def ahash_keyed_u64s(keys, xs):
    """from_random_state(RandomState{k0..k3}); write_u64(x) for x in xs; finish()."""
    ks = (C.c_uint64 * 4)(*keys)
    return int(lib().orc_ahash_keyed_u64s(ks, (C.c_uint64 * max(len(xs), 1))(*xs), len(xs)))


def partition_indices(hashes, n_partitions):
    n = len(hashes)
    counts = np.zeros(n_partitions, dtype=np.uint64)
    idx = np.zeros(n, dtype=np.uint64)
    hashes = np.ascontiguousarray(hashes, dtype=np.uint64)
    rc = lib().orc_partition_indices(
        hashes.ctypes.data_as(C.c_void_p), C.c_int64(n), C.c_uint32(n_partitions),
        counts.ctypes.data_as(C.c_void_p), idx.ctypes.data_as(C.c_void_p))
    assert rc == 0
    return counts, idx


# ---- arrow-select take.rs restated -------------------------------------------------------------------
def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def take(arr, idx):
    """take(array, UInt64 indices) for the column types on the path, carrying validity."""
    if isinstance(arr, pa.ChunkedArray):
        arr = arr.combine_chunks()
    idx = np.ascontiguousarray(idx, dtype=np.uint64)
    n = len(idx)
    t = arr.type
    bufs = arr.buffers()
    L = lib()
    validity = None
    null_count = 0
    if arr.null_count:
        vb = np.zeros((n + 7) // 8, dtype=np.uint8)
        L.orc_take_bits(C.c_void_p(bufs[0].address), C.c_int64(arr.offset), _p(idx), C.c_int64(n), _p(vb))
        validity = pa.py_buffer(vb.tobytes())
        null_count = n - int(np.unpackbits(vb, bitorder="little")[:n].sum())
    if pa.types.is_dictionary(t):
        keys = take(arr.indices, idx)
        return pa.DictionaryArray.from_arrays(keys, arr.dictionary)
    if pa.types.is_string(t) or pa.types.is_binary(t):
        offs = np.zeros(n + 1, dtype=np.int32)
        total = L.orc_take_utf8_offsets(C.c_void_p(bufs[1].address), C.c_int64(arr.offset), _p(idx), C.c_int64(n),
                                        _p(offs))
        data = np.zeros(max(int(total), 1), dtype=np.uint8)
        if total:
            L.orc_take_utf8_data(C.c_void_p(bufs[1].address), C.c_void_p(bufs[2].address), C.c_int64(arr.offset),
                                 _p(idx), C.c_int64(n), _p(offs), _p(data))
        return pa.Array.from_buffers(t, n, [validity, pa.py_buffer(offs.tobytes()),
                                            pa.py_buffer(data[:int(total)].tobytes())], null_count)
    if pa.types.is_boolean(t):
        vals = np.zeros((n + 7) // 8, dtype=np.uint8)
        L.orc_take_bits(C.c_void_p(bufs[1].address), C.c_int64(arr.offset), _p(idx), C.c_int64(n), _p(vals))
        return pa.Array.from_buffers(t, n, [validity, pa.py_buffer(vals.tobytes())], null_count)
    w = _fixed_width(t)
    out = np.zeros(n * w, dtype=np.uint8)
    if n:
        L.orc_take_fixed(C.c_void_p(bufs[1].address), C.c_int32(w), C.c_int64(arr.offset), _p(idx), C.c_int64(n),
                         _p(out))
    return pa.Array.from_buffers(t, n, [validity, pa.py_buffer(out.tobytes())], null_count)


def partition(batch, key_cols, n_partitions):
    """BatchPartitioner::partition (hash arm): [(partition, RecordBatch)] ascending, empty ones skipped."""
    hashes = create_hashes([batch.column(i) for i in key_cols])
    counts, idx = partition_indices(hashes, n_partitions)
    out, start = [], 0
    for p in range(n_partitions):
        c = int(counts[p])
        if c:
            sel = idx[start:start + c]
            out.append((p, pa.RecordBatch.from_arrays([take(col, sel) for col in batch.columns],
                                                      schema=batch.schema)))
        start += c
    return out, hashes



# ---- PhysicalExpr evaluation restated (arrow-arith 49 numeric.rs decimal_op; DataFusion 34 BinaryExpr) ---------
# Expression trees are nested tuples: ("col", i) | ("lit", value, type[, precision, scale]) | (op, left, right) with
# op in + - * < <= > >= = != and; type in "int64" "float64" "decimal128" "date32". Evaluated generically (any
# nesting), one whole column at a time like the reference — the product instead compiles a restricted shape.
I128_MIN, I128_MAX = -(1 << 127), (1 << 127) - 1


class ArithmeticOverflow(Exception):
    """arrow's ArrowError::ComputeError("Overflow happened on: ...")."""


class _Val:
    def __init__(self, kind, values, valid, precision=0, scale=0, width=8):
        # kind: "int" (wrapping two's complement of `width` bytes), "float", "decimal", "bool"
        self.kind, self.values, self.valid, self.precision, self.scale, self.width = kind, values, valid, precision, scale, width


def _int_precision(width):
    return {1: 3, 2: 5, 4: 10, 8: 20}[width]


def _column_val(arr):
    if isinstance(arr, pa.ChunkedArray):
        arr = arr.combine_chunks()
    n = len(arr)
    valid = np.ones(n, dtype=bool) if arr.null_count == 0 else np.array(arr.is_valid().to_pylist(), dtype=bool)
    t = arr.type
    if pa.types.is_decimal128(t):
        raw = np.frombuffer(arr.buffers()[1], dtype=np.uint64, count=2 * (arr.offset + n))[2 * arr.offset:]
        vals = np.empty(n, dtype=object)
        for i in range(n):
            v = int(raw[2 * i]) | (int(raw[2 * i + 1]) << 64)
            vals[i] = v - (1 << 128) if v >> 127 else v
        return _Val("decimal", vals, valid, t.precision, t.scale)
    if pa.types.is_floating(t):
        return _Val("float", np.nan_to_num(arr.cast(pa.float64()).fill_null(0.0).to_numpy(zero_copy_only=False), nan=np.nan), valid)
    if pa.types.is_date32(t):
        arr = arr.cast(pa.int32())
        t = arr.type
    if pa.types.is_integer(t):
        vals = np.array([int(x) for x in arr.fill_null(0).to_pylist()], dtype=object)
        return _Val("int", vals, valid, width=t.bit_width // 8)
    raise TypeError(f"oracle: expression over {t}")


def _to_decimal(v):
    """DataFusion coercion of an integer operand next to a decimal: Decimal128(int_precision, 0)."""
    if v.kind == "decimal":
        return v
    if v.kind == "int":
        return _Val("decimal", v.values, v.valid, _int_precision(v.width), 0)
    raise TypeError("oracle: cannot mix Float64 and Decimal128")


def _checked(vals, what):
    for x in vals:
        if x < I128_MIN or x > I128_MAX:
            raise ArithmeticOverflow(f"Overflow happened on: {what}")
    return vals


def _wrap(vals, width):
    bits = 8 * width
    out = np.empty(len(vals), dtype=object)
    for i, x in enumerate(vals):
        x &= (1 << bits) - 1
        out[i] = x - (1 << bits) if x >> (bits - 1) else x
    return out


def eval_expr(batch, tree):
    op = tree[0]
    n = batch.num_rows
    if op == "col":
        return _column_val(batch.column(tree[1]))
    if op == "lit":
        value, typ = tree[1], tree[2]
        valid = np.ones(n, dtype=bool)
        if typ == "float64":
            return _Val("float", np.full(n, float(value)), valid)
        vals = np.empty(n, dtype=object)
        vals[:] = int(value)
        if typ == "decimal128":
            return _Val("decimal", vals, valid, tree[3], tree[4])
        return _Val("int", vals, valid, width=4 if typ == "date32" else 8)
    a, b = eval_expr(batch, tree[1]), eval_expr(batch, tree[2])
    valid = a.valid & b.valid
    if op == "and":
        return _Val("bool", a.values & b.values, valid)
    if a.kind == "decimal" or b.kind == "decimal":
        a, b = _to_decimal(a), _to_decimal(b)
        if op in ("+", "-"):
            s = max(a.scale, b.scale)
            p = min(38, s + max(a.precision - a.scale, b.precision - b.scale) + 1)
            x = _checked(a.values * (10 ** (s - a.scale)), "rescale")
            y = _checked(b.values * (10 ** (s - b.scale)), "rescale")
            return _Val("decimal", _checked(x + y if op == "+" else x - y, op), valid, p, s)
        if op == "*":
            s = a.scale + b.scale
            if s > 38:
                raise ValueError("Output scale of decimal multiply would exceed max scale of 38")
            return _Val("decimal", _checked(a.values * b.values, "*"), valid, min(38, a.precision + b.precision + 1), s)
        s = max(a.scale, b.scale)  # comparison at a common scale
        x, y = a.values * (10 ** (s - a.scale)), b.values * (10 ** (s - b.scale))
    elif a.kind == "float" or b.kind == "float":
        x = a.values.astype(np.float64) if a.kind != "float" else a.values
        y = b.values.astype(np.float64) if b.kind != "float" else b.values
        if op in ("+", "-", "*"):
            with np.errstate(all="ignore"):
                return _Val("float", x + y if op == "+" else x - y if op == "-" else x * y, valid)
    else:
        x, y = a.values, b.values
        if op in ("+", "-", "*"):  # DataFusion 34 BinaryExpr: add_wrapping / sub_wrapping / mul_wrapping
            w = max(a.width, b.width)
            return _Val("int", _wrap(x + y if op == "+" else x - y if op == "-" else x * y, w), valid, width=w)
    cmp = {"<": lambda: x < y, "<=": lambda: x <= y, ">": lambda: x > y, ">=": lambda: x >= y,
           "=": lambda: x == y, "!=": lambda: x != y}[op]()
    return _Val("bool", np.asarray(cmp, dtype=bool), valid)


def val_to_array(v):
    """_Val -> pyarrow array (Decimal128(p,s), Int64 or Float64)."""
    n = len(v.values)
    validity, nulls = None, 0
    if not v.valid.all():
        validity = pa.py_buffer(np.packbits(v.valid, bitorder="little").tobytes())
        nulls = int(n - v.valid.sum())
    if v.kind == "decimal":
        raw = np.zeros(2 * n, dtype=np.uint64)
        for i, x in enumerate(v.values):
            x &= (1 << 128) - 1
            raw[2 * i], raw[2 * i + 1] = x & ((1 << 64) - 1), x >> 64
        return pa.Array.from_buffers(pa.decimal128(v.precision, v.scale), n, [validity, pa.py_buffer(raw.tobytes())], nulls)
    if v.kind == "float":
        return pa.Array.from_buffers(pa.float64(), n, [validity, pa.py_buffer(np.ascontiguousarray(v.values, dtype=np.float64).tobytes())], nulls)
    if v.kind == "int":
        t = {1: pa.int8(), 2: pa.int16(), 4: pa.int32(), 8: pa.int64()}[v.width]
        return pa.array([int(x) for x in v.values], type=t, mask=~v.valid if nulls else None)
    raise TypeError(v.kind)


def filter_batch(batch, tree):
    """FilterExec: keep rows whose predicate is true (null = dropped)."""
    v = eval_expr(batch, tree)
    assert v.kind == "bool"
    return batch.filter(pa.array(v.values & v.valid))

# ---- GroupedHashAggregateStream restated --------------------------------------------
def sum_type(t):
    """Sum's result/state type (sum.rs): Int*->Int64, UInt*->UInt64, Float->Float64, Decimal(p,s)->(min(38,p+10),s)."""
    if pa.types.is_decimal128(t):
        return pa.decimal128(min(38, t.precision + 10), t.scale)
    if pa.types.is_signed_integer(t):
        return pa.int64()
    if pa.types.is_unsigned_integer(t):
        return pa.uint64()
    if pa.types.is_floating(t):
        return pa.float64()
    raise TypeError(f"sum over {t}")


def avg_types(t):
    """(sum state type, result type) for Avg (average.rs). Integer inputs are cast to Float64 by the planner."""
    if pa.types.is_decimal128(t):
        return (pa.decimal128(min(38, t.precision + 10), t.scale),
                pa.decimal128(min(38, t.precision + 4), min(38, t.scale + 4)))
    return pa.float64(), pa.float64()


def _np_of(arr, dtype):
    """Zero-copy-ish numpy view of a fixed-width arrow array's values (ignoring validity)."""
    b = arr.buffers()[1]
    w = np.dtype(dtype).itemsize
    return np.frombuffer(b, dtype=dtype, count=(arr.offset + len(arr)), offset=0)[arr.offset:] if w else None


class _Acc:
    """One GroupsAccumulator: state arrays grow with the number of groups."""

    def __init__(self, kind, t):
        self.kind, self.t = kind, t  # kind: sum_i64|sum_u64|sum_i128|sum_f64|count
        self.seen = np.zeros(0, dtype=np.uint8)
        if kind == "sum_i128":
            self.v = np.zeros((0, 2), dtype=np.uint64)
        elif kind == "sum_f64":
            self.v = np.zeros(0, dtype=np.float64)
        else:
            self.v = np.zeros(0, dtype=np.int64)

    def grow(self, n):
        k = n - len(self.seen)
        if k > 0:
            self.seen = np.concatenate([self.seen, np.zeros(k, dtype=np.uint8)])
            self.v = np.concatenate([self.v, np.zeros((k,) + self.v.shape[1:], dtype=self.v.dtype)])

    def update(self, arr, gid, n_groups):
        self.grow(n_groups)
        L = lib()
        n = len(gid)
        if isinstance(arr, pa.ChunkedArray):
            arr = arr.combine_chunks()
        validity = C.c_void_p(arr.buffers()[0].address) if (arr is not None and arr.null_count) else None
        off = C.c_int64(arr.offset if arr is not None else 0)
        if self.kind == "count":
            L.orc_acc_count(validity, off, _p(gid), C.c_int64(n), _p(self.v))
            return
        vals = C.c_void_p(arr.buffers()[1].address)
        fn = {"sum_i64": L.orc_acc_sum_i64, "sum_u64": L.orc_acc_sum_i64, "sum_i128": L.orc_acc_sum_i128,
              "sum_f64": L.orc_acc_sum_f64}[self.kind]
        fn(vals, validity, off, _p(gid), C.c_int64(n), _p(self.v), _p(self.seen))

    def array(self, out_type, nullable_by_seen=True):
        n = len(self.seen)
        validity, nulls = None, 0
        if nullable_by_seen and self.kind != "count" and not self.seen.all():
            validity = pa.py_buffer(np.packbits(self.seen, bitorder="little").tobytes())
            nulls = int(n - self.seen.sum())
        return pa.Array.from_buffers(out_type, n, [validity, pa.py_buffer(self.v.tobytes())], nulls)


def _cast_for(fn, arr):
    t = arr.type
    if fn == "avg" and not pa.types.is_decimal128(t):
        return arr.cast(pa.float64())  # planner coercion: avg(int) -> avg(CAST(x AS Float64))
    if fn == "sum" and not pa.types.is_decimal128(t):
        return arr.cast(sum_type(t))
    return arr


def _kind(t):
    if pa.types.is_decimal128(t):
        return "sum_i128"
    if pa.types.is_floating(t):
        return "sum_f64"
    return "sum_u64" if pa.types.is_unsigned_integer(t) else "sum_i64"


class GroupedAggregate:
    """AggregateExec(mode=Partial | FinalPartitioned) over a stream of batches.

    aggs: [(fn, arg_col, name)], fn in {"sum","count","avg"}; arg_col indexes the *raw input* batch in
    Partial mode (None = count(*)); in Final mode state columns follow the n group columns in order
    (Sum -> [sum]; Count -> [count] Int64; Avg -> [count] UInt64, [sum]) and `arg_types` gives each
    aggregate's original argument type. Output group order = first-seen (the reference's is unspecified).
    """

    def __init__(self, mode, group_cols, aggs, arg_types=None, filter=None):
        assert mode in ("partial", "final")
        self.mode, self.group_cols, self.aggs, self.arg_types = mode, list(group_cols), list(aggs), arg_types
        self.filter = filter  # expression tree (tuples); the FilterExec below the Partial aggregate
        self.groups = C.c_void_p(lib().orc_groups_new())
        self.accs = None
        self.batches = []  # kept to take() group keys at emit
        self.n_groups = 0

    def __del__(self):
        if getattr(self, "groups", None):
            lib().orc_groups_free(self.groups)
            self.groups = None

    def _init_accs(self, batch):
        self.accs, self.in_types = [], []
        col = len(self.group_cols)
        for i, (fn, arg, _name) in enumerate(self.aggs):
            if self.mode == "partial":
                t = self._arg(batch, arg).type if arg is not None else None
            else:
                t = self.arg_types[i]
            self.in_types.append(t)
            if fn == "count":
                self.accs.append([_Acc("sum_i64" if self.mode == "final" else "count", pa.int64())])
            elif fn == "sum":
                st = sum_type(t)
                self.accs.append([_Acc(_kind(st), st)])
            elif fn == "avg":
                st, _ = avg_types(t)
                self.accs.append([_Acc("sum_u64" if self.mode == "final" else "count", pa.uint64()),
                                  _Acc(_kind(st), st)])
            else:
                raise ValueError(fn)

    @staticmethod
    def _arg(batch, arg):
        """Aggregate argument: a column index or an expression tree evaluated over the batch."""
        return val_to_array(eval_expr(batch, arg)) if isinstance(arg, tuple) else batch.column(arg)

    def update(self, batch):
        if self.filter is not None:
            batch = filter_batch(batch, self.filter)
        if self.accs is None:
            self._init_accs(batch)
        n = batch.num_rows
        keys = [batch.column(i) for i in self.group_cols]
        gid = np.zeros(n, dtype=np.int64)
        keep = []
        self.n_groups = lib().orc_groups_intern(self.groups, _cols(keys, keep), len(keys), C.c_int64(n), _p(gid))
        self.batches.append(batch.select(self.group_cols))
        col = len(self.group_cols)
        for (fn, arg, _), accs in zip(self.aggs, self.accs):
            if self.mode == "partial":
                a = _cast_for(fn, self._arg(batch, arg)) if arg is not None else None
                if fn == "count":
                    accs[0].update(a, gid, self.n_groups)
                elif fn == "sum":
                    accs[0].update(a, gid, self.n_groups)
                else:
                    accs[0].update(a, gid, self.n_groups)
                    accs[1].update(a, gid, self.n_groups)
            else:  # merge_batch: partial counts and sums are summed
                for acc in accs:
                    acc.update(batch.column(col), gid, self.n_groups)
                    col += 1

    def _keys(self):
        first = np.zeros(self.n_groups, dtype=np.int64)
        lib().orc_groups_first_rows(self.groups, _p(first))
        tbl = pa.Table.from_batches(self.batches).combine_chunks()
        return [take(tbl.column(i), first.astype(np.uint64)) for i in range(tbl.num_columns)], tbl.schema.names

    def emit(self):
        for accs in self.accs or []:
            for a in accs:
                a.grow(self.n_groups)
        keys, names = self._keys()
        cols, out_names = list(keys), list(names)
        for (fn, _arg, name), accs, t in zip(self.aggs, self.accs, self.in_types):
            if self.mode == "partial":
                if fn == "count":
                    cols.append(accs[0].array(pa.int64())); out_names.append(f"{name}[count]")
                elif fn == "sum":
                    cols.append(accs[0].array(accs[0].t)); out_names.append(f"{name}[sum]")
                else:
                    cols.append(accs[0].array(pa.uint64())); out_names.append(f"{name}[count]")
                    cols.append(accs[1].array(accs[1].t)); out_names.append(f"{name}[sum]")
            else:
                if fn == "count":
                    cols.append(accs[0].array(pa.int64(), nullable_by_seen=False))
                elif fn == "sum":
                    cols.append(accs[0].array(accs[0].t))
                else:
                    cols.append(self._avg(accs, t))
                out_names.append(name)
        return pa.RecordBatch.from_arrays(cols, names=out_names)

    def _avg(self, accs, t):
        cnt, s = accs
        st, rt = avg_types(t)
        n = self.n_groups
        validity, nulls = None, 0
        if not s.seen.all():
            validity = pa.py_buffer(np.packbits(s.seen, bitorder="little").tobytes())
            nulls = int(n - s.seen.sum())
        if pa.types.is_decimal128(rt):
            out = np.zeros((n, 2), dtype=np.uint64)
            c = np.ascontiguousarray(cnt.v.view(np.uint64))
            ovf = lib().orc_avg_decimal_evaluate(_p(s.v), _p(c), C.c_int64(n), C.c_int32(rt.scale - st.scale), _p(out))
            assert ovf == 0, "decimal avg overflow (reference: arithmetic error)"
            return pa.Array.from_buffers(rt, n, [validity, pa.py_buffer(out.tobytes())], nulls)
        with np.errstate(divide="ignore", invalid="ignore"):
            out = s.v / cnt.v.view(np.uint64).astype(np.float64)
        return pa.Array.from_buffers(rt, n, [validity, pa.py_buffer(out.tobytes())], nulls)


# ---- synthetic lineitem + the timed CPU pipeline (oracle/pipeline.cpp) -----------------------------------------------
Q1_DATE = 10471  # 1998-09-02 = date '1998-12-01' - interval '90' day


def gen_lineitem_numpy(rows, seed=42, row0=0, threads=None):
    """Deterministic synthetic lineitem columns (pure function of seed and row number) as numpy arrays."""
    from concurrent.futures import ThreadPoolExecutor

    L = lib()
    c = {k: np.zeros((rows, 2), dtype=np.uint64) for k in ("l_quantity", "l_extendedprice", "l_discount", "l_tax")}
    c["rf"] = np.zeros(rows, dtype=np.uint8)
    c["ls"] = np.zeros(rows, dtype=np.uint8)
    c["l_shipdate"] = np.zeros(rows, dtype=np.int32)
    threads = threads or os.cpu_count() or 1
    step = max(1, (rows + threads - 1) // threads)

    def work(lo):
        n = min(step, rows - lo)
        L.orc_gen_lineitem(C.c_int64(row0 + lo), C.c_int64(n), C.c_uint64(seed),
                           *[C.c_void_p(c[k][lo:].ctypes.data) for k in ("l_quantity", "l_extendedprice", "l_discount", "l_tax")],
                           C.c_void_p(c["rf"][lo:].ctypes.data), C.c_void_p(c["ls"][lo:].ctypes.data),
                           C.c_void_p(c["l_shipdate"][lo:].ctypes.data))
    with ThreadPoolExecutor(threads) as ex:
        list(ex.map(work, range(0, rows, step)))
    return c


def lineitem_record_batch(c):
    """numpy columns of gen_lineitem_numpy -> pyarrow RecordBatch with the reference's lineitem schema
    (/root/reference/benchmarks/src/bin/tpch.rs:923-940), zero-copy."""
    n = len(c["rf"])
    d = pa.decimal128(15, 2)
    offs = np.arange(n + 1, dtype=np.int32)

    def fixed(a, t):
        return pa.Array.from_buffers(t, n, [None, pa.py_buffer(a)])

    def utf8(a):
        return pa.Array.from_buffers(pa.string(), n, [None, pa.py_buffer(offs), pa.py_buffer(a)])
    return pa.RecordBatch.from_arrays(
        [fixed(c["l_quantity"], d), fixed(c["l_extendedprice"], d), fixed(c["l_discount"], d), fixed(c["l_tax"], d),
         utf8(c["rf"]), utf8(c["ls"]), fixed(c["l_shipdate"], pa.date32())],
        names=["l_quantity", "l_extendedprice", "l_discount", "l_tax", "l_returnflag", "l_linestatus", "l_shipdate"])


def q1_pipeline(rows, seed=42, threads=None, partitions=16, cols=None):
    """Run the timed CPU restatement of the Q1 hot path over `rows` synthetic rows. Returns seconds (of the pipeline
    only; generation is outside), the number of final groups and the final sums/counts per group."""
    threads = threads or os.cpu_count() or 1
    c = cols if cols is not None else gen_lineitem_numpy(rows, seed, threads=threads)
    L = lib()
    sec, ng = C.c_double(0), C.c_int32(0)
    mg = 64
    rf, ls = np.zeros(mg, dtype=np.uint8), np.zeros(mg, dtype=np.uint8)
    sums = np.zeros((mg, 5, 2), dtype=np.uint64)
    cnt = np.zeros(mg, dtype=np.int64)
    part = np.zeros(mg, dtype=np.uint32)
    rc = L.orc_q1_pipeline(*[C.c_void_p(c[k].ctypes.data) for k in ("l_quantity", "l_extendedprice", "l_discount", "l_tax",
                                                                    "rf", "ls", "l_shipdate")],
                           C.c_int64(rows), C.c_int32(Q1_DATE), C.c_int32(threads), C.c_uint32(partitions), C.byref(sec),
                           C.byref(ng), C.c_int32(mg), _p(rf), _p(ls), _p(sums), _p(cnt), _p(part))
    assert rc == 0, f"orc_q1_pipeline rc={rc}"
    n = ng.value

    def i128(x):
        v = int(x[0]) | (int(x[1]) << 64)
        return v - (1 << 128) if v >> 127 else v
    groups = {(chr(rf[i]), chr(ls[i])): {"sums": [i128(sums[i, a]) for a in range(5)], "count": int(cnt[i]),
                                         "partition": int(part[i])} for i in range(n)}
    return {"seconds": sec.value, "groups": n, "result": groups}
