"""ORACLE (test infrastructure only — never imported by the product path).

ctypes + pyarrow glue over oracle/_build/liboracle.so, the CPU restatement of the reference's hot path
(DataFusion v34 create_hashes / BatchPartitioner / take / GroupedHashAggregateStream; SURVEY.md
Appendix A; reference call sites /root/reference/ballista/core/src/execution_plans/shuffle_writer.rs:336-346
and /root/reference/ballista/scheduler/src/planner.rs:376-427).  PARITY UNPINNED — see oracle.h.

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

    def __init__(self, mode, group_cols, aggs, arg_types=None):
        assert mode in ("partial", "final")
        self.mode, self.group_cols, self.aggs, self.arg_types = mode, list(group_cols), list(aggs), arg_types
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
                t = batch.column(arg).type if arg is not None else None
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

    def update(self, batch):
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
                a = _cast_for(fn, batch.column(arg)) if arg is not None else None
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
