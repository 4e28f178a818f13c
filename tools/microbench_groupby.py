"""Group-by microbench, device-resident input (BASELINE configs 2 and 4b shapes).

  db-benchmark G1 (h2oai groupby-datagen shape; queries of
  /root/reference/benchmarks/db-benchmark/groupby-datafusion.py:73-141): q1..q5 over N rows, K = 100
  high cardinality: SELECT l_orderkey, sum(l_quantity), sum(l_extendedprice), count(*) GROUP BY l_orderkey

usage: python tools/microbench_groupby.py [rows=1e7] [K=100] [orderkey_rows=3e7]
Prints, per query, the Partial aggregate time (update + emit, CUDA events), rows/s and algorithmic GB/s
(SURVEY.md §8d bytes per row: q1 12, q2 16, q3 20, q4 32, q5 32; orderkey 40).
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import pyarrow as pa
import torch

import ballista_b200 as bb

N = int(float(sys.argv[1])) if len(sys.argv) > 1 else 10_000_000
K = int(float(sys.argv[2])) if len(sys.argv) > 2 else 100
N_OK = int(float(sys.argv[3])) if len(sys.argv) > 3 else 30_000_000
stream = torch.cuda.Stream()
torch.cuda.set_stream(stream)
ctx = bb.Context(0, stream=stream.cuda_stream)


def timed(make_agg, batch, reps=5):
    best, path, groups = None, "", 0
    for i in range(reps + 2):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        agg = make_agg()
        e0.record(stream)
        agg.update(batch)
        out = agg.emit()
        e1.record(stream)
        e1.synchronize()
        ms = e0.elapsed_time(e1)
        path, groups = agg.last_path, out.num_rows
        if i >= 2:
            best = ms if best is None else min(best, ms)
        del out, agg
    return best, path, groups


def g1_device(n, k, seed=42):
    """The G1 table generated on the device (torch) and wrapped without copies; dictionaries are uploaded."""
    g = torch.Generator(device="cuda").manual_seed(seed)
    nk = max(1, n // k)
    keep, cols, fields = [], [], []

    def dict_col(name, card, fmt):
        idx = torch.randint(0, card, (n,), device="cuda", generator=g, dtype=torch.int32)
        d = pa.array([fmt % i for i in range(1, card + 1)], type=pa.string())
        offs = torch.from_numpy(np.frombuffer(d.buffers()[1], dtype=np.int32)[: card + 1].copy()).cuda()
        data = torch.from_numpy(np.frombuffer(d.buffers()[2], dtype=np.uint8).copy()).cuda()
        keep.extend([idx, offs, data])
        cols.append(dict(values=idx.data_ptr(), dictionary=dict(values=data.data_ptr(), offsets=offs.data_ptr(), length=card)))
        fields.append((name, pa.dictionary(pa.int32(), pa.string())))

    def int_col(name, lo, hi):
        t = torch.randint(lo, hi + 1, (n,), device="cuda", generator=g)
        keep.append(t)
        cols.append(dict(values=t.data_ptr()))
        fields.append((name, pa.int64()))
    dict_col("id1", k, "id%03d")
    dict_col("id2", k, "id%03d")
    dict_col("id3", nk, "id%010d")
    int_col("id4", 1, k)
    int_col("id5", 1, k)
    int_col("id6", 1, nk)
    int_col("v1", 1, 5)
    int_col("v2", 1, 15)
    v3 = torch.round(torch.rand(n, device="cuda", generator=g, dtype=torch.float64) * 100 * 1e6) / 1e6
    keep.append(v3)
    cols.append(dict(values=v3.data_ptr()))
    fields.append(("v3", pa.float64()))
    schema = pa.schema(fields)
    return bb.DeviceBatch.wrap_device(ctx, schema, cols, n, keepalive=keep), schema


dev, g1_schema = g1_device(N, K)
QUERIES = [
    ("q1 sum(v1) by id1", [0], [("sum", 6, "v1")], 12),
    ("q2 sum(v1) by id1,id2", [0, 1], [("sum", 6, "v1")], 16),
    ("q3 sum(v1),avg(v3) by id3", [2], [("sum", 6, "v1"), ("avg", 8, "v3")], 20),
    ("q4 avg(v1..v3) by id4", [3], [("avg", 6, "v1"), ("avg", 7, "v2"), ("avg", 8, "v3")], 32),
    ("q5 sum(v1..v3) by id6", [5], [("sum", 6, "v1"), ("sum", 7, "v2"), ("sum", 8, "v3")], 32),
]
print(f"db-benchmark G1 N={N} K={K}")
for name, gb, aggs, bpr in QUERIES:
    ms, path, groups = timed(lambda: bb.GpuAggregateExec(ctx, bb.AggregateMode.Partial, gb, aggs, g1_schema), dev)
    print(f"  {name:28s} {ms:8.3f} ms  {N / ms / 1e6:7.2f} Grows/s  {N * bpr / ms / 1e6:8.1f} GB/s ({bpr} B/row)  "
          f"path={path} groups={groups}", flush=True)
del dev

# ---- high cardinality: l_orderkey (1..7 lines per order, sparse keys as dbgen) ---------------------------------
g = torch.Generator(device="cuda").manual_seed(42)
lines = torch.randint(1, 8, (N_OK // 3 + 8,), device="cuda", generator=g)
order = torch.repeat_interleave(torch.arange(lines.numel(), device="cuda"), lines)[:N_OK]
okey = (order // 8) * 32 + order % 8 + 1          # dbgen's sparse order keys
perm = torch.randperm(N_OK, device="cuda", generator=g)
if os.environ.get("OK_SORTED"):   # dbgen's own order (clustered): the best case
    pass
elif os.environ.get("OK_BUCKETED"):  # rows grouped into B hash buckets, unsorted inside
    nb = int(os.environ["OK_BUCKETED"])
    okey = okey[perm]
    okey = okey[torch.argsort((okey * 0x9E3779B97F4A7C15) >> 40 & (nb - 1), stable=True)].contiguous()
else:
    okey = okey[perm].contiguous()                    # unsorted input = worst case for the partial aggregate
n_groups = int(torch.unique(okey).numel())


def dec(lo, hi):
    t = torch.zeros((N_OK, 2), dtype=torch.int64, device="cuda")
    t[:, 0] = torch.randint(lo, hi + 1, (N_OK,), generator=g, device="cuda")
    return t


qty, price = dec(100, 5000), dec(90000, 10495000)
d15 = pa.decimal128(15, 2)
schema = pa.schema([("l_orderkey", pa.int64()), ("l_quantity", d15), ("l_extendedprice", d15)])
b = bb.DeviceBatch.wrap_device(ctx, schema, [dict(values=okey.data_ptr()), dict(values=qty.data_ptr()),
                                             dict(values=price.data_ptr())], N_OK, keepalive=(okey, qty, price))
aggs = [("sum", 1, "q"), ("sum", 2, "p"), ("count", None, "c")]
ms, path, groups = timed(lambda: bb.GpuAggregateExec(ctx, bb.AggregateMode.Partial, [0], aggs, schema), b, reps=3)
assert groups == n_groups, (groups, n_groups)
algo = N_OK * 40 + groups * 48
print(f"orderkey N={N_OK} groups={groups}: {ms:.3f} ms  {N_OK / ms / 1e6:.2f} Grows/s  {algo / ms / 1e6:.1f} GB/s "
      f"(40 B/row in + 48 B/group out)  path={path}", flush=True)
