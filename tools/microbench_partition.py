"""Repartition microbench (BASELINE config 3 shape: Int64 key + 4 x 8-byte payload), device-resident.
usage: python tools/microbench_partition.py [rows] [n_partitions ...]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pyarrow as pa
import torch

import ballista_b200 as bb

rows = int(float(sys.argv[1])) if len(sys.argv) > 1 else 100_000_000
parts = [int(x) for x in sys.argv[2:]] or [8, 64, 512]
stream = torch.cuda.Stream()
ctx = bb.Context(0, stream=stream.cuda_stream)
g = torch.Generator(device="cuda").manual_seed(42)
key = torch.randint(0, 2**63 - 1, (rows,), device="cuda", generator=g)
pay = [torch.randint(0, 2**40, (rows,), device="cuda", generator=g) for _ in range(4)]
schema = pa.schema([("k", pa.int64())] + [(f"p{i}", pa.int64()) for i in range(4)])
cols = [dict(values=t.data_ptr()) for t in [key] + pay]
d = bb.DeviceBatch.wrap_device(ctx, schema, cols, rows, (key, pay))
torch.cuda.synchronize()
for n in parts:
    part = bb.BatchPartitioner.try_new(bb.Partitioning.Hash([0], n), ctx)
    for _ in range(2):
        out, offs = part.partition_device(d)
        del out
    ts = []
    for _ in range(5):
        with torch.cuda.stream(stream):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            out, offs = part.partition_device(d)
            e1.record(stream)
        e1.synchronize()
        ts.append(e0.elapsed_time(e1))
        del out
    ms = min(ts)
    print(f"n={n:4d} rows={rows} best {ms:.3f} ms  {rows / ms / 1e6:.2f} Grows/s  "
          f"{rows * 80 / ms / 1e6:.1f} GB/s algorithmic (80 B/row)  all={['%.2f' % t for t in ts]}", flush=True)
