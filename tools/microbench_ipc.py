"""Microbench of the kernels either side of the exchange (device-resident batches, CUDA events / wall clock where a
host copy is part of the operation): K7 IPC framing of partitions, K8 IPC stream reader (uncompressed and LZ4_FRAME as
Arrow's codec writes it), K9 concat, and the fused repartition + exchange kernel at world = 1 (window = local memory:
the kernel's own cost without the link).

usage: python tools/microbench_ipc.py [rows=2e7]
"""
import io
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import pyarrow as pa
import torch

import ballista_b200 as bb
from ballista_b200 import ipc

N = int(float(sys.argv[1])) if len(sys.argv) > 1 else 20_000_000
stream = torch.cuda.Stream()
torch.cuda.set_stream(stream)
ctx = bb.Context(0, stream=stream.cuda_stream)
rng = np.random.default_rng(1)


def timed(fn, reps=5, warm=2):
    best = None
    for i in range(reps + warm):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        out = fn()
        ctx.synchronize()
        dt = time.perf_counter() - t0
        if i >= warm:
            best = dt if best is None else min(best, dt)
        del out
    return best


# an aggregate-state shaped batch: Int64 key, 2 x Decimal128, Int64 count = 48 B/row
key = torch.randint(0, 1 << 40, (N,), device="cuda")
s1 = torch.zeros((N, 2), dtype=torch.int64, device="cuda")
s1[:, 0] = torch.randint(0, 10**9, (N,), device="cuda")
s2 = s1.clone()
cnt = torch.randint(1, 8, (N,), device="cuda")
d = pa.decimal128(25, 2)
schema = pa.schema([("k", pa.int64()), ("s1", d), ("s2", d), ("c", pa.int64())])
state = bb.DeviceBatch.wrap_device(ctx, schema, [dict(values=t.data_ptr()) for t in (key, s1, s2, cnt)], N, keepalive=(key, s1, s2, cnt))
ROW = 48
print(f"state batch: {N} rows x {ROW} B = {N * ROW / 1e9:.2f} GB")

# fused repartition + exchange, world = 1
ex = bb.Exchange(ctx, 0, 1)
t = timed(lambda: ex.partition_exchange(state, [0], 16, window_bytes=N * ROW + (1 << 20)))
print(f"bgpu_partition_exchange world=1 n=16: {t * 1e3:.3f} ms  {N / t / 1e9:.2f} G rows/s  {2 * N * ROW / t / 1e9:.0f} GB/s (read + write; + copy-out)")
part_t = timed(lambda: bb.BatchPartitioner.try_new(bb.Partitioning.Hash([0], 16), ctx).partition_device(state))
print(f"bgpu_hash_partition n=16:              {part_t * 1e3:.3f} ms  {2 * N * ROW / part_t / 1e9:.0f} GB/s")

# K7: IPC framing of 16 partitions (includes the device-to-host copy of all messages)
part, offs = bb.BatchPartitioner.try_new(bb.Partitioning.Hash([0], 16), ctx).partition_device(state)
m = min(N, 4_000_000)
small = bb.DeviceBatch.wrap_device(ctx, schema, [dict(values=t.data_ptr()) for t in (key, s1, s2, cnt)], m, keepalive=(key, s1, s2, cnt))
sp, so = bb.BatchPartitioner.try_new(bb.Partitioning.Hash([0], 16), ctx).partition_device(small)
t = timed(lambda: ipc.encode_partitions(ctx, sp, so))
print(f"bgpu_ipc_encode_partitions {m} rows, 16 partitions: {t * 1e3:.3f} ms  {m * ROW / t / 1e9:.1f} GB/s (device layout + one D2H copy into pinned memory)")

t = timed(lambda: ipc.encode_partitions(ctx, sp, so, compression="lz4"))
enc = ipc.encode_partitions(ctx, sp, so, compression="lz4")
print(f"bgpu_ipc_encode_partitions lz4, same batch: {t * 1e3:.3f} ms  {m * ROW / t / 1e9:.1f} GB/s of input, {enc.total_bytes / 1e6:.0f} MB out of {m * ROW / 1e6:.0f} MB")
del enc

# K8: reader. The stream is what Arrow's writer produces from the host copy of `small`.
host = small.download()
for comp in (None, "lz4"):
    sink = io.BytesIO()
    with pa.ipc.new_stream(sink, host.schema, options=pa.ipc.IpcWriteOptions(compression=comp)) as w:
        for lo in range(0, m, 1 << 20):
            w.write_batch(host.slice(lo, 1 << 20))
    data = sink.getvalue()

    def read_all():
        return list(ipc.StreamReader(ctx, data))
    t = timed(read_all, reps=3)
    t_arrow = min(timeit for timeit in [_t() for _t in [lambda: (lambda t0: (pa.ipc.open_stream(io.BytesIO(data)).read_all(), time.perf_counter() - t0)[1])(time.perf_counter())] * 3])
    print(f"ipc.StreamReader {comp or 'uncompressed'}: file {len(data) / 1e6:.0f} MB -> {m * ROW / 1e6:.0f} MB on the device in {t * 1e3:.2f} ms = "
          f"{m * ROW / t / 1e9:.2f} GB/s decoded (pyarrow reader on the host: {t_arrow * 1e3:.2f} ms)")

# the reference's own framing: block-independent 64 KB LZ4 blocks (lz4_flex), one decode chain per block
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from helpers import reference_style_lz4_stream

data = reference_style_lz4_stream([host.slice(lo, 1 << 20) for lo in range(0, m, 1 << 20)])
t = timed(lambda: list(ipc.StreamReader(ctx, data)), reps=3)
print(f"ipc.StreamReader lz4, reference framing (independent 64 KB blocks): file {len(data) / 1e6:.0f} MB -> {m * ROW / 1e6:.0f} MB in "
      f"{t * 1e3:.2f} ms = {m * ROW / t / 1e9:.2f} GB/s decoded")

# low-entropy column set: what LZ4 is good at
rep = pa.RecordBatch.from_arrays([pa.array(np.repeat(np.arange(m // 1000, dtype=np.int64), 1000)), pa.array(np.zeros(m, dtype=np.int64)),
                                  pa.array(rng.integers(0, 4, m).astype(np.int64))], names=["run", "zero", "small"])
sink = io.BytesIO()
with pa.ipc.new_stream(sink, rep.schema, options=pa.ipc.IpcWriteOptions(compression="lz4")) as w:
    w.write_batch(rep)
data = reference_style_lz4_stream([rep])
t = timed(lambda: list(ipc.StreamReader(ctx, data)), reps=3)
print(f"ipc.StreamReader lz4, reference framing, compressible: file {len(data) / 1e6:.1f} MB -> {rep.nbytes / 1e6:.0f} MB in {t * 1e3:.2f} ms = {rep.nbytes / t / 1e9:.2f} GB/s decoded")

# K9: concat of 64 pieces
pieces = [bb.DeviceBatch.upload(ctx, host.slice(lo, m // 64)) for lo in range(0, m - m // 64 + 1, m // 64)]
rows = sum(p.num_rows for p in pieces)
t = timed(lambda: bb.DeviceBatch.concat(ctx, pieces))
print(f"bgpu_batch_concat {len(pieces)} x {m // 64} rows: {t * 1e3:.3f} ms  {2 * rows * ROW / t / 1e9:.0f} GB/s (read + write)")
