"""Where the step's tail goes (everything of bench.py's step that is not the fused kernel): host wall time per phase with a
stream synchronise after each, on a small resident shard (the tail does not depend on the shard size)."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import ballista_b200 as bb  # noqa: E402
import bench  # noqa: E402

n = int(os.environ.get("ROWS", 5_000_000))
dev = torch.device("cuda", 0)
stream = torch.cuda.Stream(device=dev)
torch.cuda.set_stream(stream)
ctx = bb.Context(0, stream=stream.cuda_stream)
cols = bench.gen_lineitem_torch(n, 42, dev)
batch = bench.wrap_lineitem(bb, ctx, cols, n)
hp = bench.HotPath(bb, ctx)
for _ in range(3):
    hp.step([batch]).download()
T = {}


def mark(name, t0):
    torch.cuda.synchronize()
    T[name] = T.get(name, 0.0) + (time.perf_counter() - t0) * 1e6
    return time.perf_counter()


K = 20
for _ in range(K):
    torch.cuda.synchronize()
    t = time.perf_counter()
    agg = hp.partial_op
    agg.reset()
    t = mark("partial reset", t)
    agg.update(batch)
    t = mark("partial update", t)
    state = agg.emit()
    t = mark("partial emit", t)
    part, offs = hp.partitioner.partition_device(state)
    t = mark("partition", t)
    fin = hp.final_op
    fin.reset()
    t = mark("final reset", t)
    fin.update(part)
    t = mark("final update", t)
    out = fin.emit()
    t = mark("final emit", t)
    r = out.download()
    t = mark("download", t)
    del state, part, out
    t = mark("destroy", t)
l0 = ctx.kernel_launches
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(K):
    hp.step([batch]).download()
torch.cuda.synchronize()
whole = (time.perf_counter() - t0) / K * 1e6
print(f"rows {n}; phases (us per step, each followed by a synchronise):")
for k, v in T.items():
    print(f"  {k:16s} {v / K:8.1f}")
print(f"  sum {sum(T.values()) / K:8.1f}; whole step without the extra synchronises {whole:8.1f} us; launches per step {(ctx.kernel_launches - l0) / K:.1f}")
