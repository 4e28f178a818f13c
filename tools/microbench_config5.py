"""BASELINE config 5: db-benchmark group-by on 1e9 rows with Zipf-skewed keys, k = 1e2 and k = 1e7, across the GPUs of
one node (rows sharded evenly; Partial aggregate -> fused repartition + exchange over NVLink -> FinalPartitioned).

  q4  avg(v1), avg(v2), avg(v3) by id4     id4 ~ Zipf over 1..K            (k = 1e2: 100 groups)
  q5  sum(v1), sum(v2), sum(v3) by id6     id6 ~ Zipf over 1..N/K          (k = 1e7 domain)
(/root/reference/benchmarks/db-benchmark/groupby-datafusion.py:124, 141; 32 B/row each) Zipf(s = 1) is
drawn as floor((D + 1)^u), u ~ U(0,1): P(k) = ln((k+1)/k) / ln(D+1) ~ 1/k. ZIPF=0 gives uniform keys for comparison.

Launch: python -m torch.distributed.run --nnodes=1 --nproc-per-node G --master-addr 127.0.0.1 --master-port 29531 \
            tools/microbench_config5.py [total_rows=1e9] [K=100]
Phases: CUDA events on the ctx stream, max over ranks, best of 3 after 2 warm-ups.
"""
import json
import math
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pyarrow as pa
import torch
import torch.distributed as dist

import ballista_b200 as bb

TOTAL = int(float(sys.argv[1])) if len(sys.argv) > 1 else 1_000_000_000
K = int(float(sys.argv[2])) if len(sys.argv) > 2 else 100
ZIPF = bool(int(os.environ.get("ZIPF", "1")))
NP = 16
world = int(os.environ.get("WORLD_SIZE", "1"))
rank = int(os.environ.get("RANK", "0"))
local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
stream = torch.cuda.Stream(device=dev)
torch.cuda.set_stream(stream)
ctx = bb.Context(local, stream=stream.cuda_stream)
comm = bb.Exchange(ctx, rank, world)
N = TOTAL // world
g = torch.Generator(device=dev).manual_seed(42 + rank)


def keys(domain):
    if not ZIPF:
        return torch.randint(1, domain + 1, (N,), device=dev, generator=g)
    u = torch.rand(N, device=dev, generator=g, dtype=torch.float64)
    k = torch.exp(u * math.log(domain + 1.0)).floor_().clamp_(1, domain).to(torch.int64)
    return k


id4, id6 = keys(K), keys(max(1, TOTAL // K))
v1 = torch.randint(1, 6, (N,), device=dev, generator=g)
v2 = torch.randint(1, 16, (N,), device=dev, generator=g)
v3 = torch.round(torch.rand(N, device=dev, generator=g, dtype=torch.float64) * 1e8) / 1e6
schema = pa.schema([("id4", pa.int64()), ("id6", pa.int64()), ("v1", pa.int64()), ("v2", pa.int64()), ("v3", pa.float64())])
cols = (id4, id6, v1, v2, v3)
batch = bb.DeviceBatch.wrap_device(ctx, schema, [dict(values=t.data_ptr()) for t in cols], N, keepalive=cols)
top = float((id6 == 1).sum().item()) / N
QUERIES = [("q4 avg(v1..v3) by id4", [0], [("avg", 2, "v1"), ("avg", 3, "v2"), ("avg", 4, "v3")]),
           ("q5 sum(v1..v3) by id6", [1], [("sum", 2, "v1"), ("sum", 3, "v2"), ("sum", 4, "v3")])]


def ev():
    return torch.cuda.Event(enable_timing=True)


def step(gb, aggs):
    e = [ev() for _ in range(4)]
    e[0].record(stream)
    agg = bb.GpuAggregateExec(ctx, bb.AggregateMode.Partial, gb, aggs, schema)
    agg.update(batch)
    state = agg.emit()
    e[1].record(stream)
    part = comm.partition_exchange(state, [0], NP, window_bytes=1 << 30)
    sent = comm.last_sent_bytes or 0
    e[2].record(stream)
    fin = bb.GpuAggregateExec(ctx, bb.AggregateMode.FinalPartitioned, gb, aggs, schema)
    fin.update(part)
    out = fin.emit()
    e[3].record(stream)
    e[3].synchronize()
    return [e[i].elapsed_time(e[i + 1]) for i in range(3)], state.num_rows, sent, out.num_rows, agg.last_path, comm.last_path


for name, gb, aggs in QUERIES:
    best = None
    for i in range(5):
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        ms, pgroups, sent, fgroups, path, xpath = step(gb, aggs)
        t = torch.tensor(ms + [sum(ms)], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        if i >= 2 and (best is None or t[3].item() < best[3]):
            best = t.tolist()
    cnt = torch.tensor([pgroups, sent, fgroups], device=dev, dtype=torch.int64)
    if world > 1:
        dist.all_reduce(cnt, op=dist.ReduceOp.SUM)
    pg, sb, fg = [int(x) for x in cnt.tolist()]
    if rank == 0:
        print(json.dumps({"config": "5 db-benchmark group-by, " + ("Zipf(1) keys" if ZIPF else "uniform keys"), "query": name,
                          "n_gpus": world, "rows_total": N * world, "partial_groups_total": pg, "final_groups_total": fg,
                          "agg_path": path, "exchange_path": xpath, "hottest_id6_share": top,
                          "ms": {"partial": best[0], "repartition+exchange": best[1], "final": best[2], "step": best[3]},
                          "rows_per_s": N * world / (best[3] * 1e-3), "algorithmic_GBps_per_gpu": N * 32 / (best[3] * 1e-3) / 1e9,
                          "nvlink_bytes_per_gpu": sb / world}), flush=True)
comm.close()
if world > 1:
    dist.destroy_process_group()
