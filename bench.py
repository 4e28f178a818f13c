"""bench.py — TPC-H Q1 aggregation + shuffle hot path on B200 (BASELINE.json metric: input rows/s and HBM GB/s).

One "step" = one pass of the hot path over one synthetic `lineitem` shard per GPU:
    Partial GroupedHashAggregate (Q1 filter + decimal expressions fused)  ->  hash repartition of the state batch
    (create_hashes, seeds 0,0,0,0, hash % n)  ->  [N>1: all-to-all over NVLink]  ->  FinalPartitioned aggregate.
`value` times it with the shard resident in HBM; `e2e` runs the same path from pinned HOST buffers through the C ABI
(upload inside the timed region, result downloaded). `--impl reference` times the CPU restatement of the reference's
path (oracle/pipeline.cpp; the Rust reference cannot be built here) on the host cores.

usage: python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--rows R]
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

SF100_ROWS = 600_037_902  # lineitem rows at SF100
Q1_BYTES_PER_ROW = 78     # 4 x Decimal128 + 2 x (4 B offset + 1 B) Utf8 + Date32 filter column (fused filter)
N_PARTITIONS = 16         # ballista.shuffle.partitions default (/root/reference/ballista/core/src/config.rs:177-179)
SHIP_LO, SHIP_HI, CUTOFF, Q1_DATE = 8036, 10561, 9298, 10471  # 1992-01-02, 1998-12-01, 1995-06-17, 1998-09-02


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--rows", type=int, default=int(os.environ.get("BENCH_ROWS", SF100_ROWS)), help="lineitem rows per GPU")
    ap.add_argument("--e2e-rows", type=int, default=int(os.environ.get("BENCH_E2E_ROWS", 60_000_000)),
                    help="rows of the shard kept in pinned host memory for the end-to-end leg")
    ap.add_argument("--e2e-batch-rows", type=int, default=8192, help="rows per host RecordBatch of the end-to-end leg")
    ap.add_argument("--e2e-threads", type=int, default=0, help="feeder threads of the end-to-end leg (0 = up to 16)")
    ap.add_argument("--cpu-rows", type=int, default=int(os.environ.get("BENCH_CPU_ROWS", 60_000_000)),
                    help="bounded sample for the CPU baseline")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-configs", action="store_true", help="skip BASELINE configs 2, 3 (N = 1) and 5 (N > 1)")
    ap.add_argument("--cfg3-rows", type=int, default=int(os.environ.get("BENCH_CFG3_ROWS", 1_000_000_000)))
    ap.add_argument("--cfg5-rows", type=int, default=int(os.environ.get("BENCH_CFG5_ROWS", 1_000_000_000)))
    ap.add_argument("--no-orderkey", action="store_true", help="skip the GROUP BY l_orderkey (exchange-bound) leg")
    ap.add_argument("--orderkey-rows", type=int, default=int(os.environ.get("BENCH_ORDERKEY_ROWS", SF100_ROWS)),
                    help="total rows of the GROUP BY l_orderkey leg (sharded over the ranks)")
    return ap.parse_args()


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def load_traffic(rows):
    """dram__bytes_read + dram__bytes_write of one bgpu_fused launch from the committed ncu capture (profiles/), or None
    when this run's row count is not the one that was captured."""
    p = os.path.join(ROOT, "profiles", "r2_fused_traffic.json")
    if not os.path.exists(p):
        return None
    t = json.load(open(p))
    return t["dram_bytes_read"] + t["dram_bytes_write"] if t.get("rows") == rows else None


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.stop, self.t = index, [], threading.Event(), None

    def _run(self):
        while not self.stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            self.stop.wait(0.1)

    def __enter__(self):
        self.t = threading.Thread(target=self._run, daemon=True)
        self.t.start()
        return self

    def __exit__(self, *a):
        self.stop.set()
        self.t.join(timeout=6)

    def summary(self):
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unsampled"]}
        sm = sorted(int(float(r[0])) for r in self.rows)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[3 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": int(float(self.rows[0][1])), "reasons": reasons,
                "samples": len(sm), "power_w_max": max(float(r[2]) for r in self.rows)}


# ---- synthetic lineitem (schema of /root/reference/benchmarks/src/bin/tpch.rs:923-940) -------------
def gen_lineitem_torch(n, seed, device):
    """Columns as torch tensors: Decimal128 as int64 [n, 2] (lo, hi), Utf8 as int32 offsets + uint8 bytes, Date32."""
    import torch

    g = torch.Generator(device=device).manual_seed(seed)

    def dec(lo, hi):
        t = torch.zeros((n, 2), dtype=torch.int64, device=device)
        t[:, 0] = torch.randint(lo, hi + 1, (n,), generator=g, device=device)
        return t
    qty = dec(1, 50)
    qty[:, 0] *= 100
    price, disc, tax = dec(90000, 10495000), dec(0, 10), dec(0, 8)
    ship = torch.randint(SHIP_LO, SHIP_HI + 1, (n,), generator=g, device=device, dtype=torch.int32)
    receipt = ship + torch.randint(1, 31, (n,), generator=g, device=device, dtype=torch.int32)
    coin = torch.rand(n, generator=g, device=device) < 0.5
    rf = torch.where(receipt <= CUTOFF, torch.where(coin, ord("R"), ord("A")), ord("N")).to(torch.uint8)
    ls = torch.where(ship > CUTOFF, ord("O"), ord("F")).to(torch.uint8)
    # one offsets buffer per string column, as Arrow arrays have them (a shared tensor would make the second column's
    # offset loads L2 hits and overstate the bytes the kernel moves)
    rf_offs = torch.arange(n + 1, dtype=torch.int32, device=device)
    ls_offs = torch.arange(n + 1, dtype=torch.int32, device=device)
    del receipt, coin
    return {"l_quantity": qty, "l_extendedprice": price, "l_discount": disc, "l_tax": tax, "rf_offsets": rf_offs, "rf_data": rf,
            "ls_offsets": ls_offs, "ls_data": ls, "l_shipdate": ship}


def lineitem_schema():
    import pyarrow as pa

    d = pa.decimal128(15, 2)
    return pa.schema([("l_quantity", d), ("l_extendedprice", d), ("l_discount", d), ("l_tax", d),
                      ("l_returnflag", pa.string()), ("l_linestatus", pa.string()), ("l_shipdate", pa.date32())])


def q1_plan():
    """Q1's Partial aggregate (/root/reference/benchmarks/queries/q1.sql) over lineitem_schema()."""
    from ballista_b200.expr import col, lit

    qty, price, disc, tax = col(0), col(1), col(2), col(3)
    disc_price = price * (lit(1) - disc)
    aggs = [("sum", qty, "sum_qty"), ("sum", price, "sum_base_price"), ("sum", disc_price, "sum_disc_price"),
            ("sum", disc_price * (lit(1) + tax), "sum_charge"), ("avg", qty, "avg_qty"), ("avg", price, "avg_price"),
            ("avg", disc, "avg_disc"), ("count", None, "count_order")]
    return [4, 5], aggs, col(6) <= lit(Q1_DATE, 3)


def wrap_lineitem(bb, ctx, cols, n):
    c = cols
    columns = [dict(values=c[k].data_ptr()) for k in ("l_quantity", "l_extendedprice", "l_discount", "l_tax")]
    columns += [dict(values=c["rf_data"].data_ptr(), offsets=c["rf_offsets"].data_ptr()),
                dict(values=c["ls_data"].data_ptr(), offsets=c["ls_offsets"].data_ptr()),
                dict(values=c["l_shipdate"].data_ptr())]
    return bb.DeviceBatch.wrap_device(ctx, lineitem_schema(), columns, n, keepalive=cols)


def host_lineitem_batches(cols_host, n, chunk):
    """pyarrow RecordBatches (zero-copy over pinned torch host tensors) of <= chunk rows."""
    import pyarrow as pa

    schema = lineitem_schema()
    out = []
    for lo in range(0, n, chunk):
        m = min(chunk, n - lo)

        def fixed(name, typ, width):
            t = cols_host[name]
            buf = pa.foreign_buffer(t.data_ptr() + lo * width, m * width, base=t)
            return pa.Array.from_buffers(typ, m, [None, buf])

        def utf8(prefix):
            o, d = cols_host[prefix + "_offsets"], cols_host[prefix + "_data"]
            ob = pa.foreign_buffer(o.data_ptr() + lo * 4, (m + 1) * 4, base=o)
            db = pa.foreign_buffer(d.data_ptr(), d.numel(), base=d)
            return pa.Array.from_buffers(pa.string(), m, [None, ob, db])
        d = pa.decimal128(15, 2)
        arrays = [fixed("l_quantity", d, 16), fixed("l_extendedprice", d, 16), fixed("l_discount", d, 16), fixed("l_tax", d, 16),
                  utf8("rf"), utf8("ls"), fixed("l_shipdate", pa.date32(), 4)]
        out.append(pa.RecordBatch.from_arrays(arrays, schema=schema))
    return out


def upload_bytes(batch):
    """Bytes bgpu_batch_upload copies for a host batch: fixed-width buffers whole, utf8 as (rows + 1) offsets plus the
    byte range the offsets reference (arrow_io.cu import_column)."""
    import numpy as np
    import pyarrow as pa

    total = 0
    for col in batch.columns:
        bufs = col.buffers()
        if pa.types.is_string(col.type) or pa.types.is_binary(col.type):
            offs = np.frombuffer(bufs[1], dtype=np.int32, count=len(col) + 1, offset=col.offset * 4)
            total += (len(col) + 1) * 4 + int(offs[-1] - offs[0]) + (bufs[0].size if bufs[0] is not None else 0)
        else:
            total += sum(b.size for b in bufs if b is not None)
    return total


# ---- the hot path through the host mirror of the reference operators ---------------------------------------------
class HotPath:
    """Partial aggregate -> hash repartition -> (exchange) -> FinalPartitioned aggregate on one GPU/rank."""

    def __init__(self, bb, ctx, rank=0, world=1, comm=None):
        self.bb, self.ctx, self.rank, self.world, self.comm = bb, ctx, rank, world, comm
        self.schema = lineitem_schema()
        self.group_by, self.aggs, self.filter = q1_plan()
        self.partitioner = bb.BatchPartitioner.try_new(bb.Partitioning.Hash(list(range(len(self.group_by))), N_PARTITIONS), ctx)
        self.last_path = ""
        # the operators are made once per task, as the reference's plan nodes are; every step is one `execute`: a new input
        # stream over the same plan (bgpu_agg_reset)
        self.partial_op = bb.GpuAggregateExec(ctx, bb.AggregateMode.Partial, self.group_by, self.aggs, self.schema, filter=self.filter)
        self.final_op = bb.GpuAggregateExec(ctx, bb.AggregateMode.FinalPartitioned, self.group_by, self.aggs, self.schema)

    def partial(self, device_batches):
        agg = self.partial_op
        agg.reset()
        for b in device_batches:
            agg.update(b)
        self.last_path = agg.last_path
        self.hot_kernel_ms = agg.hot_kernel_ms
        return agg.emit()

    def shuffle_and_final(self, state):
        bb = self.bb
        keys = list(range(len(self.group_by)))
        if self.world > 1:  # repartition + exchange: one all-gather for a state of a few rows, NVLink scatter / NCCL otherwise
            part = self.comm.exchange_state(state, keys, N_PARTITIONS)
            self.exchange_path = self.comm.last_path
        else:
            part, _ = self.partitioner.partition_device(state)
        # the partitions this rank owns hold disjoint keys: one merged FinalPartitioned table gives the union of the
        # per-partition results
        fin = self.final_op
        fin.reset()
        fin.update(part)
        return fin.emit()

    def step(self, device_batches):
        return self.shuffle_and_final(self.partial(device_batches))


_ORIGINAL_AFFINITY = None


def bind_to_gpu_numa_node(index):
    """Run this rank on the CPUs next to its GPU (NVML's affinity mask), so that the pinned host buffers of the end-to-end
    leg are first-touched on that NUMA node: with 8 ranks uploading at once, buffers that all sit on one socket share one
    memory controller and one inter-socket link. Best effort (a rank that cannot be bound just stays where it is)."""
    global _ORIGINAL_AFFINITY
    try:
        import pynvml

        _ORIGINAL_AFFINITY = os.sched_getaffinity(0)
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(index)
        n_cpu = os.cpu_count() or 1
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (n_cpu + 63) // 64)
        cpus = {64 * w + b for w, m in enumerate(words) for b in range(64) if (m >> b) & 1}
        cpus &= set(os.sched_getaffinity(0))
        if cpus:
            os.sched_setaffinity(0, cpus)
    except Exception:
        pass


# ---- verification: exact integer arithmetic in torch, independent of the library ------------------------------------------
SPLIT = 24  # sums that can pass 2^63 at SF100 (sum_charge) are taken as two int64 sums of the value's high and low bits


def _exact_sum(t):
    return (int((t >> SPLIT).sum()) << SPLIT) + int((t & ((1 << SPLIT) - 1)).sum())


Q1_GROUPS = [(rf, ls) for rf in "ARN" for ls in "FO"]


def q1_exact_parts(cols, m):
    """Per Q1 group over rows [0, m): int64 parts [count, sum_qty, sum_price, sum_disc, disc_price hi, lo, charge hi, lo]
    (unscaled integers; hi/lo = the sums of the values' bits above / below 2^SPLIT). Shape [6, 8], on the device."""
    import torch

    qty, price, disc, tax = (cols[k][:m, 0] for k in ("l_quantity", "l_extendedprice", "l_discount", "l_tax"))
    keep = cols["l_shipdate"][:m] <= Q1_DATE
    out = torch.zeros((len(Q1_GROUPS), 8), dtype=torch.int64, device=qty.device)
    lo_mask = (1 << SPLIT) - 1
    for gi, (rf, ls) in enumerate(Q1_GROUPS):
        sel = torch.nonzero(keep & (cols["rf_data"][:m] == ord(rf)) & (cols["ls_data"][:m] == ord(ls))).squeeze(1)
        if sel.numel() == 0:
            continue
        q, p, d, t = qty[sel], price[sel], disc[sel], tax[sel]
        dp = p * (100 - d)
        ch = dp * (100 + t)
        out[gi] = torch.stack([torch.tensor(sel.numel(), device=qty.device), q.sum(), p.sum(), d.sum(), (dp >> SPLIT).sum(),
                               (dp & lo_mask).sum(), (ch >> SPLIT).sum(), (ch & lo_mask).sum()])
        del sel, q, p, d, t, dp, ch
    return out


def verify_q1(result, parts, expect_all_groups):
    """`result`: the final Q1 batch this rank holds (a pyarrow RecordBatch: the groups of the partitions it owns);
    `parts`: q1_exact_parts summed over every rank's rows. Every group present must equal the exact arithmetic (sums,
    counts, truncating Decimal128 averages); returns (ok, groups seen here)."""
    import decimal

    parts = parts.cpu().tolist()
    want = {}
    for (rf, ls), (cnt, sq, sp, sd, dph, dpl, chh, chl) in zip(Q1_GROUPS, parts):
        if cnt:
            want[(rf, ls)] = (cnt, sq, sp, sd, (dph << SPLIT) + dpl, (chh << SPLIT) + chl)
    ok, seen = True, 0
    D = decimal.Decimal
    with decimal.localcontext() as c:
        c.prec = 60
        for r in result.to_pylist():
            key = (r["l_returnflag"], r["l_linestatus"])
            if key not in want:
                ok = False
                continue
            seen += 1
            cnt, sq, sp, sd, sdp, sch = want[key]
            ok &= r["count_order"] == cnt
            ok &= r["sum_qty"] == D(sq).scaleb(-2) and r["sum_base_price"] == D(sp).scaleb(-2)
            ok &= r["sum_disc_price"] == D(sdp).scaleb(-4) and r["sum_charge"] == D(sch).scaleb(-6)
            ok &= r["avg_qty"] == D(sq * 10**4 // cnt).scaleb(-6) and r["avg_price"] == D(sp * 10**4 // cnt).scaleb(-6)
            ok &= r["avg_disc"] == D(sd * 10**4 // cnt).scaleb(-6)
    if expect_all_groups:
        ok &= seen == len(want)
    return bool(ok), seen


def run_ours(args):
    import torch
    import torch.distributed as dist

    import ballista_b200 as bb

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    bind_to_gpu_numa_node(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)
    ctx = bb.Context(local, stream=stream.cuda_stream)
    comm = bb.Exchange(ctx, rank, world) if world > 1 else None
    n = args.rows
    cols = gen_lineitem_torch(n, 42 + rank, dev)
    batch = wrap_lineitem(bb, ctx, cols, n)
    hp = HotPath(bb, ctx, rank, world, comm)
    verified = {}

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def allreduce(t, op="sum"):
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.SUM if op == "sum" else dist.ReduceOp.MAX)
        return t

    def check_q1(result, m, name):
        """Every rank's rows [0, m) went in; `result` holds the groups of the partitions this rank owns."""
        parts = allreduce(q1_exact_parts(cols, m))
        ok, seen = verify_q1(result, parts, expect_all_groups=world == 1)
        t = allreduce(torch.tensor([0 if ok else 1, seen, result.num_rows], device=dev, dtype=torch.int64))
        bad, seen_all, rows_all = [int(x) for x in t.tolist()]
        groups = int((parts[:, 0] != 0).sum())
        verified[name] = bad == 0 and seen_all == groups and rows_all == groups
        return rows_all

    def time_q1(b, steps, warmup):
        """(ms per step as max over ranks, fused kernel ms, Partial stage ms, launches per step, last result)"""
        res = None
        for _ in range(warmup):
            res = hp.step([b]).download()
        barrier()
        launches0 = ctx.kernel_launches
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        k_ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        fused = []
        ev0.record(stream)
        for i in range(steps):
            k_ev[i][0].record(stream)
            state = hp.partial([b])
            k_ev[i][1].record(stream)
            fused.append(hp.hot_kernel_ms)
            res = hp.shuffle_and_final(state)
        ev1.record(stream)
        barrier()
        ms = float(allreduce(torch.tensor([ev0.elapsed_time(ev1)], device=dev, dtype=torch.float64), "max").item()) / steps
        return (ms, sum(fused) / len(fused), sum(a.elapsed_time(b_) for a, b_ in k_ev) / steps,
                (ctx.kernel_launches - launches0) / steps, res.download())

    # ---- weak scaling (the headline `value`): every GPU holds a whole SF100-sized shard ----------------------------------
    with ClockSampler(local) as clocks:
        ms_per_step, kernel_ms, partial_ms, launches_per_step, result = time_q1(batch, args.steps, max(args.warmup, 3))
    total_rows = n * world
    value = total_rows / (ms_per_step * 1e-3)
    result_rows = check_q1(result, n, "q1_weak")

    # ---- strong scaling: SF100 sharded over the ranks (rank r takes rows [0, n / world) of its shard) --------------------
    strong = None
    if world > 1:
        m = n // world
        sub = wrap_lineitem(bb, ctx, cols, m)
        s_ms, s_kernel, _, _, s_res = time_q1(sub, args.steps, 3)
        check_q1(s_res, m, "q1_strong")
        strong = {"rows_total": m * world, "rows_per_gpu": m, "ms_per_step": s_ms, "rows_per_s": m * world / (s_ms * 1e-3),
                  "fused_kernel_ms": s_kernel}
        del sub
    else:
        strong = {"rows_total": n, "rows_per_gpu": n, "ms_per_step": ms_per_step, "rows_per_s": value, "fused_kernel_ms": kernel_ms}

    # ---- the exchange-bound configuration (BASELINE config 4b): GROUP BY l_orderkey, SF100 sharded -----------------------
    orderkey = None
    if not args.no_orderkey:
        orderkey = run_orderkey(args, bb, ctx, comm, cols, n, rank, world, dev, stream, barrier, allreduce, verified)
    # ---- end to end: HOST RecordBatches of the reference's batch size -> C ABI -> host result -----------------------------
    # The operator is fed as the reference feeds it: one bgpu_agg_push_host call per 8192-row RecordBatch
    # (/root/reference/ballista/core/src/config.rs:180-182), from one thread per input partition
    # (/root/reference/ballista/core/src/execution_plans/coalesce_tasks.rs:183-216). The FFI structs of the batches are made
    # once, outside the timed region (a Rust caller's FFI_ArrowArray costs nothing; pyarrow's export does).
    import ctypes as C
    from concurrent.futures import ThreadPoolExecutor

    from ballista_b200._lib import lib as bgpu_lib
    from ballista_b200.aggregate import export_batch

    ne = min(args.e2e_rows, n)
    host = {k: v[: (ne + 1 if k.endswith("offsets") else ne)].cpu() for k, v in cols.items()}  # pageable, as Arrow buffers are
    hbatches = host_lineitem_batches(host, ne, args.e2e_batch_rows)
    exported = [export_batch(hb) for hb in hbatches]
    refs = [C.byref(a) for a, _ in exported]
    n_feed = max(1, min(args.e2e_threads or 16, (os.cpu_count() or 1) // world))
    pool = ThreadPoolExecutor(n_feed)
    push = bgpu_lib().bgpu_agg_push_host
    e2e_stats = {}

    def e2e_step():
        agg = hp.partial_op
        agg.reset()
        h = agg._h

        def feed(t):
            for r in refs[t::n_feed]:
                rc = push(h, r, None)
                if rc:
                    return rc
            return 0
        rcs = list(pool.map(feed, range(n_feed)))
        if any(rcs):
            ctx.check(max(rcs))
        state = agg.emit()
        e2e_stats.update(agg.ingest_stats)
        return hp.shuffle_and_final(state).download()

    for _ in range(2):
        r = e2e_step()
    barrier()
    e2e_steps = max(2, min(args.steps, 5))
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        r = e2e_step()
    barrier()
    e2e_s = (time.perf_counter() - t0) / e2e_steps
    if world > 1:
        t = torch.tensor([e2e_s], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t.item())
    h2d = e2e_stats.get("wire_bytes", 0)
    d2h = sum(sum(b.size for b in col.buffers() if b is not None) for col in r.columns)
    check_q1(r, ne, "q1_e2e")
    # the same rows as a few large batches through bgpu_batch_upload (round 1's end-to-end leg: every Arrow byte crosses PCIe)
    big = host_lineitem_batches({k: v.pin_memory() for k, v in host.items()}, ne, 8_000_000)

    def upload_step():
        return hp.step(bb.DeviceBatch.upload(ctx, hb) for hb in big).download()
    upload_step()
    barrier()
    t0 = time.perf_counter()
    upload_step()
    barrier()
    upload_s = time.perf_counter() - t0
    pool.shutdown()

    # ---- the other BASELINE configurations (bench_configs.py), once the lineitem shard is gone ---------------------------
    peak, peak_src = load_peaks()
    other = {}
    if not args.no_configs:
        import bench_configs as BC

        del batch, big, hbatches, exported, refs
        cols.clear()
        torch.cuda.empty_cache()
        if world == 1:
            other["cfg2"], verified["cfg2"] = BC.run_cfg2(bb, ctx, torch, dev, stream, peak)
            torch.cuda.empty_cache()
            other["cfg3"], verified["cfg3"] = BC.run_cfg3(bb, ctx, torch, dev, stream, peak, rows=args.cfg3_rows)
        else:
            other["cfg5"], verified["cfg5"] = BC.run_cfg5(bb, ctx, comm, torch, dist, dev, stream, rank, world, total=args.cfg5_rows)
        torch.cuda.empty_cache()

    achieved = n * Q1_BYTES_PER_ROW / (kernel_ms * 1e-3) / 1e9
    line = {
        "metric": "input rows/s & HBM GB/s (% roofline), TPC-H Q1 agg+shuffle", "value": value, "unit": "rows/s",
        "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_per_step,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "i128", "data": "synthetic",
        "config": {"workload": "TPC-H Q1 Partial agg (fused filter+decimal exprs) + hash repartition n=16 + FinalPartitioned, "
                               f"lineitem {n} rows/GPU (SF100 = {SF100_ROWS})", "rows_per_gpu": n, "partitions": N_PARTITIONS,
                   "l2": "input (>= 46 GB at SF100) is larger than L2; no flush needed", "agg_path": hp.last_path,
                   "exchange_path": getattr(hp, "exchange_path", "local"),
                   "result_rows": result_rows, "launches_per_step": launches_per_step,
                   "tail_ms": ms_per_step - kernel_ms},
        "verified": all(verified.values()), "verified_legs": verified,
        "strong": strong, "orderkey": orderkey, "configs": other,
        "clocks": clocks.summary(),
        "e2e": {"value": ne * world / e2e_s, "unit": "rows/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "rows_per_gpu": ne, "batch_rows": args.e2e_batch_rows, "feeder_threads": n_feed,
                "arrow_bytes_per_step": e2e_stats.get("arrow_bytes"), "chunks_per_step": e2e_stats.get("chunks"),
                "launches_not_read_back": e2e_stats.get("async_launches"), "replays": e2e_stats.get("replays"),
                "upload_path_rows_per_s": ne * world / upload_s,
                "note": "host Arrow RecordBatches of batch_rows rows (pageable memory) -> bgpu_agg_push_host from feeder_threads "
                        "threads (narrow wire form in pinned staging, h2d_bytes_per_step = the bytes actually copied) -> "
                        "Partial -> repartition -> Final -> bgpu_batch_download; upload_path = 8M-row pinned batches through "
                        "bgpu_batch_upload (every Arrow byte crosses PCIe)"},
        "gpu_launches": int(round(launches_per_step * args.steps)),
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": load_traffic(n), "kernel": "bgpu_fused (Partial aggregate, one launch per step)", "kernel_ms": kernel_ms,
                     "partial_stage_ms": partial_ms,
                     "bytes_per_row": Q1_BYTES_PER_ROW, "peak_source": peak_src},
    }
    if rank == 0 and world == 1 and not args.no_cpu:  # the CPU baseline is reported at N = 1 only
        if _ORIGINAL_AFFINITY:
            os.sched_setaffinity(0, _ORIGINAL_AFFINITY)  # the CPU baseline gets every host core back
        line["cpu_baseline"] = cpu_baseline(args.cpu_rows)
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()
    if not all(verified.values()):
        sys.exit(f"bench.py: a result differs from exact integer arithmetic: {verified}")


def run_orderkey(args, bb, ctx, comm, cols, n, rank, world, dev, stream, barrier, allreduce, verified):
    """BASELINE config 4b: SELECT l_orderkey, sum(l_quantity), sum(l_extendedprice), count(*) GROUP BY l_orderkey over an
    SF100-sized lineitem SHARDED over the ranks (contiguous row ranges of a globally shuffled table: the worst case, every
    rank sees every order): Partial aggregate -> hash repartition n=16 + exchange over NVLink (fused: the scatter kernel
    stores into the owners' peer-mapped windows) -> FinalPartitioned. Phases timed with CUDA events, max over ranks."""
    import pyarrow as pa
    import torch

    m = (min(args.orderkey_rows, n * world) // world)
    total = m * world
    g = torch.Generator(device=dev).manual_seed(4242)  # the same global table on every rank
    lines = torch.randint(1, 8, (total // 3 + 8,), device=dev, generator=g)
    order = torch.repeat_interleave(torch.arange(lines.numel(), device=dev), lines)[:total]
    del lines
    okey_all = (order // 8) * 32 + order % 8 + 1  # dbgen's sparse order keys: 1-7 lines per order
    del order
    perm = torch.randperm(total, device=dev, generator=g)
    okey_all = okey_all[perm]
    del perm
    distinct_total = int(torch.unique(okey_all).numel())
    sample = torch.unique(okey_all[:: max(1, total // 4096)])  # keys checked exactly, the same on every rank
    okey = okey_all[rank * m:(rank + 1) * m].clone()
    del okey_all
    torch.cuda.empty_cache()
    qty, price = cols["l_quantity"], cols["l_extendedprice"]  # this rank's first m rows of those columns
    d15 = pa.decimal128(15, 2)
    schema = pa.schema([("l_orderkey", pa.int64()), ("l_quantity", d15), ("l_extendedprice", d15)])
    batch = bb.DeviceBatch.wrap_device(ctx, schema, [dict(values=okey.data_ptr()), dict(values=qty.data_ptr()),
                                                     dict(values=price.data_ptr())], m, keepalive=(okey, qty, price))
    aggs = [("sum", 1, "q"), ("sum", 2, "p"), ("count", None, "c")]
    partial_op = bb.GpuAggregateExec(ctx, bb.AggregateMode.Partial, [0], aggs, schema)
    final_op = bb.GpuAggregateExec(ctx, bb.AggregateMode.FinalPartitioned, [0], aggs, schema)
    partitioner = bb.BatchPartitioner.try_new(bb.Partitioning.Hash([0], N_PARTITIONS), ctx)
    STATE = 48  # key 8 + 2 x i128 sum + i64 count
    window = int(STATE * m * 1.25) + (64 << 20)

    def ev():
        return torch.cuda.Event(enable_timing=True)

    def step():
        e = [ev() for _ in range(4)]
        e[0].record(stream)
        partial_op.reset()
        partial_op.update(batch)
        state = partial_op.emit()
        e[1].record(stream)
        if world > 1:
            part = comm.partition_exchange(state, [0], N_PARTITIONS, window_bytes=window)
            sent = comm.last_sent_bytes
            path = comm.last_path
        else:
            part, _ = partitioner.partition_device(state)
            sent, path = 0, "local"
        groups = state.num_rows
        del state
        e[2].record(stream)
        final_op.reset()
        final_op.update(part)
        del part
        out = final_op.emit()
        e[3].record(stream)
        e[3].synchronize()
        return [e[i].elapsed_time(e[i + 1]) for i in range(3)], groups, sent, path, out

    best, out = None, None
    for i in range(2 + max(1, min(args.steps, 3))):
        del out
        barrier()
        ms, groups, sent, path, out = step()
        t = allreduce(torch.tensor(ms + [sum(ms)], device=dev, dtype=torch.float64), "max").tolist()
        if i >= 2 and (best is None or t[3] < best[3]):
            best = t
    # ---- verification: group count, checksums over everything, exact sums for a sample of keys -------------------------
    keys = out.column_tensor(0)[:, 0]
    q, p, c = out.column_tensor(1), out.column_tensor(2), out.column_tensor(3)[:, 0]
    lo_q, lo_p = qty[:m, 0], price[:m, 0]
    tot = allreduce(torch.stack([torch.tensor(keys.numel(), device=dev), c.sum(), q[:, 0].sum(), p[:, 0].sum(),
                                 torch.tensor(m, device=dev), lo_q.sum(), lo_p.sum()]))
    n_groups, cnt_sum, q_sum, p_sum, rows_in, q_in, p_in = [int(x) for x in tot.tolist()]
    ok = n_groups == distinct_total and cnt_sum == rows_in and q_sum == q_in and p_sum == p_in
    ok &= bool((q[:, 1] == 0).all()) and bool((p[:, 1] == 0).all())  # every sum here is positive and below 2^63
    pos = torch.searchsorted(sample, okey).clamp_(max=sample.numel() - 1)
    hit = sample[pos] == okey
    want = torch.zeros((sample.numel(), 3), dtype=torch.int64, device=dev)
    want.index_add_(0, pos[hit], torch.stack([lo_q[hit], lo_p[hit], torch.ones_like(lo_q[hit])], dim=1))
    want = allreduce(want)
    rpos = torch.searchsorted(sample, keys).clamp_(max=sample.numel() - 1)
    rhit = sample[rpos] == keys
    got = torch.stack([q[rhit, 0], p[rhit, 0], c[rhit]], dim=1)
    ok &= bool((got == want[rpos[rhit]]).all())
    found = int(allreduce(torch.tensor([int(rhit.sum())], device=dev, dtype=torch.int64)).item())
    ok &= found == sample.numel()
    verified["orderkey"] = bool(int(allreduce(torch.tensor([0 if ok else 1], device=dev, dtype=torch.int64)).item()) == 0)
    stats = allreduce(torch.tensor([groups, sent], device=dev, dtype=torch.int64)).tolist()
    del out, keys, q, p, c, batch, okey
    torch.cuda.empty_cache()
    return {"workload": "GROUP BY l_orderkey (sum, sum, count), globally shuffled rows sharded over the ranks",
            "rows_total": total, "rows_per_gpu": m, "groups": distinct_total, "partial_groups_total": int(stats[0]),
            "partial_ms": best[0], "exchange_ms": best[1], "final_ms": best[2], "ms_per_step": best[3],
            "rows_per_s": total / (best[3] * 1e-3), "agg_path": partial_op.last_path, "exchange_path": path,
            "nvlink": {"bytes_sent_per_gpu": stats[1] / world, "GBps": (stats[1] / world) / (best[1] * 1e-3) / 1e9 if world > 1 else None,
                       "frac_of_900": (stats[1] / world) / (best[1] * 1e-3) / 1e9 / 900.0 if world > 1 else None}}


# ---- CPU arm: the oracle's restatement of the reference path, timed on the host cores --------------------------------
def cpu_baseline(rows, passes=4, warm=0):
    """The CPU restatement of the hot path over a bounded sample: the sample is generated once (untimed), the pipeline
    runs `passes` times over it with every host thread (passes x rows / value = the CPU seconds spent: a few seconds of
    wall clock, tens of core-seconds); `value` is the best pass, the mean is reported next to it."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle as O

    cores = os.cpu_count() or 1
    cols = O.gen_lineitem_numpy(rows, 42, threads=cores)
    secs = []
    res = None
    for i in range(warm + max(1, passes)):
        res = O.q1_pipeline(rows, seed=42, threads=cores, partitions=N_PARTITIONS, cols=cols)
        if i >= warm:
            secs.append(res["seconds"])
    best = min(secs)
    return {"value": rows / best, "unit": "rows/s", "cores": cores, "kind": "port",
            "sample": f"{rows} synthetic lineitem rows (in host memory), {cores} threads = {cores} input partitions of "
                      f"8192-row batches, Partial agg + hash repartition n={N_PARTITIONS} + FinalPartitioned; oracle/pipeline.cpp; "
                      f"best of {len(secs)} passes over the sample",
            "seconds": best, "mean_value": rows * len(secs) / sum(secs), "passes": len(secs), "result_rows": res["groups"]}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    # a step = one pass of the pipeline over the bounded sample (generated once, untimed)
    n_done, n_warm = max(1, args.steps), max(0, args.warmup)  # the driver's --steps / --warmup, as given
    rows = args.cpu_rows
    t0 = time.perf_counter()
    base = cpu_baseline(rows, passes=n_done, warm=n_warm)
    sec = (time.perf_counter() - t0) / n_done
    world = int(os.environ.get("WORLD_SIZE", "1"))
    line = {"impl": "reference", "metric": "input rows/s & HBM GB/s (% roofline), TPC-H Q1 agg+shuffle",
            "value": base["value"], "unit": "rows/s", "n_gpus": world, "steps": n_done, "warmup": n_warm,
            # the step's timed region is the pipeline itself (oracle/pipeline.cpp times it after generating its sample
            # in host memory); wall_ms_per_step also holds process start-up and data generation
            "ms_per_step": base["seconds"] * 1e3, "wall_ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "i128",
            "data": "synthetic",
            "config": {"workload": f"TPC-H Q1 Partial agg + hash repartition n=16 + FinalPartitioned, CPU restatement "
                                   f"(Ballista standalone cannot be built here: no Rust toolchain), bounded sample {rows} rows",
                       "rows": rows, "partitions": N_PARTITIONS},
            "cpu_baseline": base, "gpu_launches": 0,
            "e2e": {"value": base["value"], "unit": "rows/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


if __name__ == "__main__":
    a = parse_args()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)
