"""Several tasks in one process (SURVEY.md §8b "Threading": the executor's DedicatedExecutor runs tasks of the same and of
different jobs concurrently, /root/reference/ballista/executor/src/executor_server.rs:1027-1032): four threads, each with
its own bgpu_ctx (stream + pool), run the hot path at once - the same plan shapes with different input sizes and group
counts, so that they share the run-time compiled kernels while asking for different amounts of shared memory - and must
each get exactly what a single thread gets."""
import os
import sys
import threading

import numpy as np
import pyarrow as pa
import pytest

from helpers import lineitem_batch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def q1(bb, ctx, batch):
    sys.path.insert(0, ROOT)
    import bench

    gi, aggs, filt = bench.q1_plan()
    part = bb.GpuAggregateExec(ctx, bb.AggregateMode.Partial, gi, aggs, batch.schema, filter=filt)
    part.update(batch)
    state, _ = bb.BatchPartitioner.try_new(bb.Partitioning.Hash([0, 1], 16), ctx).partition_device(part.emit())
    fin = bb.GpuAggregateExec(ctx, bb.AggregateMode.FinalPartitioned, gi, aggs, batch.schema)
    fin.update(state)
    return pa.Table.from_batches([fin.emit().download()]).sort_by([("l_returnflag", "ascending"), ("l_linestatus", "ascending")])


def group_by(bb, ctx, batch):
    aggs = [("sum", 1, "s"), ("count", None, "c"), ("avg", 2, "a")]
    part = bb.GpuAggregateExec(ctx, bb.AggregateMode.Partial, [0], aggs, batch.schema)
    part.update(batch)
    state, _ = bb.BatchPartitioner.try_new(bb.Partitioning.Hash([0], 8), ctx).partition_device(part.emit())
    fin = bb.GpuAggregateExec(ctx, bb.AggregateMode.FinalPartitioned, [0], aggs, batch.schema)
    fin.update(state)
    return pa.Table.from_batches([fin.emit().download()]).sort_by("k")


@pytest.mark.timeout(900)
def test_concurrent_tasks_each_on_their_own_context(bb):
    rng = np.random.default_rng(31)
    work = []
    for i, (rows, groups) in enumerate([(40_000, 50), (150_000, 700), (90_000, 5_000), (60_000, 90_000)]):
        li = lineitem_batch(rng, rows)
        gb = pa.RecordBatch.from_arrays([pa.array(rng.integers(0, groups, rows), type=pa.int64()),
                                         pa.array(rng.integers(-1000, 1000, rows), type=pa.int64()),
                                         pa.array(rng.random(rows))], names=["k", "v", "f"])
        work.append((li, gb))
    # what one thread gets
    with bb.Context(0) as ctx:
        want = [(q1(bb, ctx, li), group_by(bb, ctx, gb)) for li, gb in work]
    got, errors = [None] * len(work), []

    def task(i):
        try:
            with bb.Context(0) as ctx:
                for _ in range(3):
                    got[i] = (q1(bb, ctx, work[i][0]), group_by(bb, ctx, work[i][1]))
        except Exception as e:  # noqa: BLE001 - collected and reported below
            errors.append((i, repr(e)))
    # daemon threads + bounded joins: a deadlock must fail this test, not hang the suite
    threads = [threading.Thread(target=task, args=(i,), daemon=True) for i in range(len(work))]
    for t in threads:
        t.start()
    for t in threads:
        t.join(timeout=300)
    assert not any(t.is_alive() for t in threads), "a task is still running after 300 s"
    assert errors == []
    for i, ((wq, wg), (gq, gg)) in enumerate(zip(want, got)):
        assert gq.equals(wq), f"task {i}: Q1 result differs under concurrency"
        # integer sums and counts exact; Float64 averages within the tolerance the suite states for float sums
        assert gg.column("k").equals(wg.column("k")) and gg.column("s").equals(wg.column("s")) and gg.column("c").equals(wg.column("c"))
        assert np.allclose(gg.column("a").to_numpy(), wg.column("a").to_numpy(), rtol=1e-9, atol=0)
