"""Re-entrancy of the host side (SURVEY.md §8b "Threading": an executor runs several tasks per process on a thread pool,
/root/reference/ballista/executor/src/executor_server.rs:1027-1032): the host-only entry points - schema framing, the IPC
reader's stream parser, the aggregate planner and its NVRTC compile - called from several threads at once, each on its
own context (ctypes drops the GIL for the duration of a call, so these really overlap). No GPU needed."""
import io
import os
import sys
import threading

import numpy as np
import pyarrow as pa

from helpers import random_batch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_host_entry_points_from_concurrent_threads(bb):
    sys.path.insert(0, ROOT)
    import bench
    from ballista_b200 import ipc
    from ballista_b200.aggregate import debug_compile

    gi, aggs, filt = bench.q1_plan()
    schema = bench.lineitem_schema()
    batch = random_batch(np.random.default_rng(0), 100, with_dict=False)
    sink = io.BytesIO()
    with pa.ipc.new_stream(sink, batch.schema) as w:
        w.write_batch(batch)
    data = sink.getvalue()
    want_msg = ipc.schema_message(batch.schema)
    errors = []

    def task(i):
        try:
            for _ in range(2):
                assert ipc.schema_message(batch.schema) == want_msg
                r = ipc.StreamReader(ipc._HostContext(), data)
                assert r.schema.equals(batch.schema)
                r.close()
                if i % 2 == 0:
                    debug_compile(0, gi, aggs, schema, filter=filt)
                else:
                    s2 = pa.schema([("k", pa.int64()), ("v", pa.float64()), ("d", pa.decimal128(20, 4))])
                    debug_compile(0, [0], [("sum", 1, "s"), ("avg", 2, "a"), ("count", None, "c")], s2, flags=i % 8)
        except Exception as e:  # noqa: BLE001 - collected and reported below
            errors.append((i, repr(e)))
    threads = [threading.Thread(target=task, args=(i,)) for i in range(4)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert errors == []
