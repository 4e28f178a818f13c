"""The ordered fan-in of CoalesceTasksExec (/root/reference/ballista/core/src/execution_plans/coalesce_tasks.rs:148-173:
`streaming_merge` of the partitions' sorted streams, ties to the lower stream index) = concat in stream order + a STABLE
sort on the device (bgpu_batch_sort). Checked against Arrow's own stable sort, bit for bit incl. the order of ties."""
import numpy as np
import pyarrow as pa
import pyarrow.compute as pc
import pytest

from helpers import random_batch

pytestmark = pytest.mark.gpu


def _arrow_sorted(table, keys):
    """keys: [(name, descending, nulls_first)]. Arrow takes ONE null placement per call: sort key by key, least significant
    first, each pass stable (the way the device does it)."""
    idx = pa.array(np.arange(table.num_rows))
    for name, desc, nulls_first in reversed(keys):
        col = table[name].take(idx)
        order = pc.sort_indices(pa.table({"k": col}), sort_keys=[("k", "descending" if desc else "ascending")],
                                null_placement="at_start" if nulls_first else "at_end")
        idx = idx.take(order)
    return table.take(idx)


@pytest.mark.parametrize("keys", [[("k64", False, False)], [("small", False, False), ("k32", True, True)], [("s", False, True)],
                                  [("flag", True, False), ("dec", False, False)], [("f", True, True), ("u8", False, False)],
                                  [("b", False, False), ("date", True, False), ("i16", False, True)], [("s", True, False), ("k64", False, False)]])
def test_sort_matches_arrows_stable_sort(bb, ctx, keys):
    rng = np.random.default_rng(len(keys) * 7 + len(keys[0][0]))
    b = random_batch(rng, 50_000, null_frac=0.15)
    want = _arrow_sorted(pa.Table.from_batches([b]), keys)
    names = b.schema.names
    got = bb.DeviceBatch.upload(ctx, b).sort([names.index(k) for k, _, _ in keys], [d for _, d, _ in keys], [n for _, _, n in keys]).download()
    gt = pa.Table.from_batches([got])
    for name in names:
        a, w = gt[name].combine_chunks(), want[name].combine_chunks()
        if pa.types.is_dictionary(a.type):
            a, w = a.cast(a.type.value_type), w.cast(w.type.value_type)
        assert a.equals(w), f"column {name} differs under keys {keys}"


def test_ordered_fan_in_equals_a_merge_of_sorted_partitions(bb, ctx):
    """Three sorted partitions with many equal keys across them: the fan-in yields the merged order, ties by partition."""
    rng = np.random.default_rng(5)
    parts = []
    for p in range(3):
        n = 20_000 + 1000 * p
        t = pa.table({"k": pa.array(np.sort(rng.integers(0, 500, n)), type=pa.int64()),
                      "name": pa.array([f"n{int(x):04d}" for x in rng.integers(0, 50, n)]),
                      "src": pa.array(np.full(n, p), type=pa.int32()), "row": pa.array(np.arange(n), type=pa.int64())})
        t = t.take(pc.sort_indices(t, sort_keys=[("k", "ascending"), ("name", "descending")]))  # stable: sorted by (k, name desc)
        parts.append(t.combine_chunks().to_batches()[0].slice(0, n))
    src = {p: [parts[p].slice(0, 7000), parts[p].slice(7000)] for p in range(3)}
    fan = bb.CoalesceTasksExec(lambda p: iter(src[p]), [0, 1, 2], schema=parts[0].schema,
                               order_by=[(0, False, False), (1, True, False)], ctx=ctx)
    assert fan.output_ordering() == [(0, False, False), (1, True, False)]
    out = list(fan.execute())
    assert len(out) == 1
    got = pa.Table.from_batches([out[0].download()])
    whole = pa.concat_tables([pa.Table.from_batches([b]) for b in parts])
    want = whole.take(pc.sort_indices(whole, sort_keys=[("k", "ascending"), ("name", "descending")]))  # stable = merge with ties by stream
    assert got.equals(want.combine_chunks())
    # a single partition passes through untouched (coalesce_tasks.rs:144-146)
    one = bb.CoalesceTasksExec(lambda p: iter(src[p]), [1], order_by=[(0, False, False)], ctx=ctx)
    assert [b.num_rows for b in one.execute()] == [7000, parts[1].num_rows - 7000]


def test_sort_edge_cases(bb, ctx):
    empty = pa.RecordBatch.from_arrays([pa.array([], type=pa.int64()), pa.array([], type=pa.string())], names=["a", "s"])
    assert bb.DeviceBatch.upload(ctx, empty).sort([0]).download().num_rows == 0
    b = pa.RecordBatch.from_arrays([pa.array(["ab", "a", "", None, "ab\0", "abc" * 20, "a"]), pa.array([1.0, -0.0, 0.0, float("nan"), -1.5, None, 2.0])],
                                   names=["s", "f"])
    d = bb.DeviceBatch.upload(ctx, b)
    assert d.sort([0]).download().column(0).to_pylist() == ["", "a", "a", "ab", "ab\0", "abc" * 20, None]
    got = d.sort([1], [False], [True]).download().column(1).to_pylist()
    assert got[0] is None and got[1:5] == [-1.5, -0.0, 0.0, 1.0] and got[5] == 2.0 and got[6] != got[6]  # NaN last (total order)
