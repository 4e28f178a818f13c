"""Malformed bytes never take the process down (SURVEY.md §8b "Errors": never abort, never throw across the ABI):
a short run of tools/fuzz_ipc.py in a child process, plus the specific malformations it found."""
import io
import os
import subprocess
import sys

import pyarrow as pa
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("seed", [1, 2])
def test_fuzzed_streams_and_frames_are_errors_not_crashes(seed):
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "fuzz_ipc.py"), str(seed), "1500"], capture_output=True,
                       text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    assert r.stdout.startswith("ok ")


def _stream(schema):
    sink = io.BytesIO()
    with pa.ipc.new_stream(sink, schema):
        pass
    return sink.getvalue()


def test_schema_values_arrow_would_reject_are_rejected_here(bb):
    from ballista_b200 import ipc

    hc = ipc._HostContext()
    good = _stream(pa.schema([("abc", pa.decimal128(15, 2))]))
    assert ipc.StreamReader(hc, good).schema.field("abc").type == pa.decimal128(15, 2)
    # a field name that is not UTF-8: an error whose text is still readable
    with pytest.raises(bb.BgpuError, match="not UTF-8"):
        ipc.StreamReader(hc, good.replace(b"abc", b"\xff\xfe\xfd"))
    # Decimal128 precision 99: one of the places the int32 15 occurs is the precision
    hits = 0
    at = good.find((15).to_bytes(4, "little"))
    while at >= 0:
        bad = good[:at] + (99).to_bytes(4, "little") + good[at + 4:]
        try:
            ipc.StreamReader(hc, bad).schema
        except bb.BgpuError as e:
            hits += "precision / scale out of range" in str(e)
        at = good.find((15).to_bytes(4, "little"), at + 1)
    assert hits == 1
    # a field count no message of this size can hold must not size an allocation
    words = [i for i in range(0, len(good) - 4, 4) if good[i:i + 4] == (1).to_bytes(4, "little")]
    for at in words:
        bad = good[:at] + (0x7FFFFFF0).to_bytes(4, "little") + good[at + 4:]
        try:
            ipc.StreamReader(hc, bad).schema
        except bb.BgpuError:
            pass


def test_fuzzed_aggregate_plans_are_errors_not_crashes():
    """tools/fuzz_expr.py: random expression programs, column indices and modes against the aggregate planner's host side
    (reached without a device through bgpu_agg_debug_compile); what passes validation is compiled by NVRTC."""
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "fuzz_expr.py"), "3", "250"], capture_output=True, text=True,
                       timeout=900)
    assert r.returncode == 0, r.stderr[-2000:]
    rows = [line.split(" ", 2) for line in r.stdout.splitlines()]
    assert sum(int(n) for n, _, _ in rows) == 250
    # 0 = compiled, 1 = BGPU_EINVAL, 2 = BGPU_EUNSUPPORTED; anything else (a CUDA / NVRTC failure) is a bug
    assert {int(rc) for _, rc, _ in rows} <= {0, 1, 2}, r.stdout
