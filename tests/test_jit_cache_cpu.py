"""The on-disk cache of run-time compiled kernels (csrc/jit.cu): the second PROCESS that sees a plan compiles nothing.
No device needed (bgpu_agg_debug_compile is NVRTC only)."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

CHILD = r"""
import ctypes as C, json, sys, time
sys.path.insert(0, %r)
import ballista_b200 as bb
from ballista_b200.aggregate import debug_compile
sys.path.insert(0, %r)
from helpers import q1_plan
import bench
gi, aggs, filt = q1_plan()
t0 = time.perf_counter()
debug_compile(bb.AggregateMode.Partial, gi, aggs, bench.lineitem_schema(), filter=filt, threads=640, rows_per_thread=1, min_ctas=1)
dt = time.perf_counter() - t0
out = (C.c_uint64 * 4)()
bb.lib().bgpu_jit_stats(out)
print(json.dumps({"stats": [int(x) for x in out], "seconds": dt}))
"""


def _run(cache_dir):
    env = dict(os.environ, BGPU_CACHE_DIR=cache_dir)
    out = subprocess.run([sys.executable, "-c", CHILD % (ROOT, os.path.join(ROOT, "tests"))], env=env, capture_output=True,
                         text=True, timeout=300)
    assert out.returncode == 0, out.stderr[-2000:]
    return json.loads(out.stdout.strip().splitlines()[-1])


def test_second_process_compiles_nothing(tmp_path):
    first = _run(str(tmp_path))
    assert first["stats"][0] == 1 and first["stats"][1] == 0 and first["stats"][3] == 1   # compiled once, stored once
    files = [f for f in os.listdir(tmp_path) if f.endswith(".cubin")]
    assert len(files) == 1
    second = _run(str(tmp_path))
    assert second["stats"][0] == 0 and second["stats"][1] == 1                            # nothing compiled: read from disk
    assert second["seconds"] < first["seconds"]
    # a damaged file is not trusted: it costs a recompile, never a wrong kernel
    path = os.path.join(tmp_path, files[0])
    blob = bytearray(open(path, "rb").read())
    blob[40] ^= 0xFF  # inside the stored tag / source
    open(path, "wb").write(bytes(blob))
    third = _run(str(tmp_path))
    assert third["stats"][0] == 1 and third["stats"][1] == 0


def test_cache_can_be_switched_off(tmp_path):
    env_dir = "off"
    r = _run(env_dir)
    assert r["stats"][0] == 1 and r["stats"][3] == 0
