"""Fuzz of the host-side parsers that face bytes from other processes: libbgpu's IPC stream reader
(schema message, flatbuffer walking) and its LZ4 frame walker + block decoder (the kernel's logic run on the host through
the test hook). Mutates good inputs (byte flips, truncation, random words in the metadata); every call must either
succeed or return an error code - no crash, no exception across the C ABI, no unreadable error text.

    python tools/fuzz_ipc.py <seed> <iterations>
"""
import sys, io, random
import os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, pyarrow as pa
import ballista_b200 as bb
from ballista_b200 import ipc
from helpers import random_batch, independent_frame
rng = np.random.default_rng(1)
b = random_batch(rng, 200, with_dict=False)
sink = io.BytesIO()
with pa.ipc.new_stream(sink, b.schema, options=pa.ipc.IpcWriteOptions(compression="lz4")) as w:
    w.write_batch(b)
good = sink.getvalue()
hc = ipc._HostContext()
R = random.Random(int(sys.argv[1]))
n_ok = n_err = 0
for it in range(int(sys.argv[2])):
    data = bytearray(good)
    kind = R.randrange(3)
    if kind == 0:
        for _ in range(R.randrange(1, 8)):
            data[R.randrange(len(data))] = R.randrange(256)
    elif kind == 1:
        data = data[:R.randrange(len(data))]
    else:
        i = R.randrange(min(len(data), 1200)); data[i:i+4] = R.randrange(2**32).to_bytes(4, "little")
    try:
        r = ipc.StreamReader(hc, bytes(data)); _ = r.schema; r.close(); n_ok += 1
    except bb.BgpuError:
        n_err += 1
# lz4 frames
raw = bytes(rng.integers(0, 4, 100000, dtype=np.uint8))
import pyarrow as pa
frame = pa.compress(raw, codec="lz4", asbytes=True)
frame2 = independent_frame(raw)
for it in range(int(sys.argv[2])):
    f = bytearray(R.choice([frame, frame2]))
    for _ in range(R.randrange(1, 6)):
        f[R.randrange(len(f))] = R.randrange(256)
    if R.randrange(4) == 0: f = f[:R.randrange(len(f))]
    try:
        out = ipc.lz4_frame_decode_host(bytes(f), len(raw), hc); n_ok += 1
    except bb.BgpuError:
        n_err += 1
print("ok", n_ok, "err", n_err)
