"""Fuzz of the aggregate planner's host side (bgpu_agg_create_expr's validation, reached without a device through
bgpu_agg_debug_compile): random postfix expression programs, column indices, group columns and aggregate functions over
a fixed schema. Every call must return a code - no crash; programs that pass validation are compiled by NVRTC, which must
succeed (a plan the validator accepts and the code generator cannot compile is a bug).

    python tools/fuzz_expr.py <seed> <iterations> [wellformed]
"""
import collections
import ctypes as C
import os
import random
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import pyarrow as pa  # noqa: E402

import ballista_b200 as bb  # noqa: E402
from ballista_b200._lib import ArrowSchema, ExprNode, lib  # noqa: E402

schema = pa.schema([("i", pa.int64()), ("f", pa.float64()), ("d", pa.decimal128(15, 2)), ("e", pa.decimal128(38, 10)),
                    ("s", pa.string()), ("t", pa.date32()), ("k", pa.int32()), ("b", pa.bool_()),
                    ("x", pa.dictionary(pa.int32(), pa.string()))])
R = random.Random(int(sys.argv[1]))
L = lib()
h = C.c_void_p()
L.bgpu_ctx_create(0, C.byref(h))
out = collections.Counter()


def node():
    k = R.randrange(10)
    if k < 4:
        return ExprNode(0, R.choice([0, 1, 2, 3, 4, 5, 6, 7, 8, -1, 9, 1 << 30]), 0, 0, 0, 0)
    if k < 6:
        return ExprNode(1, R.choice([0, 1, 2, 3, 4, -1]), R.choice([0, 1, 15, 38, 39, -5]), R.choice([0, 2, 10, 38, 40, -3]),
                        R.getrandbits(64) if R.randrange(3) == 0 else R.randrange(200), R.choice([0, 0, (1 << 64) - 1, R.getrandbits(64)]))
    return ExprNode(R.choice([2, 3, 4, 4, 4, 5, 6, 7, 8, 9, 10, 11, 12, -1, 99]), 0, 0, 0, 0, 0)


def tree(depth, out_nodes):
    """A well-formed random expression over the numeric columns (postfix)."""
    if depth == 0 or R.randrange(3) == 0:
        if R.randrange(4) == 0:
            t = R.choice([0, 1, 2])
            out_nodes.append(ExprNode(1, t, 0 if t != 2 else R.choice([1, 3, 10]), 0 if t != 2 else R.choice([0, 2]),
                                      R.randrange(1, 100) if t != 1 else 0x3FF8000000000000, 0))
        else:
            out_nodes.append(ExprNode(0, R.choice([0, 1, 2, 3, 5, 6]), 0, 0, 0, 0))
        return
    tree(depth - 1, out_nodes)
    tree(depth - 1, out_nodes)
    out_nodes.append(ExprNode(R.choice([2, 3, 4, 4]), 0, 0, 0, 0, 0))


def predicate(out_nodes):
    n = R.randrange(1, 6)
    for i in range(n):
        c = R.choice([0, 1, 2, 3, 5, 6])
        out_nodes.append(ExprNode(0, c, 0, 0, 0, 0))
        t = {0: 0, 1: 1, 2: 2, 3: 2, 5: R.choice([3, 0]), 6: 0}[c] if R.randrange(8) else R.choice([0, 1, 2, 3])
        out_nodes.append(ExprNode(1, t, 10 if t == 2 else 0, R.choice([0, 2, 4]) if t == 2 else 0,
                                  R.randrange(1, 20000) if t != 1 else 0x4059000000000000, 0))
        out_nodes.append(ExprNode(R.choice([5, 6, 7, 8, 9, 10]), 0, 0, 0, 0, 0))
        if i:
            out_nodes.append(ExprNode(11, 0, 0, 0, 0, 0))


WELL_FORMED = len(sys.argv) > 3 and sys.argv[3] == "wellformed"
for it in range(int(sys.argv[2])):
    n_agg = R.randrange(0, 5)
    nodes, offs, fns = [], [0], []
    if WELL_FORMED:
        n_agg = R.randrange(1, 7)
        for _ in range(n_agg):
            fns.append(R.choice([2, 2, 3, 4]))
            if fns[-1] != 4 or R.randrange(2):
                tree(R.randrange(0, 4), nodes)
            offs.append(len(nodes))
        filt = []
        mode = R.choice([0, 0, 2, 3])
        if mode != 2 and R.randrange(2):   # a filter applies to raw input rows only
            predicate(filt)
        groups = R.sample([0, 4, 5, 6, 7, 8], R.randrange(1, 4))
    else:
        for _ in range(n_agg):
            fns.append(R.choice([2, 2, 3, 4, 0, 1, 7, -1]))
            for _ in range(R.choice([0, 1, 1, 3, 3, 5, 7])):
                nodes.append(node())
            offs.append(len(nodes) if R.randrange(20) else R.choice([-1, len(nodes) + 3, 0]))
        filt = [node() for _ in range(R.choice([0, 0, 3, 3, 7, 5]))]
        groups = [R.choice([0, 4, 5, 6, 7, 8, 2, 1, -1, 9]) for _ in range(R.choice([0, 1, 1, 2, 3, 9]))]
    s = ArrowSchema()
    schema._export_to_c(C.addressof(s))
    rc = L.bgpu_agg_debug_compile(h, mode if WELL_FORMED else R.choice([0, 0, 2, 1, 3, 5]), C.byref(s), (C.c_int32 * max(len(groups), 1))(*groups), len(groups),
                                  (C.c_int32 * max(n_agg, 1))(*fns), (ExprNode * max(len(nodes), 1))(*nodes),
                                  (C.c_int32 * len(offs))(*offs), n_agg, (ExprNode * max(len(filt), 1))(*filt), len(filt),
                                  R.randrange(16), R.choice([256, 256, 640, 1024]), R.choice([1, 4, 4, 8]), R.choice([1, 2, 2]), None)
    bb.runtime._release(s)
    msg = L.bgpu_last_error(h).decode(errors="replace") if rc else "compiled"
    out[(rc, msg[:70])] += 1
for (rc, msg), v in out.most_common(40):
    print(v, rc, msg)
