"""Generates tests/golden/ahash_vectors.json with a pure-Python big-int implementation of the ahash 0.8
fallback hasher (seeds 0,0,0,0) written directly from the field mapping read off compiled ahash-0.8.12
(oracle/evidence/README.md). It is independent of both oracle/ahash_fallback.h (C) and csrc/hash.cuh (CUDA):
the three must agree. Run: python tests/golden/make_ahash_vectors.py
"""
import json
import os
import struct

M = (1 << 64) - 1
PI2 = [0x452821E638D01377, 0xBE5466CF34E90C6C, 0xC0AC29B7C97C50DD, 0x3F84D5B5B5470917]
MULT = 6364136223846793005
PAD, BUF, E0, E1 = PI2[0], PI2[1], PI2[2], PI2[3]


def fm(a, b):
    r = a * b
    return (r & M) ^ (r >> 64)


def rotl(x, r):
    r &= 63
    return ((x << r) | (x >> (64 - r))) & M if r else x


def upd(buf, x):
    return fm(x ^ buf, MULT)


def large(buf, lo, hi):
    return rotl(((buf + PAD) & M) ^ fm(lo ^ E0, hi ^ E1), 23)


def fin(buf):
    return rotl(fm(buf, PAD), buf & 63)


def le(b):
    return int.from_bytes(b, "little")


def write(buf, d):
    n = len(d)
    buf = ((buf + n) * MULT) & M
    if n > 16:
        buf = large(buf, le(d[n - 16:n - 8]), le(d[n - 8:]))
        while len(d) > 16:
            buf = large(buf, le(d[:8]), le(d[8:16]))
            d = d[16:]
    elif n > 8:
        buf = large(buf, le(d[:8]), le(d[n - 8:]))
    elif n >= 4:
        buf = large(buf, le(d[:4]), le(d[n - 4:]))
    elif n >= 2:
        buf = large(buf, le(d[:2]), d[n - 1])
    elif n == 1:
        buf = large(buf, d[0], d[0])
    else:
        buf = large(buf, 0, 0)
    return buf


def h_int(x, width):  # zero-extended bit pattern of a width-byte integer
    return fin(upd(BUF, x & ((1 << (8 * width)) - 1)))


def h_i128(x):
    x &= (1 << 128) - 1
    return fin(large(BUF, x & M, x >> 64))


def h_str(s):
    return fin(upd(write(BUF, s.encode()), 0xFF))


def h_bytes(b):
    return fin(write(upd(BUF, len(b)), b))


def combine(l, r):
    return ((17 * 37 + l) * 37 + r) & M


if __name__ == "__main__":
    ints = [0, 1, 2, -1, 42, 2**31 - 1, -2**31, 2**63 - 1, -2**63, 1234567890123]
    strs = ["", "a", "ab", "abc", "abcd", "hello", "abcdefgh", "abcdefghi", "0123456789abcdef",
            "0123456789abcdefg", "x" * 31, "y" * 32, "z" * 33, "A", "F", "N", "O", "R", "id042", "id0000001234",
            "héllo wörld ✓"]
    decs = [0, 1, -1, 12345, -99999999999999999999, 10**37, -(10**37)]
    out = {
        "i64": [[x, h_int(x, 8)] for x in ints],
        "i32": [[x, h_int(x, 4)] for x in ints if -2**31 <= x < 2**31],
        "u32": [[x, h_int(x, 4)] for x in (1, 2, 3, 4, 2**32 - 1)],
        "i16": [[x, h_int(x, 2)] for x in (0, 1, -1, 32767, -32768)],
        "i8": [[x, h_int(x, 1)] for x in (0, 1, -1, 127, -128)],
        "f64": [[x, h_int(struct.unpack("<Q", struct.pack("<d", x))[0], 8)] for x in (0.0, -0.0, 1.5, -3.25, 1e300)],
        "bool": [[int(x), h_int(int(x), 1)] for x in (False, True)],
        "utf8": [[s, h_str(s)] for s in strs],
        "binary": [[s.encode().hex(), h_bytes(s.encode())] for s in strs[:12]],
        "decimal128": [[str(x), h_i128(x)] for x in decs],
        # multi-column: hash(col0) then combine_hashes(hash(col1), running)
        "i64_utf8": [[a, s, combine(h_str(s), h_int(a, 8))] for a, s in zip(ints, strs)],
    }
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "ahash_vectors.json")
    with open(path, "w") as f:
        json.dump(out, f, indent=1, ensure_ascii=False)
    print("wrote", path)
