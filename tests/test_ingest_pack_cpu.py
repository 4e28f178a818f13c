"""Host side of the streamed input (csrc/ingest_pack.cpp): the wire form of a fixed-width column segment decodes back to
the column bit for bit (base + unsigned delta, mod 2^64), for both builds of the packer, and is as narrow as the value
range allows. No device needed."""
import ctypes as C

import numpy as np
import pytest

from ballista_b200._lib import lib

RAW = 255


def _pack(fmt, arr, which):
    n = len(arr)
    out = np.zeros(max(n, 1) * 16, dtype=np.uint8)
    base, code = C.c_int64(), C.c_uint32()
    rc = lib().bgpu_debug_pack_fixed_host(fmt.encode(), arr.ctypes.data_as(C.c_void_p), n if not fmt.startswith("d") else n // 2,
                                          which, C.byref(base), C.byref(code), out.ctypes.data_as(C.c_void_p))
    if rc == 2 and which in (0, 2):
        pytest.skip("host without AVX2 / AVX-512")
    assert rc == 0
    return base.value, code.value, out


def _decode(base, code, packed, n, dtype):
    if code == 0:
        delta = np.zeros(n, dtype=np.uint64)
    else:
        delta = packed[: n * code].view({1: np.uint8, 2: np.uint16, 4: np.uint32, 8: np.uint64}[code]).astype(np.uint64)
    return (np.uint64(base & (2**64 - 1)) + delta).astype(np.uint64).view(np.int64).astype(dtype)


CASES = [("c", np.int8, -128, 127), ("C", np.uint8, 0, 255), ("s", np.int16, -3000, 3000), ("S", np.uint16, 0, 65535),
         ("i", np.int32, -5, 200), ("i", np.int32, -2**31, 2**31 - 1), ("I", np.uint32, 2**31, 2**32 - 1),
         ("l", np.int64, 10**12, 10**12 + 60000), ("l", np.int64, -2**63, 2**63 - 1), ("L", np.uint64, 2**63, 2**63 + 100),
         ("tdD", np.int32, 8036, 10561), ("tsu:", np.int64, 1_600_000_000_000_000, 1_600_000_000_000_000 + 10**9)]


@pytest.mark.parametrize("which", [0, 1, 2])
@pytest.mark.parametrize("fmt,dtype,lo,hi", CASES)
def test_fixed_roundtrip_and_width(which, fmt, dtype, lo, hi):
    rng = np.random.default_rng(1)
    for n in (1, 7, 8192):
        arr = rng.integers(lo, hi, n, dtype=dtype, endpoint=True)
        arr[0], arr[-1] = (lo, hi) if n > 1 else (lo, lo)
        base, code, packed = _pack(fmt, arr, which)
        width = arr.dtype.itemsize
        if code == RAW:
            assert (packed[: n * width].view(dtype) == arr).all()
            span = int(arr.max()) - int(arr.min())
            assert span >= 2 ** (8 * (width // 2 if width > 1 else 1)) or width == 1  # raw only when nothing is gained
        else:
            assert code < width
            span = int(arr.max()) - int(arr.min())
            assert span < 2 ** (8 * code) if code else span == 0
            assert (code == 0 or span >= 2 ** (8 * (code // 2))) or code == 1
            got = _decode(base, code, packed, n, dtype)
            assert (got == arr).all()


@pytest.mark.parametrize("which", [0, 1, 2])
def test_decimal128_fits64_and_wide(which):
    rng = np.random.default_rng(2)
    n = 5000
    lo = rng.integers(90000, 10495000, n).astype(np.int64)
    lo[5] = -7  # negative values keep their sign extension
    dec = np.empty(2 * n, dtype=np.int64)
    dec[0::2], dec[1::2] = lo, lo >> 63
    base, code, packed = _pack("d:15,2", dec, which)
    assert code == 4 and base == -7
    assert (_decode(base, code, packed, n, np.int64) == lo).all()
    dec[1001] = 3  # one value beyond 64 bits: the segment travels as it is
    base, code, packed = _pack("d:38,2", dec, which)
    assert code == RAW and (packed[: n * 16].view(np.int64) == dec).all()


@pytest.mark.parametrize("which", [0, 1, 2])
def test_floats_travel_raw(which):
    arr = np.random.default_rng(3).random(1000)
    base, code, packed = _pack("g", arr, which)
    assert code == RAW and (packed[:8000].view(np.float64) == arr).all()
