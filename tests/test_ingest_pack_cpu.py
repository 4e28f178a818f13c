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


def _pack_checked(fmt, arr, which, base, code):
    n = len(arr) if not fmt.startswith("d") else len(arr) // 2
    out = np.zeros(max(n, 1) * 16, dtype=np.uint8)
    rc = lib().bgpu_debug_pack_checked_host(fmt.encode(), arr.ctypes.data_as(C.c_void_p), n, which, base, code,
                                            out.ctypes.data_as(C.c_void_p))
    if rc == -2 and which in (0, 2):
        pytest.skip("host without AVX2 / AVX-512")
    assert rc in (0, 1), rc
    return rc == 1, out


@pytest.mark.parametrize("which", [0, 1, 2])
@pytest.mark.parametrize("fmt,dtype", [("l", np.int64), ("i", np.int32), ("I", np.uint32), ("s", np.int16), ("tdD", np.int32)])
def test_one_pass_pack_accepts_what_fits_and_nothing_else(which, fmt, dtype):
    """The one-pass route packs against a GIVEN (base, width) and verifies on the way: when every value fits the packed
    bytes decode to the input bit for bit; one value below the base or beyond the width makes it say no."""
    rng = np.random.default_rng(5)
    for n in (1, 7, 8, 63, 1000, 8192):
        for code, span in ((1, 200), (2, 60_000), (4, 2_000_000_000)):
            if code >= np.dtype(dtype).itemsize:
                continue
            base = int(rng.integers(-1000, 1000)) if np.dtype(dtype).kind == "i" else int(rng.integers(0, 1000))
            hi = min(base + span, np.iinfo(dtype).max)
            arr = rng.integers(base, hi, n, endpoint=True).astype(dtype)
            ok, packed = _pack_checked(fmt, arr, which, base, code)
            assert ok
            got = _decode(base, code, packed, n, dtype)
            assert np.array_equal(got, arr)
            # one value that does not fit: below the base, or beyond the width
            for bad in ([base - 1] if base - 1 >= np.iinfo(dtype).min else []) + ([base + (1 << (8 * code))] if base + (1 << (8 * code)) <= np.iinfo(dtype).max else []):
                arr2 = arr.copy()
                arr2[int(rng.integers(0, n))] = bad
                ok2, _ = _pack_checked(fmt, arr2, which, base, code)
                assert not ok2, (n, code, base, bad)


@pytest.mark.parametrize("which", [0, 1, 2])
def test_one_pass_pack_of_decimal128(which):
    rng = np.random.default_rng(6)
    for n in (1, 8, 9, 4000):
        lo = rng.integers(-5000, 5000, n).astype(np.int64)
        dec = np.empty(2 * n, dtype=np.int64)
        dec[0::2], dec[1::2] = lo, lo >> 63
        base = int(lo.min())
        ok, packed = _pack_checked("d:15,2", dec, which, base, 2)
        assert ok and np.array_equal(_decode(base, 2, packed, n, np.int64), lo)
        assert not _pack_checked("d:15,2", dec, which, base + 1, 2)[0]      # the minimum is below the base
        assert not _pack_checked("d:15,2", dec, which, base, 1)[0] or lo.max() - base < 256
        wide = dec.copy()
        wide[2 * int(rng.integers(0, n)) + 1] = 7                            # a value beyond 64 bits
        assert not _pack_checked("d:38,2", wide, which, base, 2)[0]
