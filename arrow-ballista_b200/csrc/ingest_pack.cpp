// Host side of the streamed input (ingest.cu): value-range statistics and narrowing of one column segment into its
// wire form. Plain loops the compiler vectorises (plus explicit 512-bit code for Decimal128); this file is compiled three
// times (PACK_NS = pack_avx512 with -mavx512f/bw/vl/dq and PACK_AVX512, pack_avx2 with -mavx2, pack_base without) and
// ingest.cu picks one at run time with __builtin_cpu_supports.
//
// Wire form of a fixed-width integer-like column segment (Int8..Int64, UInt8..UInt64, Date32/64, Time, Timestamp,
// Decimal128 whose values fit 64 bits): frame of reference - `base` = the segment's minimum, then (value - base) as
// unsigned integers of the narrowest of 0 (all values equal), 1, 2, 4 or 8 bytes. Everything is mod-2^64 arithmetic on
// the sign- (or zero-) extended value, so the device's base + delta reproduces every value bit for bit.
#include <stddef.h>
#include <stdint.h>
#include <string.h>

#include "ingest_pack.h"
#ifdef PACK_AVX512
#include <immintrin.h>
#endif

namespace PACK_NS {

template <typename T>
static void range_of(const T* v, int64_t n, int64_t* lo, int64_t* hi) {
  int64_t mn = INT64_MAX, mx = INT64_MIN;
  for (int64_t i = 0; i < n; i++) {
    const int64_t x = (int64_t)v[i];
    mn = x < mn ? x : mn;
    mx = x > mx ? x : mx;
  }
  *lo = mn;
  *hi = mx;
}

// Decimal128 (little-endian i128): range of the low words and whether every high word is the sign extension of its low
// word (the value fits i64)
static bool range_of_dec(const int64_t* v, int64_t n, int64_t* lo, int64_t* hi) {
  int64_t mn = INT64_MAX, mx = INT64_MIN, bad = 0;
  int64_t i = 0;
#ifdef PACK_AVX512
  // Decimal128 columns are most of a TPC-H row (64 of Q1's 78 bytes): explicit 512-bit code, four values per register
  // (even qwords = low words): native 64-bit min / max under a lane mask, high words checked against the sign of the low
  // word next to them. Two registers per iteration keep two independent min / max chains going.
  if (n >= 16) {
    __m512i mn0 = _mm512_set1_epi64(INT64_MAX), mx0 = _mm512_set1_epi64(INT64_MIN), mn1 = mn0, mx1 = mx0, bd = _mm512_setzero_si512();
    for (; i + 8 <= n; i += 8) {
      const __m512i a = _mm512_loadu_si512(v + 2 * i), b = _mm512_loadu_si512(v + 2 * i + 8);
      mn0 = _mm512_mask_min_epi64(mn0, 0x55, mn0, a);
      mx0 = _mm512_mask_max_epi64(mx0, 0x55, mx0, a);
      mn1 = _mm512_mask_min_epi64(mn1, 0x55, mn1, b);
      mx1 = _mm512_mask_max_epi64(mx1, 0x55, mx1, b);
      // lane 2k holds l_k, lane 2k + 1 holds h_k: after the swap lane 2k holds h_k next to sign(l_k)
      const __m512i ha = _mm512_shuffle_epi32(a, (_MM_PERM_ENUM)0x4E), hb = _mm512_shuffle_epi32(b, (_MM_PERM_ENUM)0x4E);
      bd = _mm512_or_si512(bd, _mm512_xor_si512(ha, _mm512_srai_epi64(a, 63)));
      bd = _mm512_or_si512(bd, _mm512_xor_si512(hb, _mm512_srai_epi64(b, 63)));
    }
    mn = _mm512_mask_reduce_min_epi64(0x55, _mm512_mask_min_epi64(mn0, 0x55, mn0, mn1));
    mx = _mm512_mask_reduce_max_epi64(0x55, _mm512_mask_max_epi64(mx0, 0x55, mx0, mx1));
    bad = _mm512_mask_reduce_or_epi64(0x55, bd);
  }
#endif
  for (; i < n; i++) {
    const int64_t l = v[2 * i], h = v[2 * i + 1];
    bad |= h ^ (l >> 63);
    mn = l < mn ? l : mn;
    mx = l > mx ? l : mx;
  }
  *lo = mn;
  *hi = mx;
  return bad == 0;
}

static uint32_t code_for_range(int64_t lo, int64_t hi) {
  const uint64_t r = (uint64_t)hi - (uint64_t)lo;
  if (r == 0) return 0;
  if (r < (1ULL << 8)) return 1;
  if (r < (1ULL << 16)) return 2;
  if (r < (1ULL << 32)) return 4;
  return 8;
}

void fixed_stats(const void* values, int64_t n, int width, int is_signed, int is_decimal, int is_float,
                 bgpu_pack::FixedStats* out) {
  out->base = 0;
  out->code = bgpu_pack::kRaw;
  out->lo = out->hi = 0;
  out->has_range = 0;
  if (is_float || n == 0) {
    if (n == 0) out->code = 0;
    return;
  }
  int64_t lo = 0, hi = 0;
  if (is_decimal) {
    if (!range_of_dec((const int64_t*)values, n, &lo, &hi)) return;  // raw 16 bytes
  } else {
    switch (width) {
      case 1: is_signed ? range_of((const int8_t*)values, n, &lo, &hi) : range_of((const uint8_t*)values, n, &lo, &hi); break;
      case 2: is_signed ? range_of((const int16_t*)values, n, &lo, &hi) : range_of((const uint16_t*)values, n, &lo, &hi); break;
      case 4: is_signed ? range_of((const int32_t*)values, n, &lo, &hi) : range_of((const uint32_t*)values, n, &lo, &hi); break;
      case 8: range_of((const int64_t*)values, n, &lo, &hi); break;  // UInt64 as its two's-complement reading: still exact
      default: return;
    }
  }
  const uint32_t code = code_for_range(lo, hi);
  out->lo = lo;
  out->hi = hi;
  out->has_range = 1;
  if ((int)code >= width) return;  // nothing gained: raw copy
  out->base = lo;
  out->code = code;
}

template <typename Tin, typename Tout>
static void pack_loop(const Tin* in, int64_t n, int64_t stride, int64_t base, Tout* out) {
  int64_t i = 0;
#ifdef PACK_AVX512
  if (sizeof(Tin) == 8 && stride == 2 && sizeof(Tout) <= 4) {  // Decimal128 low words -> 1 / 2 / 4-byte deltas, 8 per step
    const __m512i vb = _mm512_set1_epi64(base), even = _mm512_set_epi64(14, 12, 10, 8, 6, 4, 2, 0);
    const int64_t* p = (const int64_t*)in;
    for (; i + 8 <= n; i += 8) {
      const __m512i a = _mm512_loadu_si512(p + 2 * i), b = _mm512_loadu_si512(p + 2 * i + 8);
      const __m512i d = _mm512_sub_epi64(_mm512_permutex2var_epi64(a, even, b), vb);
      if (sizeof(Tout) == 1) _mm_storel_epi64((__m128i*)(out + i), _mm512_cvtepi64_epi8(d));
      else if (sizeof(Tout) == 2) _mm_storeu_si128((__m128i*)(out + i), _mm512_cvtepi64_epi16(d));
      else _mm256_storeu_si256((__m256i*)(out + i), _mm512_cvtepi64_epi32(d));
    }
  }
#endif
  for (; i < n; i++) out[i] = (Tout)((uint64_t)(int64_t)in[i * stride] - (uint64_t)base);
}
template <typename Tin>
static void pack_to(const Tin* in, int64_t n, int64_t stride, int64_t base, uint32_t code, void* out) {
  switch (code) {
    case 1: pack_loop(in, n, stride, base, (uint8_t*)out); break;
    case 2: pack_loop(in, n, stride, base, (uint16_t*)out); break;
    case 4: pack_loop(in, n, stride, base, (uint32_t*)out); break;
    case 8: pack_loop(in, n, stride, base, (uint64_t*)out); break;
    default: break;
  }
}

// One pass instead of range + pack: pack against a GIVEN base and width (the column's running range over the segments
// seen so far) and check, while packing, that every value fits - every delta below 2^(8 * code), and for Decimal128 every
// high word the sign extension of its low word. Returns false when some value does not fit (the caller then takes the
// exact two-pass route); `out` holds garbage in that case.
template <typename Tin, typename Tout>
static bool pack_checked_loop(const Tin* in, int64_t n, int64_t stride, int64_t base, Tout* out) {
  uint64_t bad = 0;
  int64_t i = 0;
#ifdef PACK_AVX512
  if (sizeof(Tin) == 8 && stride == 2) {  // Decimal128: 8 values per step
    const __m512i vb = _mm512_set1_epi64(base), even = _mm512_set_epi64(14, 12, 10, 8, 6, 4, 2, 0);
    const __m512i odd = _mm512_set_epi64(15, 13, 11, 9, 7, 5, 3, 1);
    __m512i vbad = _mm512_setzero_si512();
    const int64_t* p = (const int64_t*)in;
    for (; i + 8 <= n; i += 8) {
      const __m512i a = _mm512_loadu_si512(p + 2 * i), b = _mm512_loadu_si512(p + 2 * i + 8);
      const __m512i l = _mm512_permutex2var_epi64(a, even, b), h = _mm512_permutex2var_epi64(a, odd, b);
      const __m512i d = _mm512_sub_epi64(l, vb);
      vbad = _mm512_or_si512(vbad, _mm512_srli_epi64(d, 8 * (int)sizeof(Tout)));
      vbad = _mm512_or_si512(vbad, _mm512_xor_si512(h, _mm512_srai_epi64(l, 63)));
      if (sizeof(Tout) == 1) _mm_storel_epi64((__m128i*)(out + i), _mm512_cvtepi64_epi8(d));
      else if (sizeof(Tout) == 2) _mm_storeu_si128((__m128i*)(out + i), _mm512_cvtepi64_epi16(d));
      else _mm256_storeu_si256((__m256i*)(out + i), _mm512_cvtepi64_epi32(d));
    }
    bad = (uint64_t)_mm512_reduce_or_epi64(vbad);
  }
#endif
  for (; i < n; i++) {
    const uint64_t d = (uint64_t)(int64_t)in[i * stride] - (uint64_t)base;
    bad |= d >> (8 * sizeof(Tout));
    if (stride == 2) bad |= (uint64_t)((int64_t)in[i * stride + 1] ^ ((int64_t)in[i * stride] >> 63));
    out[i] = (Tout)d;
  }
  return bad == 0;
}
template <typename Tin>
static bool pack_checked_to(const Tin* in, int64_t n, int64_t stride, int64_t base, uint32_t code, void* out) {
  switch (code) {
    case 1: return pack_checked_loop(in, n, stride, base, (uint8_t*)out);
    case 2: return pack_checked_loop(in, n, stride, base, (uint16_t*)out);
    case 4: return pack_checked_loop(in, n, stride, base, (uint32_t*)out);
    default: return false;
  }
}
int fixed_pack_checked(const void* values, int64_t n, int width, int is_signed, int is_decimal, int64_t base, uint32_t code,
                       void* out) {
  if (code != 1 && code != 2 && code != 4) return 0;
  if (is_decimal) return pack_checked_to((const int64_t*)values, n, 2, base, code, out);
  if ((int)code >= width) return 0;
  switch (width) {
    case 2: return is_signed ? pack_checked_to((const int16_t*)values, n, 1, base, code, out) : pack_checked_to((const uint16_t*)values, n, 1, base, code, out);
    case 4: return is_signed ? pack_checked_to((const int32_t*)values, n, 1, base, code, out) : pack_checked_to((const uint32_t*)values, n, 1, base, code, out);
    case 8: return pack_checked_to((const int64_t*)values, n, 1, base, code, out);
    default: return 0;
  }
}

void fixed_pack(const void* values, int64_t n, int width, int is_signed, int is_decimal, int64_t base, uint32_t code,
                void* out) {
  if (code == bgpu_pack::kRaw) {
    memcpy(out, values, (size_t)n * (size_t)width);
    return;
  }
  if (code == 0) return;
  if (is_decimal) { pack_to((const int64_t*)values, n, 2, base, code, out); return; }
  switch (width) {
    case 2: is_signed ? pack_to((const int16_t*)values, n, 1, base, code, out) : pack_to((const uint16_t*)values, n, 1, base, code, out); break;
    case 4: is_signed ? pack_to((const int32_t*)values, n, 1, base, code, out) : pack_to((const uint32_t*)values, n, 1, base, code, out); break;
    case 8: pack_to((const int64_t*)values, n, 1, base, code, out); break;
    default: break;
  }
}

// utf8 / binary: lengths from the offsets
void utf8_stats(const int32_t* offs, int64_t n, int32_t* min_len, int32_t* max_len, int* monotonic) {
  int32_t mn = INT32_MAX, mx = INT32_MIN;
  for (int64_t i = 0; i < n; i++) {
    const int32_t l = offs[i + 1] - offs[i];
    mn = l < mn ? l : mn;
    mx = l > mx ? l : mx;
  }
  if (n == 0) mn = mx = 0;
  *min_len = mn;
  *max_len = mx;
  *monotonic = mn >= 0;
}
void utf8_pack_lens(const int32_t* offs, int64_t n, uint32_t code, void* out) {
  switch (code) {
    case 1: { uint8_t* o = (uint8_t*)out; for (int64_t i = 0; i < n; i++) o[i] = (uint8_t)(offs[i + 1] - offs[i]); } break;
    case 2: { uint16_t* o = (uint16_t*)out; for (int64_t i = 0; i < n; i++) o[i] = (uint16_t)(offs[i + 1] - offs[i]); } break;
    case 4: { uint32_t* o = (uint32_t*)out; for (int64_t i = 0; i < n; i++) o[i] = (uint32_t)(offs[i + 1] - offs[i]); } break;
    default: break;
  }
}

// n bits starting at bit `off` of src -> dst aligned at bit 0 (trailing bits of the last byte are zero)
void bits_realign(const uint8_t* src, int64_t off, int64_t n, uint8_t* dst) {
  const size_t bytes = (size_t)((n + 7) / 8);
  if (bytes == 0) return;
  const int sh = (int)(off & 7);
  const uint8_t* s = src + (off >> 3);
  if (sh == 0) {
    memcpy(dst, s, bytes);
  } else {
    const size_t src_bytes = (size_t)((sh + n + 7) / 8);
    for (size_t i = 0; i < bytes; i++) {
      const uint32_t a = s[i], b = i + 1 < src_bytes ? s[i + 1] : 0;
      dst[i] = (uint8_t)((a >> sh) | (b << (8 - sh)));
    }
  }
  if (n & 7) dst[bytes - 1] &= (uint8_t)((1u << (n & 7)) - 1);
}

const bgpu_pack::Ops ops = {fixed_stats, fixed_pack, utf8_stats, utf8_pack_lens, bits_realign, fixed_pack_checked};
}  // namespace PACK_NS
