// Device-side create_hashes: ahash 0.8 fallback hasher with RandomState::with_seeds(0,0,0,0), as used by
// DataFusion's BatchPartitioner (reference call site ballista/core/src/execution_plans/shuffle_writer.rs:336-346;
// algorithm restated in SURVEY.md A.1 and pinned structurally in oracle/evidence/README.md).
// This is the product's own implementation; the oracle has an independent one in oracle/ahash_fallback.h.
#pragma once
#include <stdint.h>

#include "internal.h"

namespace bgpu {

constexpr uint64_t kMultiple = 6364136223846793005ULL;
// with_seeds(0,0,0,0): k_i = 0 ^ PI2[i]; AHasher{buffer: k1, pad: k0, extra_keys: [k2, k3]}
constexpr uint64_t kPad = 0x452821e638d01377ULL;
constexpr uint64_t kBuffer = 0xbe5466cf34e90c6cULL;
constexpr uint64_t kExtra0 = 0xc0ac29b7c97c50ddULL;
constexpr uint64_t kExtra1 = 0x3f84d5b5b5470917ULL;

__device__ __forceinline__ uint64_t folded_multiply(uint64_t s, uint64_t by) {
  return (s * by) ^ __umul64hi(s, by);
}
__device__ __forceinline__ uint64_t rotl64(uint64_t x, unsigned r) {
  r &= 63;
  return (x << r) | (x >> ((64 - r) & 63));
}
__device__ __forceinline__ uint64_t ah_update(uint64_t buffer, uint64_t x) {
  return folded_multiply(x ^ buffer, kMultiple);
}
__device__ __forceinline__ uint64_t ah_large_update(uint64_t buffer, uint64_t lo, uint64_t hi) {
  uint64_t combined = folded_multiply(lo ^ kExtra0, hi ^ kExtra1);
  return rotl64((buffer + kPad) ^ combined, 23);
}
__device__ __forceinline__ uint64_t ah_finish(uint64_t buffer) {
  return rotl64(folded_multiply(buffer, kPad), (unsigned)(buffer & 63));
}
// hash_one of an integer of width <= 8 bytes (zero-extended bit pattern), bool, float bits
__device__ __forceinline__ uint64_t hash_u64(uint64_t x) { return ah_finish(ah_update(kBuffer, x)); }
// hash_one of i128 (Decimal128): write_u128 -> large_update(lo, hi)
__device__ __forceinline__ uint64_t hash_u128(uint64_t lo, uint64_t hi) {
  return ah_finish(ah_large_update(kBuffer, lo, hi));
}

__device__ __forceinline__ uint64_t load_le(const uint8_t* p, int n) {  // unaligned little-endian read, n <= 8
  uint64_t v = 0;
  for (int i = 0; i < n; i++) v |= (uint64_t)p[i] << (8 * i);
  return v;
}

// Hasher::write(bytes)
__device__ __forceinline__ uint64_t ah_write(uint64_t buffer, const uint8_t* d, uint64_t len) {
  buffer = (buffer + len) * kMultiple;
  if (len > 8) {
    if (len > 16) {
      buffer = ah_large_update(buffer, load_le(d + len - 16, 8), load_le(d + len - 8, 8));
      while (len > 16) {
        buffer = ah_large_update(buffer, load_le(d, 8), load_le(d + 8, 8));
        d += 16;
        len -= 16;
      }
    } else {
      buffer = ah_large_update(buffer, load_le(d, 8), load_le(d + len - 8, 8));
    }
  } else if (len >= 4) {
    buffer = ah_large_update(buffer, load_le(d, 4), load_le(d + len - 4, 4));
  } else if (len >= 2) {
    buffer = ah_large_update(buffer, load_le(d, 2), (uint64_t)d[len - 1]);
  } else if (len == 1) {
    buffer = ah_large_update(buffer, (uint64_t)d[0], (uint64_t)d[0]);
  } else {
    buffer = ah_large_update(buffer, 0, 0);
  }
  return buffer;
}
// <str as Hash>::hash: write(bytes); write_u8(0xff)
__device__ __forceinline__ uint64_t hash_str(const uint8_t* d, uint64_t len) {
  return ah_finish(ah_update(ah_write(kBuffer, d, len), 0xff));
}
// <[u8] as Hash>::hash: write_usize(len); write(bytes)
__device__ __forceinline__ uint64_t hash_bytes(const uint8_t* d, uint64_t len) {
  return ah_finish(ah_write(ah_update(kBuffer, len), d, len));
}
// datafusion-common hash_utils.rs combine_hashes
__device__ __forceinline__ uint64_t combine_hashes(uint64_t l, uint64_t r) { return (17ULL * 37ULL + l) * 37ULL + r; }

__device__ __forceinline__ bool bit_is_set(const uint8_t* bits, int64_t i) { return (bits[i >> 3] >> (i & 7)) & 1; }

// hash_one of row r of one key column (the caller has checked validity)
__device__ __forceinline__ uint64_t hash_value(const KCol& c, int64_t r) {
  switch (c.kind) {
    case K_FIXED:
      switch (c.width) {
        case 8: return hash_u64(((const uint64_t*)c.values)[r]);
        case 4: return hash_u64(((const uint32_t*)c.values)[r]);
        case 16: {
          const ulonglong2 v = ((const ulonglong2*)c.values)[r];
          return hash_u128(v.x, v.y);
        }
        case 2: return hash_u64(((const uint16_t*)c.values)[r]);
        default: return hash_u64(((const uint8_t*)c.values)[r]);
      }
    case K_BOOL: return hash_u64(bit_is_set((const uint8_t*)c.values, r) ? 1 : 0);
    case K_UTF8: {
      const int32_t b = c.offsets[r], e = c.offsets[r + 1];
      return hash_str((const uint8_t*)c.values + b, (uint64_t)(e - b));
    }
    case K_BINARY: {
      const int32_t b = c.offsets[r], e = c.offsets[r + 1];
      return hash_bytes((const uint8_t*)c.values + b, (uint64_t)(e - b));
    }
    default:  // K_DICT32: hash of the dictionary value, computed once per dictionary entry
      return c.dict_hashes[((const int32_t*)c.values)[r]];
  }
}

constexpr int kMaxKeys = 8;
struct KeySet {
  KCol cols[kMaxKeys];
  int32_t n;
  int32_t fast64;  // 1: exactly one key column, K_FIXED width 8, no validity
};

// create_hashes for row r over all key columns: column 0 sets, columns >= 1 combine; null rows are skipped.
static __device__ __noinline__ uint64_t row_hash_generic(const KeySet& ks, int64_t r) {
  uint64_t h = 0;
#pragma unroll 1
  for (int c = 0; c < ks.n; c++) {
    const KCol& k = ks.cols[c];
    if (k.validity && !bit_is_set(k.validity, r)) continue;
    const uint64_t vh = hash_value(k, r);
    h = (c == 0) ? vh : combine_hashes(vh, h);
  }
  return h;
}
// Fast path: a single non-null 8-byte key (Int64/UInt64/Float64/Timestamp) is hashed inline.
__device__ __forceinline__ uint64_t row_hash(const KeySet& ks, int64_t r) {
  if (ks.fast64) return hash_u64(((const uint64_t*)ks.cols[0].values)[r]);
  return row_hash_generic(ks, r);
}

// hash % n without a 64-bit division on the hot path.
struct Modulo {
  uint32_t n;
  uint32_t mask;      // n-1 when n is a power of two, else 0 (and n != 1)
  uint32_t pow2;      // 1 when n is a power of two
  uint32_t small;     // 1 when n < 65536 (three 32-bit fastmods)
  uint64_t magic;     // ceil(2^64 / n) for Lemire's 32-bit fastmod
  uint32_t two32_mod; // 2^32 % n
};
__device__ __forceinline__ uint32_t fastmod32(uint32_t a, const Modulo& m) {
  const uint64_t low = m.magic * a;
  return (uint32_t)__umul64hi(low, m.n);
}
__device__ __forceinline__ uint32_t partition_of(uint64_t h, const Modulo& m) {
  if (m.pow2) return (uint32_t)h & m.mask;
  if (m.small) {
    const uint32_t r1 = fastmod32((uint32_t)(h >> 32), m), r2 = fastmod32((uint32_t)h, m);
    return fastmod32(r1 * m.two32_mod + r2, m);  // < n*n + n <= 2^32
  }
  return (uint32_t)(h % m.n);
}

}  // namespace bgpu
