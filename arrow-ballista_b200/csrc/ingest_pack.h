// Host-side column narrowing of the streamed input (see ingest_pack.cpp / ingest.cu).
#pragma once
#include <stdint.h>

namespace bgpu_pack {
constexpr uint32_t kRaw = 255;  // wire code: the segment's values as they are (full width)
struct FixedStats {
  int64_t base;   // frame-of-reference base (the segment's minimum)
  uint32_t code;  // bytes per packed value: 0 (all equal to base), 1, 2, 4, 8, or kRaw
  int64_t lo, hi;     // the segment's exact value range (valid when has_range: integers, and Decimal128 that fit 64 bits)
  int32_t has_range;
};
struct Ops {
  void (*fixed_stats)(const void* values, int64_t n, int width, int is_signed, int is_decimal, int is_float, FixedStats* out);
  void (*fixed_pack)(const void* values, int64_t n, int width, int is_signed, int is_decimal, int64_t base, uint32_t code, void* out);
  void (*utf8_stats)(const int32_t* offs, int64_t n, int32_t* min_len, int32_t* max_len, int* monotonic);
  void (*utf8_pack_lens)(const int32_t* offs, int64_t n, uint32_t code, void* out);
  void (*bits_realign)(const uint8_t* src, int64_t bit_off, int64_t n, uint8_t* dst);
  // one pass: pack against a given (base, code in {1, 2, 4}) and verify that every value fits; 0 = some value did not
  int (*fixed_pack_checked)(const void* values, int64_t n, int width, int is_signed, int is_decimal, int64_t base, uint32_t code,
                            void* out);
};
}  // namespace bgpu_pack
namespace pack_avx2 { extern const bgpu_pack::Ops ops; }
namespace pack_base { extern const bgpu_pack::Ops ops; }
namespace pack_avx512 { extern const bgpu_pack::Ops ops; }
