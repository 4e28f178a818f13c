// Device-side pieces shared by the ahead-of-time aggregate kernels (agg_kernels.cuh). The run-time specialised
// kernels carry their own copy of the table code in agg_jit_prelude.cuh; key_hash must stay identical in both.
#pragma once
#include "agg_plan.h"

namespace bgpu {

__device__ __forceinline__ uint64_t mix64(uint64_t x) {  // murmur3 fmix64: table hash (not the partition hash)
  x ^= x >> 33; x *= 0xff51afd7ed558ccdULL; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ULL; x ^= x >> 33;
  return x;
}

struct PackedKey { uint64_t k0, k1; uint32_t meta; };

__device__ __forceinline__ uint64_t key_hash(const PackedKey& k) {
  return mix64(k.k0 ^ mix64(k.k1 + 0x9e3779b97f4a7c15ULL * (k.meta + 1)));
}

// slot accessors (GroupTable layout, agg_plan.h)
__device__ __forceinline__ uint64_t* slot_ptr(const GroupTable& t, uint32_t i) { return t.slots + (size_t)i * t.slot_words; }
__device__ __forceinline__ int slot_sm_word(const GroupTable& t) { return t.slot_words == 2 ? 1 : 2; }

}  // namespace bgpu
