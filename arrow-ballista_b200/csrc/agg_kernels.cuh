// Ahead-of-time kernels of the grouped aggregate: the small plan-independent ones (table rebuild, commit of the fused
// path's per-CTA slabs, emit). The hot kernels (fused / intern / accumulate) are specialised per plan at run time:
// agg_jit_prelude.cuh (kernel templates) + agg_codegen.cu (the generated `Plan`).
#pragma once
#include "agg_device.cuh"

namespace bgpu {

constexpr int kAggThreads = 256;
constexpr uint32_t kNoGroup = 0xffffffffu;  // row dropped by the filter

// Rebuild the entry table from the dense keys (after growing).
__global__ void __launch_bounds__(kAggThreads) rehash_kernel(const __grid_constant__ GroupTable t, uint32_t n, int words) {
  for (uint32_t g = blockIdx.x * blockDim.x + threadIdx.x; g < n; g += gridDim.x * blockDim.x) {
    PackedKey k{t.key0[g], words > 1 ? t.key1[g] : 0, t.kmeta[g]};
    const uint64_t h = key_hash(k);
    const unsigned long long tag = (h >> 32) << 32;
    uint32_t i = (uint32_t)h & t.mask;
    while (atomicCAS(&t.entries[i], 0ULL, tag | (unsigned long long)(g + 2u)) != 0ULL) i = (i + 1) & t.mask;
  }
}

// Scratch slab of fused-path CTA b: u64[g_cap * W] partial accumulators, then u32[g_cap] seen masks (padded to 8 B).
__host__ __device__ __forceinline__ size_t fused_slab_words(uint32_t g_cap, int W) {
  return (size_t)g_cap * W + ((size_t)g_cap + 1) / 2;
}

// Fold the per-CTA slabs into the dense accumulator arrays: one WARP per (group, accumulator); lanes stride over the
// slabs and the partials are combined by shuffles in a fixed order (Float64 sums are reproducible for a fixed grid).
__global__ void __launch_bounds__(kAggThreads)
    commit_kernel(const __grid_constant__ AccPlan plan, const uint64_t* __restrict__ scratch, uint32_t n_ctas,
                  uint32_t g_cap, uint32_t n_groups) {
  const int W = plan.total_words;
  const size_t slab = fused_slab_words(g_cap, W);
  const uint32_t items = n_groups * (uint32_t)plan.n;
  const uint32_t lane = threadIdx.x & 31, warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, n_warps = (gridDim.x * blockDim.x) >> 5;
  for (uint32_t item = warp; item < items; item += n_warps) {
    const uint32_t grp = item / plan.n;
    const AccK& a = plan.a[item % plan.n];
    const AccK& src = a.alias >= 0 ? plan.a[a.alias] : a;
    uint64_t lo = 0, hi = 0;
    double fsum = 0.0;
    for (uint32_t b = lane; b < n_ctas; b += 32) {
      const uint64_t* p = scratch + (size_t)b * slab + (size_t)grp * W + src.word0;
      if (src.op == OP_ADD_F64) {
        fsum += __longlong_as_double((long long)p[0]);
      } else {
        const uint64_t x = p[0];
        lo += x;
        hi += (lo < x) + (src.op == OP_ADD_I128 ? p[1] : 0ULL);
      }
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
      fsum += __shfl_down_sync(0xffffffffu, fsum, d);
      const uint64_t olo = __shfl_down_sync(0xffffffffu, lo, d), ohi = __shfl_down_sync(0xffffffffu, hi, d);
      lo += olo;
      hi += ohi + (lo < olo);
    }
    if (lane == 0) {
      uint64_t* o = a.out + (size_t)grp * a.words;
      if (a.op == OP_ADD_F64) {
        *(double*)o += fsum;
      } else {
        const uint64_t cur = o[0], nlo = cur + lo;
        o[0] = nlo;
        if (a.words > 1) o[1] += hi + (nlo < cur);
      }
    }
  }
  if (plan.seen) {
    for (uint32_t grp = warp; grp < n_groups; grp += n_warps) {
      uint32_t m = 0;
      for (uint32_t b = lane; b < n_ctas; b += 32) m |= ((const uint32_t*)(scratch + (size_t)b * slab + (size_t)g_cap * W))[grp];
#pragma unroll
      for (int d = 16; d > 0; d >>= 1) m |= __shfl_down_sync(0xffffffffu, m, d);
      if (lane == 0) {
        // aliased accumulators saw whatever their source saw
        for (int i = 0; i < plan.n; i++)
          if (plan.a[i].alias >= 0 && ((m >> plan.a[plan.a[i].alias].seen_bit) & 1)) m |= 1u << plan.a[i].seen_bit;
        if (m) plan.seen[grp] |= m;
      }
    }
  }
}

// ---- emit kernels -----------------------------------------------------------------------------------------------
// Unpack one fixed-width / bool / dictionary key column of the dense key arrays into an Arrow column.
__global__ void __launch_bounds__(kAggThreads)
    unpack_fixed_kernel(const uint64_t* __restrict__ key0, const uint64_t* __restrict__ key1,
                        const uint32_t* __restrict__ kmeta, uint32_t n, int col, int kind, int width, int byte_off,
                        void* __restrict__ out_values, uint32_t* __restrict__ out_valid) {
  const uint32_t n_pad = (n + 31) & ~31u;
  for (uint32_t g = blockIdx.x * blockDim.x + threadIdx.x; g < n_pad; g += gridDim.x * blockDim.x) {
    bool valid = false, bit = false;
    if (g < n) {
      valid = !((kmeta[g] >> col) & 1);
      const uint64_t k0 = key0[g], k1 = key1 ? key1[g] : 0;
      const uint64_t word = byte_off < 8 ? (k0 >> (8 * byte_off)) : (k1 >> (8 * (byte_off - 8)));
      bit = (word & 1) != 0;
      if (kind == K_BOOL) {
        // bit-packed output: written by the ballot below
      } else if (width == 16) {
        ((ulonglong2*)out_values)[g] = make_ulonglong2(k0, k1);
      } else if (width == 8) ((uint64_t*)out_values)[g] = word;
      else if (width == 4) ((uint32_t*)out_values)[g] = (uint32_t)word;
      else if (width == 2) ((uint16_t*)out_values)[g] = (uint16_t)word;
      else ((uint8_t*)out_values)[g] = (uint8_t)word;
    }
    if (kind == K_BOOL) {
      const uint32_t bits = __ballot_sync(0xffffffffu, bit);
      if ((threadIdx.x & 31) == 0) ((uint32_t*)out_values)[g >> 5] = bits;
    }
    if (out_valid) {
      const uint32_t vb = __ballot_sync(0xffffffffu, valid);
      if ((threadIdx.x & 31) == 0) out_valid[g >> 5] = vb;
    }
  }
}

// utf8 key column: lengths (pass 0) or bytes (pass 1, offsets known)
__global__ void __launch_bounds__(kAggThreads)
    unpack_utf8_kernel(const uint64_t* __restrict__ key0, const uint64_t* __restrict__ key1,
                       const uint32_t* __restrict__ kmeta, uint32_t n, int col, int byte_off, int pass,
                       int32_t* __restrict__ lens_or_offsets, uint8_t* __restrict__ out_data,
                       uint32_t* __restrict__ out_valid) {
  const uint32_t n_pad = (n + 31) & ~31u;
  for (uint32_t g = blockIdx.x * blockDim.x + threadIdx.x; g < n_pad; g += gridDim.x * blockDim.x) {
    bool valid = false;
    if (g < n) {
      valid = !((kmeta[g] >> col) & 1);
      const uint64_t k0 = key0[g], k1 = key1 ? key1[g] : 0;
      auto byte_at = [&](int off) -> uint32_t {
        return (uint32_t)((off < 8 ? (k0 >> (8 * off)) : (k1 >> (8 * (off - 8)))) & 0xff);
      };
      const int len = valid ? (int)byte_at(byte_off) : 0;
      if (pass == 0) {
        lens_or_offsets[g] = len;
      } else {
        const int32_t b = lens_or_offsets[g];
        for (int i = 0; i < len; i++) out_data[b + i] = (uint8_t)byte_at(byte_off + 1 + i);
      }
    }
    if (pass == 0 && out_valid) {
      const uint32_t vb = __ballot_sync(0xffffffffu, valid);
      if ((threadIdx.x & 31) == 0) out_valid[g >> 5] = vb;
    }
  }
}

// validity bitmap of an accumulator from the seen masks
__global__ void __launch_bounds__(kAggThreads)
    seen_bitmap_kernel(const uint32_t* __restrict__ seen, uint32_t n, int bit, uint32_t* __restrict__ out_valid) {
  const uint32_t n_pad = (n + 31) & ~31u;
  for (uint32_t g = blockIdx.x * blockDim.x + threadIdx.x; g < n_pad; g += gridDim.x * blockDim.x) {
    const uint32_t vb = __ballot_sync(0xffffffffu, g < n && ((seen[g] >> bit) & 1));
    if ((threadIdx.x & 31) == 0) out_valid[g >> 5] = vb;
  }
}

__global__ void __launch_bounds__(kAggThreads) fill_u32_kernel(uint32_t* p, uint32_t n, uint32_t v) {
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) p[i] = v;
}

// AvgGroupsAccumulator::evaluate. Float64: sum / count. Decimal128: (sum * 10^scale_up) / count, truncating;
// an overflowing scaled multiply is an error in the reference -> flags[2].
__global__ void __launch_bounds__(kAggThreads)
    avg_evaluate_kernel(const uint64_t* __restrict__ sum, const uint64_t* __restrict__ count, uint32_t n, int is_decimal,
                        int scale_up, uint64_t* __restrict__ out, uint32_t* __restrict__ out_valid, int32_t* flags) {
  const uint32_t n_pad = (n + 31) & ~31u;
  for (uint32_t g = blockIdx.x * blockDim.x + threadIdx.x; g < n_pad; g += gridDim.x * blockDim.x) {
    bool valid = false;
    if (g < n) {
      const uint64_t c = count[g];
      valid = c != 0;
      if (is_decimal) {
        __int128 s = (__int128)(((unsigned __int128)sum[2 * g + 1] << 64) | sum[2 * g]);
        __int128 q = 0;
        if (valid) {
          bool ovf = false;
          for (int i = 0; i < scale_up; i++) {
            const __int128 lim = (((__int128)1 << 126) - 1 + ((__int128)1 << 126)) / 10;  // i128::MAX / 10
            if (s > lim || s < -lim) ovf = true;
            s *= 10;
          }
          if (ovf) atomicOr(&flags[2], 1);
          q = s / (__int128)c;
        }
        out[2 * g] = (uint64_t)q;
        out[2 * g + 1] = (uint64_t)((unsigned __int128)q >> 64);
      } else {
        const double s = __longlong_as_double((long long)sum[g]);
        out[g] = (uint64_t)__double_as_longlong(valid ? s / (double)c : 0.0);
      }
    }
    const uint32_t vb = __ballot_sync(0xffffffffu, valid);
    if ((threadIdx.x & 31) == 0) out_valid[g >> 5] = vb;
  }
}
}  // namespace bgpu
