// bgpu_batch_sort: a stable multi-column sort of a device batch - the ordered fan-in of CoalesceTasksExec.
//
// When the child of a task group is ordered the reference merges its partitions' sorted streams with `streaming_merge`
// (/root/reference/ballista/core/src/execution_plans/coalesce_tasks.rs:148-173; ties go to the lower stream index). On
// the device that is: concatenate the partitions' batches in stream order (bgpu_batch_concat) and sort STABLY by the sort
// expressions - for sorted inputs the two give the same rows in the same order. The sort is least-significant-key-first
// passes of cub's (stable) radix sort over a row permutation: one 64-bit pass per fixed-width key column (two for
// Decimal128), ceil(max length / 8) + 1 passes per utf8 column, one 1-bit pass per nullable column for NULLS FIRST / LAST;
// then one gather of every column through the permutation (`take`).
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>

#include "internal.h"

namespace bgpu {
namespace {
constexpr int kThreads = 256;
int grid_for(bgpu_ctx* ctx, int64_t n) {
  return (int)std::max<int64_t>(1, std::min<int64_t>(ceil_div64(n, kThreads), (int64_t)ctx->sm_count * 8));
}
__device__ __forceinline__ bool bit_at(const uint8_t* bits, int64_t i) { return (bits[i >> 3] >> (i & 7)) & 1; }

enum KeyMode : int32_t { M_SIGNED = 0, M_UNSIGNED = 1, M_FLOAT64 = 2, M_FLOAT32 = 3, M_DEC_LO = 4, M_DEC_HI = 5, M_BOOL = 6,
                         M_NULLFLAG = 7, M_STR_LEN = 8, M_STR_CHUNK = 9 };
struct KeyArgs {
  const void* values;
  const int32_t* offsets;
  const uint8_t* validity;
  int32_t width, mode, chunk, descending, nulls_first;
};

// key[i] = order-preserving 64-bit image of the chosen piece of the key column at row perm[i] (0 for null rows)
__global__ void __launch_bounds__(kThreads)
    sort_key_kernel(const __grid_constant__ KeyArgs a, const uint32_t* __restrict__ perm, int64_t n, uint64_t* __restrict__ key) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = perm[i];
    const bool valid = !a.validity || bit_at(a.validity, r);
    uint64_t k = 0;
    if (a.mode == M_NULLFLAG) {
      key[i] = valid ? (a.nulls_first ? 1 : 0) : (a.nulls_first ? 0 : 1);
      continue;
    }
    if (valid) {
      switch (a.mode) {
        case M_SIGNED: {
          int64_t v;
          switch (a.width) {
            case 1: v = ((const int8_t*)a.values)[r]; break;
            case 2: v = ((const int16_t*)a.values)[r]; break;
            case 4: v = ((const int32_t*)a.values)[r]; break;
            default: v = ((const int64_t*)a.values)[r]; break;
          }
          k = (uint64_t)v ^ 0x8000000000000000ULL;
        } break;
        case M_UNSIGNED:
          switch (a.width) {
            case 1: k = ((const uint8_t*)a.values)[r]; break;
            case 2: k = ((const uint16_t*)a.values)[r]; break;
            case 4: k = ((const uint32_t*)a.values)[r]; break;
            default: k = ((const uint64_t*)a.values)[r]; break;
          }
          break;
        case M_FLOAT64: {  // IEEE total order, as arrow's sort (total_cmp)
          const uint64_t b = ((const uint64_t*)a.values)[r];
          k = (b >> 63) ? ~b : b ^ 0x8000000000000000ULL;
        } break;
        case M_FLOAT32: {
          const uint32_t b = ((const uint32_t*)a.values)[r];
          k = (b >> 31) ? (uint32_t)~b : b ^ 0x80000000u;
        } break;
        case M_DEC_LO: k = ((const uint64_t*)a.values)[2 * r]; break;
        case M_DEC_HI: k = ((const uint64_t*)a.values)[2 * r + 1] ^ 0x8000000000000000ULL; break;
        case M_BOOL: k = bit_at((const uint8_t*)a.values, r) ? 1 : 0; break;
        case M_STR_LEN: k = (uint64_t)(a.offsets[r + 1] - a.offsets[r]); break;
        default: {  // M_STR_CHUNK: bytes [8 chunk, 8 chunk + 8) big-endian, zero padded
          const int32_t b = a.offsets[r], len = a.offsets[r + 1] - b;
          const uint8_t* p = (const uint8_t*)a.values + b;
          for (int j = 0; j < 8; j++) {
            const int32_t at = a.chunk * 8 + j;
            k = (k << 8) | (at < len ? p[at] : 0);
          }
        } break;
      }
      if (a.descending) k = ~k;
    }
    key[i] = k;
  }
}

__global__ void __launch_bounds__(kThreads) iota_u32_kernel(uint32_t* p, int64_t n) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) p[i] = (uint32_t)i;
}
__global__ void __launch_bounds__(kThreads) max_len_kernel(const int32_t* __restrict__ offs, int64_t n, int32_t* __restrict__ out) {
  int32_t m = 0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    m = max(m, offs[i + 1] - offs[i]);
  m = __reduce_max_sync(0xffffffffu, m);
  if ((threadIdx.x & 31) == 0 && m) atomicMax(out, m);
}

// ---- take: out[i] = in[perm[i]] for every column kind --------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads)
    take_fixed_kernel(const uint8_t* __restrict__ in, int width, const uint32_t* __restrict__ perm, int64_t n, uint8_t* __restrict__ out) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = perm[i];
    switch (width) {
      case 16: ((uint4*)out)[i] = ((const uint4*)in)[r]; break;
      case 8: ((uint64_t*)out)[i] = ((const uint64_t*)in)[r]; break;
      case 4: ((uint32_t*)out)[i] = ((const uint32_t*)in)[r]; break;
      case 2: ((uint16_t*)out)[i] = ((const uint16_t*)in)[r]; break;
      default: out[i] = in[r]; break;
    }
  }
}
__global__ void __launch_bounds__(kThreads)
    take_bits_kernel(const uint8_t* __restrict__ in, const uint32_t* __restrict__ perm, int64_t n, uint32_t* __restrict__ out) {
  const int64_t n_pad = (n + 31) & ~31LL;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_pad; i += (int64_t)gridDim.x * blockDim.x) {
    const uint32_t vb = __ballot_sync(0xffffffffu, i < n && bit_at(in, perm[i < n ? i : 0]));
    if ((threadIdx.x & 31) == 0) out[i >> 5] = vb;
  }
}
__global__ void __launch_bounds__(kThreads)
    take_lens_kernel(const int32_t* __restrict__ offs, const uint32_t* __restrict__ perm, int64_t n, int32_t* __restrict__ lens) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = perm[i];
    lens[i] = offs[r + 1] - offs[r];
  }
}
__global__ void __launch_bounds__(kThreads)
    take_str_kernel(const int32_t* __restrict__ in_offs, const uint8_t* __restrict__ in_data, const uint32_t* __restrict__ perm,
                    const int32_t* __restrict__ out_offs, uint8_t* __restrict__ out_data, int64_t n) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int32_t b = out_offs[i], len = out_offs[i + 1] - b;
    const uint8_t* src = in_data + in_offs[perm[i]];
    for (int32_t j = 0; j < len; j++) out_data[b + j] = src[j];
  }
}

int take_column(bgpu_ctx* ctx, const Column& c, const uint32_t* perm, int64_t n, Column* out) {
  cudaStream_t st = ctx->stream;
  *out = c;
  out->owned.clear();
  out->length = n;
  const int64_t n1 = std::max<int64_t>(n, 1);
  const size_t bitmap = (size_t)ceil_div64(n1, 32) * 4;
  if (c.validity) {
    DevAllocPtr v;
    BGPU_TRY(dev_alloc(ctx, bitmap, &v));
    if (n) take_bits_kernel<<<grid_for(ctx, n), kThreads, 0, st>>>(c.validity, perm, n, (uint32_t*)v->p);
    out->validity = (uint8_t*)v->p;
    out->owned.push_back(v);
  }
  if (c.kind == K_UTF8 || c.kind == K_BINARY) {
    DevAllocPtr lens, offs, data;
    BGPU_TRY(dev_alloc(ctx, (size_t)n1 * 4, &lens));
    BGPU_TRY(dev_alloc(ctx, (size_t)(n + 1) * 4, &offs));
    BGPU_CUDA(ctx, cudaMemsetAsync(offs->p, 0, 4, st));
    int32_t total = 0;
    if (n) {
      take_lens_kernel<<<grid_for(ctx, n), kThreads, 0, st>>>(c.offsets, perm, n, (int32_t*)lens->p);
      size_t tmp_bytes = 0;
      cub::DeviceScan::InclusiveSum(nullptr, tmp_bytes, (const int32_t*)lens->p, (int32_t*)offs->p + 1, (int)n, st);
      DevAllocPtr tmp;
      BGPU_TRY(dev_alloc(ctx, tmp_bytes, &tmp));
      BGPU_CUDA(ctx, cub::DeviceScan::InclusiveSum(tmp->p, tmp_bytes, (const int32_t*)lens->p, (int32_t*)offs->p + 1, (int)n, st));
      BGPU_CUDA(ctx, cudaMemcpyAsync(&total, (int32_t*)offs->p + n, 4, cudaMemcpyDeviceToHost, st));
      BGPU_CUDA(ctx, cudaStreamSynchronize(st));
    }
    BGPU_TRY(dev_alloc(ctx, (size_t)total + 16, &data));
    if (total)
      take_str_kernel<<<grid_for(ctx, n), kThreads, 0, st>>>(c.offsets, (const uint8_t*)c.values, perm, (const int32_t*)offs->p,
                                                            (uint8_t*)data->p, n);
    out->offsets = (int32_t*)offs->p; out->values = data->p; out->data_bytes = total; out->data_first = 0;
    out->owned.push_back(offs); out->owned.push_back(data);
  } else if (c.kind == K_BOOL) {
    DevAllocPtr v;
    BGPU_TRY(dev_alloc(ctx, bitmap, &v));
    if (n) take_bits_kernel<<<grid_for(ctx, n), kThreads, 0, st>>>((const uint8_t*)c.values, perm, n, (uint32_t*)v->p);
    out->values = v->p;
    out->owned.push_back(v);
  } else {  // fixed width, dictionary indices (the dictionary is shared)
    DevAllocPtr v;
    BGPU_TRY(dev_alloc(ctx, (size_t)n1 * c.width, &v));
    if (n) take_fixed_kernel<<<grid_for(ctx, n), kThreads, 0, st>>>((const uint8_t*)c.values, c.width, perm, n, (uint8_t*)v->p);
    out->values = v->p;
    out->owned.push_back(v);
    if (c.kind == K_DICT32) out->owned.insert(out->owned.end(), c.owned.begin(), c.owned.end());
  }
  ctx->launches += 2;
  BGPU_CUDA(ctx, cudaGetLastError());
  return BGPU_OK;
}
}  // namespace
}  // namespace bgpu

using namespace bgpu;

extern "C" int bgpu_batch_sort(bgpu_ctx* ctx, const bgpu_batch* in, const int32_t* sort_cols, const uint8_t* descending,
                               const uint8_t* nulls_first, int32_t n_keys, bgpu_batch** out) try {
  if (!ctx || !in || !out || !sort_cols || n_keys < 1) return BGPU_EINVAL;
  *out = nullptr;
  if (!ctx->has_device) return ctx->fail(BGPU_ENODEVICE, "context has no device (libbgpu has no CPU fallback)");
  const int64_t n = in->n_rows;
  if (n > 0xfffffff0LL) return ctx->fail(BGPU_EUNSUPPORTED, "sort: more than 2^32 rows");
  BGPU_CUDA(ctx, cudaSetDevice(ctx->device));
  GpuTimeScope timed(ctx, T_KERNEL);
  cudaStream_t st = ctx->stream;
  const int64_t n1 = std::max<int64_t>(n, 1);
  DevAllocPtr perm[2], key[2], tmp;
  for (int i = 0; i < 2; i++) {
    BGPU_TRY(dev_alloc(ctx, (size_t)n1 * 4, &perm[i]));
    BGPU_TRY(dev_alloc(ctx, (size_t)n1 * 8, &key[i]));
  }
  size_t tmp_bytes = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, (const uint64_t*)key[0]->p, (uint64_t*)key[1]->p, (const uint32_t*)perm[0]->p,
                                  (uint32_t*)perm[1]->p, (int)n1, 0, 64, st);
  BGPU_TRY(dev_alloc(ctx, tmp_bytes, &tmp));
  int cur = 0;
  if (n) {
    iota_u32_kernel<<<grid_for(ctx, n), kThreads, 0, st>>>((uint32_t*)perm[0]->p, n);
    ctx->launches++;
  }
  auto pass = [&](const KeyArgs& ka, int bits) -> int {
    if (!n) return BGPU_OK;
    sort_key_kernel<<<grid_for(ctx, n), kThreads, 0, st>>>(ka, (const uint32_t*)perm[cur]->p, n, (uint64_t*)key[0]->p);
    size_t tb = tmp_bytes;
    BGPU_CUDA(ctx, cub::DeviceRadixSort::SortPairs(tmp->p, tb, (const uint64_t*)key[0]->p, (uint64_t*)key[1]->p,
                                                   (const uint32_t*)perm[cur]->p, (uint32_t*)perm[cur ^ 1]->p, (int)n, 0, bits, st));
    cur ^= 1;
    ctx->launches += 2;
    return BGPU_OK;
  };
  for (int32_t k = n_keys - 1; k >= 0; k--) {  // least significant key first: every pass is stable
    const int32_t ci = sort_cols[k];
    if (ci < 0 || ci >= (int32_t)in->cols.size()) return ctx->fail(BGPU_EINVAL, "sort column out of range");
    const Column& c = in->cols[(size_t)ci];
    KeyArgs ka{c.values, c.offsets, c.validity, c.width, 0, 0, descending && descending[k] ? 1 : 0, nulls_first && nulls_first[k] ? 1 : 0};
    const char f0 = c.format.empty() ? '?' : c.format[0];
    if (c.kind == K_UTF8 || c.kind == K_BINARY) {
      DevAllocPtr d_max;
      BGPU_TRY(dev_alloc(ctx, 4, &d_max));
      BGPU_CUDA(ctx, cudaMemsetAsync(d_max->p, 0, 4, st));
      int32_t max_len = 0;
      if (n) {
        max_len_kernel<<<grid_for(ctx, n), kThreads, 0, st>>>(c.offsets, n, (int32_t*)d_max->p);
        ctx->launches++;
        BGPU_CUDA(ctx, cudaMemcpyAsync(&max_len, d_max->p, 4, cudaMemcpyDeviceToHost, st));
        BGPU_CUDA(ctx, cudaStreamSynchronize(st));
      }
      if (max_len > 4096) return ctx->fail(BGPU_EUNSUPPORTED, "sort: string keys longer than 4096 bytes");
      ka.mode = M_STR_LEN;  // a prefix sorts before its extensions (also when the rest is NUL bytes)
      BGPU_TRY(pass(ka, 32));
      for (int32_t chunk = (max_len + 7) / 8 - 1; chunk >= 0; chunk--) {
        ka.mode = M_STR_CHUNK; ka.chunk = chunk;
        BGPU_TRY(pass(ka, 64));
      }
    } else if (c.kind == K_BOOL) {
      ka.mode = M_BOOL;
      BGPU_TRY(pass(ka, 1));
    } else if (c.kind == K_FIXED && c.width == 16) {
      ka.mode = M_DEC_LO;
      BGPU_TRY(pass(ka, 64));
      ka.mode = M_DEC_HI;
      BGPU_TRY(pass(ka, 64));
    } else if (c.kind == K_FIXED) {
      ka.mode = f0 == 'g' ? M_FLOAT64 : f0 == 'f' ? M_FLOAT32 : (f0 == 'C' || f0 == 'S' || f0 == 'I' || f0 == 'L') ? M_UNSIGNED : M_SIGNED;
      if (f0 == 'e') return ctx->fail(BGPU_EUNSUPPORTED, "sort: Float16 key");
      BGPU_TRY(pass(ka, ka.mode == M_FLOAT32 ? 32 : 64));
    } else {
      return ctx->fail(BGPU_EUNSUPPORTED, "sort: dictionary-encoded key column " + c.name);
    }
    if (c.validity) {
      ka.mode = M_NULLFLAG;
      BGPU_TRY(pass(ka, 1));
    }
  }
  std::unique_ptr<bgpu_batch> res(new bgpu_batch(ctx));
  res->n_rows = n;
  res->cols.resize(in->cols.size());
  for (size_t i = 0; i < in->cols.size(); i++)
    BGPU_TRY(take_column(ctx, in->cols[i], (const uint32_t*)perm[cur]->p, n, &res->cols[i]));
  BGPU_CUDA(ctx, cudaGetLastError());
  *out = res.release();
  return BGPU_OK;
} catch (const std::exception& e) {
  return bgpu::abi_guard(ctx, e);  // nothing is thrown across the C ABI
}
