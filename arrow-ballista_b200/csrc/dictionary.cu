// Dictionary<Int32, Utf8 | Binary | fixed> columns on the move.
//
// The reference's db-benchmark keys arrive dictionary-encoded (/root/reference/benchmarks/db-benchmark/groupby-datafusion.py:56)
// and a Partial aggregate's state keeps them so; its shuffle writer / reader carry them as Arrow IPC dictionary batches
// (/root/reference/ballista/core/src/async_reader/mod.rs:168-258). A dictionary is only meaningful next to its batch, so
// whenever rows of SEVERAL batches meet - concat, the exchange between ranks, the per-partition messages of a shuffle
// file - this library moves the VALUES and re-encodes on arrival:
//   dict_decode_batch    every dictionary column -> its values, row by row (a `take`; null index or null value = null row)
//   dict_identity_encode the moved values become the dictionary, the indices are 0 .. n-1 (valid Arrow: dictionaries may
//                        hold duplicates; the aggregate compares dictionary keys by VALUE, aggregate_run.cu)
// Results are the same rows with the same logical values; the schema (Dictionary<Int32, T>) is unchanged.
#include <cub/device/device_scan.cuh>

#include "internal.h"

namespace bgpu {
namespace {
constexpr int kThreads = 256;
int grid_for(bgpu_ctx* ctx, int64_t n) {
  return (int)std::max<int64_t>(1, std::min<int64_t>(ceil_div64(n, kThreads), (int64_t)ctx->sm_count * 8));
}
__device__ __forceinline__ bool bit_at(const uint8_t* bits, int64_t i) { return (bits[i >> 3] >> (i & 7)) & 1; }

// row validity of the decoded column (index valid and value valid) as a bitmap, and utf8 lengths
__global__ void __launch_bounds__(kThreads)
    dict_lens_kernel(const int32_t* __restrict__ idx, const uint8_t* __restrict__ idx_valid, const int32_t* __restrict__ d_offs,
                     const uint8_t* __restrict__ d_valid, int64_t n, int32_t* __restrict__ lens, uint32_t* __restrict__ out_valid) {
  const int64_t n_pad = (n + 31) & ~31LL;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_pad; i += (int64_t)gridDim.x * blockDim.x) {
    bool valid = false;
    if (i < n) {
      valid = !idx_valid || bit_at(idx_valid, i);
      int32_t k = 0;
      if (valid) {
        k = idx[i];
        if (d_valid && !bit_at(d_valid, k)) valid = false;
      }
      if (lens) lens[i] = valid ? d_offs[k + 1] - d_offs[k] : 0;
    }
    if (out_valid) {
      const uint32_t vb = __ballot_sync(0xffffffffu, valid);
      if ((threadIdx.x & 31) == 0) out_valid[i >> 5] = vb;
    }
  }
}
__global__ void __launch_bounds__(kThreads)
    dict_bytes_kernel(const int32_t* __restrict__ idx, const int32_t* __restrict__ d_offs, const uint8_t* __restrict__ d_data,
                      const int32_t* __restrict__ out_offs, uint8_t* __restrict__ out_data, int64_t n) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int32_t b = out_offs[i], len = out_offs[i + 1] - b;
    if (len <= 0) continue;
    const uint8_t* src = d_data + d_offs[idx[i]];
    for (int32_t j = 0; j < len; j++) out_data[b + j] = src[j];
  }
}
__global__ void __launch_bounds__(kThreads)
    dict_fixed_kernel(const int32_t* __restrict__ idx, const uint8_t* __restrict__ valid_bits, const uint8_t* __restrict__ d_values,
                      int width, uint8_t* __restrict__ out, int64_t n) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const bool valid = !valid_bits || bit_at(valid_bits, i);
    const uint8_t* src = d_values + (size_t)(valid ? idx[i] : 0) * width;
    for (int j = 0; j < width; j++) out[(size_t)i * width + j] = valid ? src[j] : 0;
  }
}
__global__ void __launch_bounds__(kThreads) iota_i32_kernel(int32_t* __restrict__ out, int64_t n) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) out[i] = (int32_t)i;
}
// index of row r inside its partition: r - offsets[p(r)] (offsets ascending, n + 1 entries)
__global__ void __launch_bounds__(kThreads)
    iota_in_partition_kernel(const uint64_t* __restrict__ offsets, uint32_t n_parts, int32_t* __restrict__ out, int64_t n) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    uint32_t lo = 0, hi = n_parts;  // the last p with offsets[p] <= i
    while (hi - lo > 1) {
      const uint32_t mid = (lo + hi) >> 1;
      if (offsets[mid] <= (uint64_t)i) lo = mid;
      else hi = mid;
    }
    out[i] = (int32_t)((uint64_t)i - offsets[lo]);
  }
}
}  // namespace

bool batch_has_dictionary(const bgpu_batch* b) {
  for (const Column& c : b->cols)
    if (c.kind == K_DICT32) return true;
  return false;
}

// The values of dictionary column `c` (n rows), row by row.
int dict_decode_column(bgpu_ctx* ctx, const Column& c, int64_t n, Column* out) {
  const Column& d = *c.dict;
  if (d.kind == K_DICT32 || d.kind == K_BOOL) return ctx->fail(BGPU_EUNSUPPORTED, "dictionary of " + d.format + " values (column " + c.name + ")");
  cudaStream_t st = ctx->stream;
  *out = Column();
  out->name = c.name; out->format = d.format; out->flags = c.flags | ARROW_FLAG_NULLABLE; out->kind = d.kind; out->width = d.width;
  out->length = n;
  const int64_t n1 = std::max<int64_t>(n, 1);
  const bool nullable = c.validity || d.validity;
  DevAllocPtr valid;
  if (nullable) {
    BGPU_TRY(dev_alloc(ctx, (size_t)ceil_div64(n1, 32) * 4, &valid));
    out->validity = (uint8_t*)valid->p;
    out->null_count = -1;
    out->owned.push_back(valid);
  }
  const int32_t* idx = (const int32_t*)c.values;
  if (d.kind == K_UTF8 || d.kind == K_BINARY) {
    DevAllocPtr lens, offs, data;
    BGPU_TRY(dev_alloc(ctx, (size_t)n1 * 4, &lens));
    BGPU_TRY(dev_alloc(ctx, (size_t)(n + 1) * 4, &offs));
    BGPU_CUDA(ctx, cudaMemsetAsync(offs->p, 0, 4, st));
    int32_t total = 0;
    if (n) {
      dict_lens_kernel<<<grid_for(ctx, n), kThreads, 0, st>>>(idx, c.validity, d.offsets, d.validity, n, (int32_t*)lens->p,
                                                            nullable ? (uint32_t*)valid->p : nullptr);
      size_t tmp_bytes = 0;
      cub::DeviceScan::InclusiveSum(nullptr, tmp_bytes, (const int32_t*)lens->p, (int32_t*)offs->p + 1, (int)n, st);
      DevAllocPtr tmp;
      BGPU_TRY(dev_alloc(ctx, tmp_bytes, &tmp));
      BGPU_CUDA(ctx, cub::DeviceScan::InclusiveSum(tmp->p, tmp_bytes, (const int32_t*)lens->p, (int32_t*)offs->p + 1, (int)n, st));
      BGPU_CUDA(ctx, cudaMemcpyAsync(&total, (int32_t*)offs->p + n, 4, cudaMemcpyDeviceToHost, st));
      BGPU_CUDA(ctx, cudaStreamSynchronize(st));
      ctx->launches += 2;
    }
    BGPU_TRY(dev_alloc(ctx, (size_t)total + 16, &data));
    if (total) {
      dict_bytes_kernel<<<grid_for(ctx, n), kThreads, 0, st>>>(idx, d.offsets, (const uint8_t*)d.values, (const int32_t*)offs->p,
                                                             (uint8_t*)data->p, n);
      ctx->launches++;
    }
    out->offsets = (int32_t*)offs->p; out->values = data->p; out->data_bytes = total;
    out->owned.push_back(offs); out->owned.push_back(data);
  } else {
    DevAllocPtr vals;
    BGPU_TRY(dev_alloc(ctx, (size_t)n1 * d.width, &vals));
    if (n) {
      if (nullable) {
        dict_lens_kernel<<<grid_for(ctx, n), kThreads, 0, st>>>(idx, c.validity, nullptr, d.validity, n, nullptr, (uint32_t*)valid->p);
        ctx->launches++;
      }
      dict_fixed_kernel<<<grid_for(ctx, n), kThreads, 0, st>>>(idx, nullable ? (const uint8_t*)valid->p : nullptr,
                                                             (const uint8_t*)d.values, d.width, (uint8_t*)vals->p, n);
      ctx->launches++;
    }
    out->values = vals->p;
    out->owned.push_back(vals);
  }
  BGPU_CUDA(ctx, cudaGetLastError());
  return BGPU_OK;
}

// A batch like `in` with every dictionary column decoded (other columns shared); dict_cols = their indices.
int dict_decode_batch(bgpu_ctx* ctx, const bgpu_batch* in, std::unique_ptr<bgpu_batch>* out, std::vector<size_t>* dict_cols) {
  std::unique_ptr<bgpu_batch> b(new bgpu_batch(ctx));
  b->n_rows = in->n_rows;
  b->cols = in->cols;
  for (size_t i = 0; i < in->cols.size(); i++) {
    if (in->cols[i].kind != K_DICT32) continue;
    BGPU_TRY(dict_decode_column(ctx, in->cols[i], in->n_rows, &b->cols[i]));
    dict_cols->push_back(i);
  }
  *out = std::move(b);
  return BGPU_OK;
}

// Column `i` of `b` (values) becomes Dictionary<Int32, values> again: the column is the dictionary, the indices are
// 0 .. n-1 (or, with `offsets`, the row's position inside its partition - the per-partition messages of a shuffle file
// carry one dictionary per partition). `proto` supplies the dictionary column's name / format / flags.
int dict_identity_encode(bgpu_ctx* ctx, bgpu_batch* b, size_t i, const Column& proto, const uint64_t* d_offsets, uint32_t n_parts) {
  Column vals = b->cols[i];
  const int64_t n = b->n_rows;
  DevAllocPtr idx;
  BGPU_TRY(dev_alloc(ctx, (size_t)std::max<int64_t>(n, 1) * 4, &idx));
  if (n) {
    if (d_offsets) iota_in_partition_kernel<<<grid_for(ctx, n), kThreads, 0, ctx->stream>>>(d_offsets, n_parts, (int32_t*)idx->p, n);
    else iota_i32_kernel<<<grid_for(ctx, n), kThreads, 0, ctx->stream>>>((int32_t*)idx->p, n);
    ctx->launches++;
    BGPU_CUDA(ctx, cudaGetLastError());
  }
  Column c;
  c.name = proto.name; c.format = proto.format; c.flags = proto.flags; c.kind = K_DICT32; c.width = 4;
  c.length = n; c.values = idx->p;
  c.validity = vals.validity; c.null_count = vals.null_count;
  c.owned = vals.owned;  // the validity bitmap lives with the values' buffers
  c.owned.push_back(idx);
  vals.name = proto.dict ? proto.dict->name : "";
  vals.length = n;
  c.dict = std::make_shared<Column>(std::move(vals));
  b->cols[i] = std::move(c);
  return BGPU_OK;
}
}  // namespace bgpu
