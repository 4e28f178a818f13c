// K6 for small states: the exchange of an aggregate state of a few rows (TPC-H Q1: four) as ONE fixed-size ncclAllGather.
//
// bgpu_hash_partition + bgpu_all_to_all move such a batch with a dozen launches, a header collective and a grouped
// ncclSend/ncclRecv per (buffer, peer, partition) - hundreds of microseconds that grow with the number of ranks, for a
// payload of a few hundred bytes. Here every rank packs its batch into a fixed 32 KB slot (one kernel), ONE ncclAllGather
// hands every rank every slot, and one kernel keeps the rows of the partitions this rank owns (create_hashes with the
// reference's seeds, hash % n_partitions, owner = partition % world) and rebuilds Arrow columns; one host round trip
// returns the row count. The rows a rank ends up with are exactly those bgpu_all_to_all(bgpu_hash_partition(batch))
// delivers, ordered by (source rank, source row) instead of (partition, source rank) - FinalPartitioned does not care.
//
// Replaces, like exchange.cu, the shuffle file write + Flight do_get + reader fan-in of
// /root/reference/ballista/core/src/execution_plans/shuffle_writer.rs:351-385,
// /root/reference/ballista/executor/src/flight_service.rs:124-166,
// /root/reference/ballista/core/src/execution_plans/shuffle_reader.rs:539-624 inside one NVSwitch node.
#include <string.h>

#include <algorithm>

#include "hash.cuh"
#include "internal.h"
#include "nccl_dl.h"

namespace bgpu {
constexpr int kSmallCols = 32;
constexpr uint32_t kSmallRows = 256;        // rows per rank the slot is meant for
constexpr uint32_t kSmallSlot = 32 << 10;   // bytes per rank
constexpr int kSmallWorld = 16;
constexpr int kSmallThreads = 1024;

struct SmallHeader {
  uint32_t n_rows, flag /* 1: does not fit, take the general path */, valid_mask, pad;
  uint32_t str_bytes[kSmallCols];
  uint32_t reserved[28];
};
static_assert(sizeof(SmallHeader) == 256, "slot header is 256 bytes");

struct SmallCol {
  const void* values;
  const int32_t* offsets;
  const uint8_t* validity;
  int32_t kind, width;
};
struct SmallPackArgs {
  SmallCol c[kSmallCols];
  int32_t n_cols;
  uint32_t n_rows, force_flag;
  uint8_t* slot;
};

struct SmallLayout {
  uint32_t valid[kSmallCols], values[kSmallCols], data[kSmallCols];  // byte offsets inside the slot
  uint32_t total;
};
__host__ __device__ inline uint32_t small_align(uint32_t x) { return (x + 15u) & ~15u; }
__host__ __device__ inline void small_layout(const SmallHeader& h, const int32_t* kind, const int32_t* width, int n_cols,
                                             SmallLayout* L) {
  uint32_t off = sizeof(SmallHeader);
  const uint32_t n = h.n_rows;
  for (int c = 0; c < n_cols; c++) {
    L->valid[c] = L->values[c] = L->data[c] = 0;
    if ((h.valid_mask >> c) & 1) { L->valid[c] = off; off += small_align((n + 7) / 8); }
    L->values[c] = off;
    if (kind[c] == K_UTF8 || kind[c] == K_BINARY) {
      off += small_align((n + 1) * 4);
      L->data[c] = off;
      off += small_align(h.str_bytes[c]);
    } else if (kind[c] == K_BOOL) {
      off += small_align((n + 7) / 8);
    } else {
      off += small_align(n * (uint32_t)width[c]);
    }
  }
  L->total = off;
}

// This rank's batch -> its slot (one CTA).
__global__ void __launch_bounds__(256) small_pack_kernel(const __grid_constant__ SmallPackArgs a) {
  __shared__ SmallHeader h;
  __shared__ SmallLayout L;
  __shared__ int32_t kind[kSmallCols], width[kSmallCols];
  const uint32_t n = a.n_rows;
  if (threadIdx.x == 0) {
    memset(&h, 0, sizeof h);
    h.n_rows = n;
    h.flag = a.force_flag;
    for (int c = 0; c < a.n_cols; c++) {
      kind[c] = a.c[c].kind;
      width[c] = a.c[c].width;
      if (a.c[c].validity) h.valid_mask |= 1u << c;
      if (!h.flag && (kind[c] == K_UTF8 || kind[c] == K_BINARY) && n) {
        const int64_t b = (int64_t)a.c[c].offsets[n] - a.c[c].offsets[0];
        if (b < 0 || b > kSmallSlot) h.flag = 1;
        else h.str_bytes[c] = (uint32_t)b;
      }
    }
    small_layout(h, kind, width, a.n_cols, &L);
    if (L.total > kSmallSlot) h.flag = 1;
    if (h.flag) { h.n_rows = 0; h.valid_mask = 0; }
    *(SmallHeader*)a.slot = h;
  }
  __syncthreads();
  if (h.flag) return;
  for (int c = 0; c < a.n_cols; c++) {
    const SmallCol& col = a.c[c];
    if (col.validity)
      for (uint32_t i = threadIdx.x; i < (n + 7) / 8; i += 256) a.slot[L.valid[c] + i] = col.validity[i];
    uint8_t* out = a.slot + L.values[c];
    if (kind[c] == K_UTF8 || kind[c] == K_BINARY) {
      const int32_t first = n ? col.offsets[0] : 0;
      for (uint32_t i = threadIdx.x; i <= n; i += 256) ((int32_t*)out)[i] = n ? col.offsets[i] - first : 0;
      const uint8_t* src = (const uint8_t*)col.values + first;
      for (uint32_t i = threadIdx.x; i < h.str_bytes[c]; i += 256) a.slot[L.data[c] + i] = src[i];
    } else {
      const uint32_t bytes = kind[c] == K_BOOL ? (n + 7) / 8 : n * (uint32_t)width[c];
      for (uint32_t i = threadIdx.x; i < bytes; i += 256) out[i] = ((const uint8_t*)col.values)[i];
    }
  }
}

struct SmallOutCol {
  void* values;
  int32_t* offsets;
  uint8_t* data;
  uint32_t* valid;  // zero-initialised
  int32_t kind, width;
};
struct SmallUnpackArgs {
  SmallOutCol c[kSmallCols];
  int32_t key_cols[kMaxKeys];
  int32_t n_cols, n_keys, rank, world;
  Modulo mod;
  const uint8_t* slots;  // world x kSmallSlot
  uint32_t* status;      // {rows kept, flag, OR of the sources' valid masks, 0}
};

__device__ __forceinline__ uint32_t block_exclusive_scan(uint32_t v, uint32_t* s_warp, uint32_t* total) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint32_t x = v;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const uint32_t y = __shfl_up_sync(0xffffffffu, x, d);
    if (lane >= d) x += y;
  }
  if (lane == 31) s_warp[warp] = x;
  __syncthreads();
  uint32_t pre = 0, all = 0;
  for (int w = 0; w < kSmallThreads / 32; w++) {
    if (w < warp) pre += s_warp[w];
    all += s_warp[w];
  }
  __syncthreads();
  *total = all;
  return pre + x - v;
}

// Every rank's slot -> the rows of the partitions this rank owns, as Arrow columns (one CTA).
__global__ void __launch_bounds__(kSmallThreads) small_unpack_kernel(const __grid_constant__ SmallUnpackArgs a) {
  __shared__ SmallLayout L[kSmallWorld];
  __shared__ uint32_t s_rows[kSmallWorld + 1], s_warp[kSmallThreads / 32];
  __shared__ int32_t kind[kSmallCols], width[kSmallCols];
  __shared__ uint32_t s_flag, s_valid_or;
  if (threadIdx.x == 0) {
    s_flag = 0;
    s_valid_or = 0;
    for (int c = 0; c < a.n_cols; c++) { kind[c] = a.c[c].kind; width[c] = a.c[c].width; }
    s_rows[0] = 0;
    for (int s = 0; s < a.world; s++) {
      const SmallHeader& h = *(const SmallHeader*)(a.slots + (size_t)s * kSmallSlot);
      if (h.flag) s_flag = 1;
      s_valid_or |= h.valid_mask;
      small_layout(h, kind, width, a.n_cols, &L[s]);
      s_rows[s + 1] = s_rows[s] + h.n_rows;
    }
  }
  __syncthreads();
  if (s_flag) {
    if (threadIdx.x == 0) { a.status[0] = 0; a.status[1] = 1; a.status[2] = 0; a.status[3] = 0; }
    return;
  }
  const uint32_t total = s_rows[a.world];
  uint32_t kept_before = 0;                       // rows kept in earlier tiles
  __shared__ uint32_t s_str_cursor[kSmallCols];
  if (threadIdx.x < kSmallCols) s_str_cursor[threadIdx.x] = 0;
  __syncthreads();
  for (uint32_t g0 = 0; g0 < total; g0 += kSmallThreads) {
    const uint32_t g = g0 + threadIdx.x;
    int s = 0;
    uint32_t r = 0;
    bool keep = false;
    const uint8_t* slot = nullptr;
    if (g < total) {
      while (g >= s_rows[s + 1]) s++;
      r = g - s_rows[s];
      slot = a.slots + (size_t)s * kSmallSlot;
      const SmallHeader& h = *(const SmallHeader*)slot;
      // create_hashes over the key columns: column 0 sets, later columns combine, null rows keep the running hash
      uint64_t hash = 0;
      for (int k = 0; k < a.n_keys; k++) {
        const int c = a.key_cols[k];
        const bool has_valid = (h.valid_mask >> c) & 1;
        if (has_valid && !bit_is_set(slot + L[s].valid[c], r)) continue;
        KCol kc{slot + L[s].values[c], nullptr, nullptr, nullptr, kind[c], width[c]};
        if (kind[c] == K_UTF8 || kind[c] == K_BINARY) {
          kc.offsets = (const int32_t*)(slot + L[s].values[c]);
          kc.values = slot + L[s].data[c];
        }
        const uint64_t vh = hash_value(kc, r);
        hash = k == 0 ? vh : combine_hashes(vh, hash);
      }
      keep = (partition_of(hash, a.mod) % (uint32_t)a.world) == (uint32_t)a.rank;
    }
    uint32_t tile_kept = 0;
    const uint32_t d = kept_before + block_exclusive_scan(keep ? 1u : 0u, s_warp, &tile_kept);
    for (int c = 0; c < a.n_cols; c++) {
      const SmallOutCol& oc = a.c[c];
      const SmallHeader* h = slot ? (const SmallHeader*)slot : nullptr;
      const bool has_valid = keep && ((h->valid_mask >> c) & 1);
      const bool valid = keep && (!has_valid || bit_is_set(slot + L[s].valid[c], r));
      if (valid) atomicOr(&oc.valid[d >> 5], 1u << (d & 31));
      if (kind[c] == K_UTF8 || kind[c] == K_BINARY) {
        const int32_t* offs = keep ? (const int32_t*)(slot + L[s].values[c]) : nullptr;
        const uint32_t len = keep ? (uint32_t)(offs[r + 1] - offs[r]) : 0;
        uint32_t tile_bytes = 0;
        const uint32_t start = s_str_cursor[c] + block_exclusive_scan(len, s_warp, &tile_bytes);
        if (keep) {
          oc.offsets[d] = (int32_t)start;
          const uint8_t* src = slot + L[s].data[c] + offs[r];
          for (uint32_t i = 0; i < len; i++) oc.data[start + i] = src[i];
        }
        __syncthreads();
        if (threadIdx.x == 0) s_str_cursor[c] += tile_bytes;
        __syncthreads();
      } else if (keep) {
        const uint8_t* src = slot + L[s].values[c];
        if (kind[c] == K_BOOL) {
          if (bit_is_set(src, r)) atomicOr(&((uint32_t*)oc.values)[d >> 5], 1u << (d & 31));
        } else if (width[c] == 16) ((uint4*)oc.values)[d] = ((const uint4*)src)[r];
        else if (width[c] == 8) ((uint64_t*)oc.values)[d] = ((const uint64_t*)src)[r];
        else if (width[c] == 4) ((uint32_t*)oc.values)[d] = ((const uint32_t*)src)[r];
        else if (width[c] == 2) ((uint16_t*)oc.values)[d] = ((const uint16_t*)src)[r];
        else ((uint8_t*)oc.values)[d] = src[r];
      }
    }
    kept_before += tile_kept;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int c = 0; c < a.n_cols; c++)
      if (kind[c] == K_UTF8 || kind[c] == K_BINARY) a.c[c].offsets[kept_before] = (int32_t)s_str_cursor[c];
    a.status[0] = kept_before;
    a.status[1] = 0;
    a.status[2] = s_valid_or;
    a.status[3] = 0;
  }
}

static Modulo small_modulo(uint32_t n) {
  Modulo m{};
  m.n = n;
  m.pow2 = (n & (n - 1)) == 0;
  m.mask = n - 1;
  m.small = n < 65536;
  m.magic = n > 1 ? (UINT64_MAX / n + 1) : 0;
  m.two32_mod = (uint32_t)((1ULL << 32) % n);
  return m;
}
}  // namespace bgpu

using namespace bgpu;

extern "C" int bgpu_small_exchange(bgpu_ctx* ctx, void* comm, int32_t rank, int32_t world, const bgpu_batch* in,
                                   const int32_t* key_cols, int32_t n_keys, uint32_t n_partitions, bgpu_batch** out) try {
  if (!ctx || !out) return BGPU_EINVAL;
  *out = nullptr;
  if (!in || !key_cols || world < 1 || rank < 0 || rank >= world || n_partitions == 0 || (world > 1 && !comm))
    return ctx->fail(BGPU_EINVAL, "bgpu_small_exchange: bad arguments");
  if (!ctx->has_device) return ctx->fail(BGPU_ENODEVICE, "context has no device");
  // schema-level limits: the same answer on every rank, before any collective
  const size_t n_cols = in->cols.size();
  if (n_cols > (size_t)kSmallCols || world > kSmallWorld || n_keys < 1 || n_keys > kMaxKeys)
    return ctx->fail(BGPU_EUNSUPPORTED, "bgpu_small_exchange: schema or world beyond the small path");
  cudaSetDevice(ctx->device);
  GpuTimeScope timed(ctx, T_KERNEL);
  if (batch_has_dictionary(in)) {  // schema-level, the same on every rank: values travel, re-encoded on arrival (dictionary.cu)
    BGPU_CUDA(ctx, cudaSetDevice(ctx->device));
    std::unique_ptr<bgpu_batch> decoded;
    std::vector<size_t> dict_cols;
    BGPU_TRY(dict_decode_batch(ctx, in, &decoded, &dict_cols));
    bgpu_batch* got = nullptr;
    BGPU_TRY(bgpu_small_exchange(ctx, comm, rank, world, decoded.get(), key_cols, n_keys, n_partitions, &got));
    std::unique_ptr<bgpu_batch> res(got);
    for (size_t i : dict_cols) BGPU_TRY(dict_identity_encode(ctx, res.get(), i, in->cols[i], nullptr, 0));
    *out = res.release();
    return BGPU_OK;
  }
  for (int32_t k = 0; k < n_keys; k++)
    if (key_cols[k] < 0 || key_cols[k] >= (int32_t)n_cols) return ctx->fail(BGPU_EINVAL, "key column index out of range");
  BGPU_CUDA(ctx, cudaSetDevice(ctx->device));
  cudaStream_t st = ctx->stream;
  DevAllocPtr send, recv, status;
  BGPU_TRY(dev_alloc(ctx, kSmallSlot, &send));
  BGPU_TRY(dev_alloc(ctx, (size_t)kSmallSlot * world, &recv));
  BGPU_TRY(dev_alloc(ctx, 16, &status));
  SmallPackArgs pa{};
  pa.n_cols = (int32_t)n_cols;
  pa.n_rows = (uint32_t)std::min<int64_t>(in->n_rows, kSmallRows + 1);
  pa.force_flag = in->n_rows > (int64_t)kSmallRows ? 1u : 0u;  // too many rows here: every rank will see the flag
  pa.slot = (uint8_t*)send->p;
  for (size_t c = 0; c < n_cols; c++) {
    const Column& col = in->cols[c];
    pa.c[c] = SmallCol{col.values, col.offsets, col.validity, (int32_t)col.kind, col.width};
  }
  small_pack_kernel<<<1, 256, 0, st>>>(pa);
  ctx->launches++;
  BGPU_CUDA(ctx, cudaGetLastError());
  if (world > 1) {
    const Nccl& nc = nccl();
    if (!nc.ok) return ctx->fail(BGPU_ENCCL, "NCCL unavailable: " + nc.why);
    BGPU_NCCL(ctx, nc.AllGather(send->p, recv->p, kSmallSlot, kNcclUint8, comm, st));
  } else {
    BGPU_CUDA(ctx, cudaMemcpyAsync(recv->p, send->p, kSmallSlot, cudaMemcpyDeviceToDevice, st));
  }
  // output columns sized for the most rows / bytes the slots can hold
  const uint32_t max_rows = kSmallRows * (uint32_t)world;
  const size_t bitmap = ((size_t)max_rows + 31) / 32 * 4;
  std::unique_ptr<bgpu_batch> ob(new bgpu_batch(ctx));
  ob->cols.resize(n_cols);
  SmallUnpackArgs ua{};
  ua.n_cols = (int32_t)n_cols; ua.n_keys = n_keys; ua.rank = rank; ua.world = world;
  for (int32_t k = 0; k < n_keys; k++) ua.key_cols[k] = key_cols[k];
  ua.mod = small_modulo(n_partitions);
  ua.slots = (const uint8_t*)recv->p;
  ua.status = (uint32_t*)status->p;
  for (size_t c = 0; c < n_cols; c++) {
    const Column& ic = in->cols[c];
    Column& oc = ob->cols[c];
    oc.name = ic.name; oc.format = ic.format; oc.flags = ic.flags; oc.kind = ic.kind; oc.width = ic.width;
    SmallOutCol& uc = ua.c[c];
    uc.kind = ic.kind; uc.width = ic.width;
    DevAllocPtr valid, vals;
    BGPU_TRY(dev_alloc(ctx, bitmap, &valid));
    BGPU_CUDA(ctx, cudaMemsetAsync(valid->p, 0, bitmap, st));
    uc.valid = (uint32_t*)valid->p;
    oc.validity = (uint8_t*)valid->p;
    oc.owned.push_back(valid);
    if (ic.kind == K_UTF8 || ic.kind == K_BINARY) {
      DevAllocPtr data;
      const size_t cap = (size_t)kSmallSlot * world;
      BGPU_TRY(dev_alloc(ctx, ((size_t)max_rows + 1) * 4, &vals));
      BGPU_TRY(dev_alloc(ctx, cap + 16, &data));
      BGPU_CUDA(ctx, cudaMemsetAsync(vals->p, 0, 4, st));
      uc.offsets = (int32_t*)vals->p; uc.data = (uint8_t*)data->p;
      oc.offsets = (int32_t*)vals->p; oc.values = data->p; oc.data_bytes = (int64_t)cap;
      oc.owned.push_back(vals); oc.owned.push_back(data);
    } else {
      const size_t bytes = ic.kind == K_BOOL ? bitmap : (size_t)max_rows * ic.width;
      BGPU_TRY(dev_alloc(ctx, bytes, &vals));
      if (ic.kind == K_BOOL) BGPU_CUDA(ctx, cudaMemsetAsync(vals->p, 0, bytes, st));
      uc.values = vals->p;
      oc.values = vals->p;
      oc.owned.push_back(vals);
    }
  }
  small_unpack_kernel<<<1, kSmallThreads, 0, st>>>(ua);
  ctx->launches++;
  BGPU_CUDA(ctx, cudaGetLastError());
  uint32_t* pin = (uint32_t*)ctx->pinned + 64;
  BGPU_CUDA(ctx, cudaMemcpyAsync(pin, status->p, 16, cudaMemcpyDeviceToHost, st));
  BGPU_CUDA(ctx, cudaStreamSynchronize(st));  // the one host round trip
  if (pin[1]) return ctx->fail(BGPU_EUNSUPPORTED, "bgpu_small_exchange: some rank's batch does not fit the small path");
  ob->n_rows = pin[0];
  for (size_t c = 0; c < n_cols; c++) {
    Column& oc = ob->cols[c];
    oc.length = ob->n_rows;
    if ((pin[2] >> c) & 1) {
      oc.null_count = -1;
    } else {
      oc.validity = nullptr;  // no rank had nulls in this column
      oc.null_count = 0;
    }
  }
  *out = ob.release();
  return BGPU_OK;
} catch (const std::exception& e) {
  return bgpu::abi_guard(ctx, e);  // nothing is thrown across the C ABI
}
