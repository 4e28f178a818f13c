// Streamed host input of GpuAggregateExec: bgpu_agg_push_host.
//
// The reference feeds its aggregate stream one RecordBatch of <= 8192 rows at a time
// (/root/reference/ballista/core/src/config.rs:180-182; the stream loop of
// /root/reference/ballista/core/src/execution_plans/shuffle_writer.rs:342-346 drives it), from as many input partitions at
// once as the task has (/root/reference/ballista/core/src/execution_plans/coalesce_tasks.rs:183-216: one spawned task per
// partition). A kernel launch, a copy per buffer and a host round trip for every 8192 rows would cap the operator far
// below PCIe, so the operator's input side is a coalescer:
//
//   push (any thread, concurrently)   each batch is cut into segments of <= 64 Ki rows; per segment and column the calling
//                                     thread takes the value range (one pass) and packs the column into its WIRE form in
//                                     a pinned staging chunk: frame-of-reference integers of 0/1/2/4/8 bytes (Decimal128
//                                     whose values fit 64 bits included), utf8 as narrow lengths (nothing at all when
//                                     every string of the segment has the same length) + bytes, bitmaps re-aligned.
//                                     Only the columns the aggregate reads are packed. TPC-H Q1: 78 B/row of Arrow
//                                     buffers cross PCIe as ~12 B/row.
//   submit (whoever fills a chunk)    ONE cudaMemcpyAsync of the chunk's bytes (+ one of its segment table) on a copy
//                                     stream; on the compute stream, after an event: widen_kernel rebuilds plain Arrow
//                                     buffers in HBM (two alternating sets), then the aggregate's fused kernel runs over
//                                     the whole chunk - launched WITHOUT a host round trip (device-side conditional
//                                     commit, AggAsyncReq); upload of chunk i + 1 overlaps the kernels of chunk i.
//   poll (before a buffer set is reused, and at emit)   the launch's status words are read from pinned memory; a launch
//                                     that did not commit (group capacity, table growth, rung change) is replayed
//                                     synchronously from its still intact device buffers.
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <condition_variable>
#include <pthread.h>
#include <deque>
#include <mutex>

#include "agg_host.h"
#include "ingest_pack.h"

using namespace bgpu;

namespace {
constexpr int kMaxIngestCols = 24;
constexpr uint32_t kSegRows = 65536;  // rows per segment: one CTA widens one (segment, column)
constexpr int kSlots = 4;             // pinned staging chunks
constexpr int kSets = 2;              // device buffer sets (wire + widened Arrow buffers)
constexpr uint64_t kNoValidity = ~0ULL;

struct WireSeg { uint32_t row_start, n_rows; };
struct WireCol {
  uint64_t off;        // wire offset of the packed values / lengths / value bits (16-byte aligned)
  uint64_t off2;       // utf8: wire offset of the string bytes
  uint64_t valid_off;  // wire offset of the validity bits, kNoValidity = no nulls in this segment
  int64_t base;        // frame-of-reference base; utf8: first byte of the segment in the column's data region
  uint32_t code;       // bytes per packed value (0 = constant), bgpu_pack::kRaw = full width
  uint32_t aux;        // utf8: the common string length when code == 0; otherwise total string bytes of the segment
};

struct WidenCol {
  void* values;        // fixed values / bool bits
  int32_t* offsets;    // utf8
  uint8_t* data;       // utf8 bytes
  uint32_t* validity;  // nullptr: no segment of this chunk has nulls in this column
  int32_t kind, width;
};
struct WidenArgs {
  WidenCol c[kMaxIngestCols];
  const uint8_t* wire;
  const uint8_t* header;
  uint32_t rec_bytes;
};

__device__ __forceinline__ uint64_t load_packed(const uint8_t* p, uint32_t code, uint32_t i) {
  switch (code) {
    case 1: return p[i];
    case 2: return ((const uint16_t*)p)[i];
    case 4: return ((const uint32_t*)p)[i];
    case 8: return ((const uint64_t*)p)[i];
    default: return 0;
  }
}
// bits [0, n) of `src` (aligned at bit 0) -> bits [r0, r0 + n) of the zero-initialised bitmap `dst`; src == nullptr: ones
__device__ void put_bits(uint32_t* dst, const uint8_t* src, uint32_t r0, uint32_t n) {
  const uint32_t w0 = r0 >> 5, w1 = (r0 + n + 31) >> 5;
  for (uint32_t w = w0 + threadIdx.x; w < w1; w += blockDim.x) {
    uint32_t bits = 0;
    for (int b = 0; b < 32; b++) {
      const int64_t row = (int64_t)w * 32 + b - (int64_t)r0;
      if (row < 0 || row >= (int64_t)n) continue;
      if (!src || ((src[row >> 3] >> (row & 7)) & 1)) bits |= 1u << b;
    }
    if (bits) atomicOr(&dst[w], bits);  // the edge words are shared with the neighbouring segments
  }
}

// One CTA per (segment, column): wire form -> rows [row_start, row_start + n_rows) of the chunk's Arrow buffers.
__global__ void __launch_bounds__(256) widen_kernel(const __grid_constant__ WidenArgs a) {
  const uint8_t* rec = a.header + (size_t)blockIdx.x * a.rec_bytes;
  const WireSeg seg = *(const WireSeg*)rec;
  const WireCol wc = ((const WireCol*)(rec + sizeof(WireSeg)))[blockIdx.y];
  const WidenCol& co = a.c[blockIdx.y];
  const uint32_t n = seg.n_rows, r0 = seg.row_start;
  if (co.validity) put_bits(co.validity, wc.valid_off == kNoValidity ? nullptr : a.wire + wc.valid_off, r0, n);
  const uint8_t* p = a.wire + wc.off;
  if (co.kind == K_BOOL) {
    put_bits((uint32_t*)co.values, p, r0, n);
  } else if (co.kind == K_UTF8 || co.kind == K_BINARY) {
    __shared__ uint32_t s_warp[8];
    __shared__ uint32_t s_carry;
    uint32_t total;
    if (wc.code == 0) {
      for (uint32_t i = threadIdx.x; i <= n; i += 256) co.offsets[r0 + i] = (int32_t)(wc.base + (int64_t)i * wc.aux);
      total = n * wc.aux;
    } else {
      if (threadIdx.x == 0) s_carry = 0;
      __syncthreads();
      const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
      for (uint32_t i0 = 0; i0 < n; i0 += 256) {
        const uint32_t i = i0 + threadIdx.x;
        const uint32_t len = i < n ? (uint32_t)load_packed(p, wc.code, i) : 0;
        uint32_t x = len;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
          const uint32_t y = __shfl_up_sync(0xffffffffu, x, d);
          if (lane >= d) x += y;
        }
        if (lane == 31) s_warp[warp] = x;
        __syncthreads();
        uint32_t pre = s_carry;
        for (int w = 0; w < warp; w++) pre += s_warp[w];
        if (i < n) co.offsets[r0 + i] = (int32_t)(wc.base + pre + x - len);
        __syncthreads();
        if (threadIdx.x == 255) s_carry = pre + x;
        __syncthreads();
      }
      total = s_carry;
      if (threadIdx.x == 0) co.offsets[r0 + n] = (int32_t)(wc.base + total);
    }
    const uint8_t* src = a.wire + wc.off2;
    uint8_t* dst = co.data + wc.base;
    for (uint32_t i = threadIdx.x; i < total; i += 256) dst[i] = src[i];
  } else if (wc.code == bgpu_pack::kRaw) {
    if (co.width == 16) {
      const uint4* s = (const uint4*)p;
      uint4* d = (uint4*)co.values + r0;
      for (uint32_t i = threadIdx.x; i < n; i += 256) d[i] = s[i];
    } else if (co.width == 8) {
      const uint64_t* s = (const uint64_t*)p;
      uint64_t* d = (uint64_t*)co.values + r0;
      for (uint32_t i = threadIdx.x; i < n; i += 256) d[i] = s[i];
    } else if (co.width == 4) {
      const uint32_t* s = (const uint32_t*)p;
      uint32_t* d = (uint32_t*)co.values + r0;
      for (uint32_t i = threadIdx.x; i < n; i += 256) d[i] = s[i];
    } else {
      uint8_t* d = (uint8_t*)co.values + (size_t)r0 * co.width;
      for (uint32_t i = threadIdx.x; i < n * (uint32_t)co.width; i += 256) d[i] = p[i];
    }
  } else {
    for (uint32_t i = threadIdx.x; i < n; i += 256) {
      const uint64_t v = (uint64_t)wc.base + load_packed(p, wc.code, i);
      switch (co.width) {
        case 16: ((ulonglong2*)co.values)[r0 + i] = make_ulonglong2(v, (uint64_t)((int64_t)v >> 63)); break;
        case 8: ((uint64_t*)co.values)[r0 + i] = v; break;
        case 4: ((uint32_t*)co.values)[r0 + i] = (uint32_t)v; break;
        case 2: ((uint16_t*)co.values)[r0 + i] = (uint16_t)v; break;
        default: ((uint8_t*)co.values)[r0 + i] = (uint8_t)v; break;
      }
    }
  }
}

// ---- host ----------------------------------------------------------------------------------------------------------
struct IngestCol {
  int32_t batch_col = 0;
  ColKind kind = K_FIXED;
  int32_t width = 0;
  bool is_signed = true, is_decimal = false, is_float = false;
};

struct Chunk {
  uint8_t* pinned = nullptr;  // [header: max_segs records][data]; from the context's pool of pinned blocks
  size_t pinned_bytes = 0;
  // cursors, under Ingest::mu
  uint32_t rows = 0, n_segs = 0;
  size_t bytes = 0;
  std::vector<size_t> str_bytes;   // per column: bytes in the utf8 data region so far
  std::vector<uint8_t> has_nulls;  // per column: some segment carries validity bits
  int writers = 0;
  bool sealed = false;
  enum State { FREE, FILLING, IN_FLIGHT } state = FREE;
  cudaEvent_t h2d_done = nullptr;
};

struct DeviceSet {
  DevAllocPtr wire, header;
  std::vector<DevAllocPtr> values, offsets, data, validity;  // per ingest column
  DevAllocPtr status;
  cudaEvent_t widen_done = nullptr;  // the wire / header buffers may be overwritten
  bool used = false;
  // the launch over this set the host has not looked at yet
  AggAsyncReq req;
  std::unique_ptr<bgpu_batch> batch;
  bool pending = false;
};

struct Ingest {
  bgpu_agg* g = nullptr;
  bgpu_ctx* ctx = nullptr;
  const bgpu_pack::Ops* ops = nullptr;
  std::vector<IngestCol> cols;
  uint32_t rec_bytes = 0, max_segs = 0, max_rows = 0;
  size_t header_bytes = 0, data_cap = 0;
  std::mutex mu;
  std::condition_variable cv;
  Chunk chunks[kSlots];
  int cur = -1;  // slot being filled
  std::mutex submit_mu;
  cudaStream_t copy_stream = nullptr;
  DeviceSet sets[kSets];
  uint32_t* h_status = nullptr;  // pinned, kSets x 8 words
  uint64_t seq = 0;
  int sticky_error = BGPU_OK;
  std::string sticky_msg;
  // running value range of every fixed-width column over the segments seen so far: the one-pass packer's guess for the
  // next segment's base and width (verified while packing; any stale or torn pair only costs the exact two-pass route)
  std::atomic<int64_t> run_lo[kMaxIngestCols], run_hi[kMaxIngestCols];
  std::atomic<int32_t> run_set[kMaxIngestCols];
  std::atomic<uint64_t> st_one_pass{0}, st_two_pass{0};
  // statistics
  uint64_t st_rows = 0, st_batches = 0, st_wire_bytes = 0, st_raw_bytes = 0, st_chunks = 0, st_replays = 0, st_async = 0;

  ~Ingest() {
    if (ctx && ctx->has_device) {
      cudaSetDevice(ctx->device);
      if (copy_stream) cudaStreamSynchronize(copy_stream);
      cudaStreamSynchronize(ctx->stream);
    }
    for (Chunk& c : chunks) {
      if (c.pinned) ctx_pinned_put(ctx, c.pinned, c.pinned_bytes);
      if (c.h2d_done) cudaEventDestroy(c.h2d_done);
    }
    for (DeviceSet& s : sets) {
      s.batch.reset();
      for (cudaEvent_t e : {s.widen_done, s.req.k0, s.req.k1, s.req.done})
        if (e) cudaEventDestroy(e);
    }
    if (h_status) cudaFreeHost(h_status);
    if (copy_stream) cudaStreamDestroy(copy_stream);
  }
};

size_t align16(size_t x) { return (x + 15) & ~(size_t)15; }

size_t env_size(const char* name, size_t dflt) {
  const char* e = getenv(name);
  if (!e) return dflt;
  const long long v = atoll(e);
  return v > 0 ? (size_t)v : dflt;
}

// which columns of the raw input the aggregate reads
int used_columns(bgpu_agg* g, std::vector<IngestCol>* out) {
  bgpu_ctx* ctx = g->ctx;
  std::vector<int32_t> idx(g->group_cols.begin(), g->group_cols.end());
  for (const HFilterTerm& t : g->filter) idx.push_back(t.col);
  for (const HValue& v : g->values)
    for (const HFactor& f : v.f) {
      if (f.x.is_col) idx.push_back(f.x.col);
      if (f.op && f.y.is_col) idx.push_back(f.y.col);
    }
  std::sort(idx.begin(), idx.end());
  idx.erase(std::unique(idx.begin(), idx.end()), idx.end());
  if (idx.size() > (size_t)kMaxIngestCols) return ctx->fail(BGPU_EUNSUPPORTED, "aggregate reads more than 24 columns");
  for (int32_t c : idx) {
    if (c < 0 || c >= (int32_t)g->fields.size()) return ctx->fail(BGPU_EINVAL, "aggregate column out of range");
    const Column& f = g->fields[(size_t)c];
    IngestCol ic;
    ic.batch_col = c; ic.kind = f.kind; ic.width = f.width;
    if (f.kind == K_DICT32)
      return ctx->fail(BGPU_EUNSUPPORTED, "bgpu_agg_push_host: dictionary column " + f.name + " (use bgpu_batch_upload + bgpu_agg_update)");
    const char c0 = f.format.empty() ? '?' : f.format[0];
    ic.is_float = c0 == 'f' || c0 == 'g' || c0 == 'e';
    ic.is_decimal = c0 == 'd';
    ic.is_signed = !(c0 == 'C' || c0 == 'S' || c0 == 'I' || c0 == 'L');
    out->push_back(ic);
  }
  return BGPU_OK;
}

// The widest build of the host packer this CPU runs, starting at `from` (0: AVX-512, 1: AVX2, 2: plain).
static const bgpu_pack::Ops* pack_ops_for(int from) {
  if (from <= 0 && __builtin_cpu_supports("avx512f") && __builtin_cpu_supports("avx512bw") && __builtin_cpu_supports("avx512vl") &&
      __builtin_cpu_supports("avx512dq"))
    return &pack_avx512::ops;
  if (from <= 1 && __builtin_cpu_supports("avx2")) return &pack_avx2::ops;
  return &pack_base::ops;
}

// BGPU_INGEST_TRACE=1: one stderr line per state change of the staging pipeline (thread id, chunk slot, sequence number)
static bool ingest_trace() {
  static const bool on = getenv("BGPU_INGEST_TRACE") != nullptr;
  return on;
}
#define INGEST_TRACE(...)                                                                   \
  do {                                                                                      \
    if (ingest_trace()) {                                                                   \
      fprintf(stderr, "[bgpu ingest %lx] ", (unsigned long)pthread_self());                 \
      fprintf(stderr, __VA_ARGS__);                                                         \
      fputc('\n', stderr);                                                                  \
    }                                                                                       \
  } while (0)

int ingest_create(bgpu_agg* g) {
  bgpu_ctx* ctx = g->ctx;
  if (g->final_mode) return ctx->fail(BGPU_EINVAL, "bgpu_agg_push_host takes raw input rows (Partial / Single modes)");
  auto in = std::make_shared<Ingest>();
  in->g = g; in->ctx = ctx;
  for (int c = 0; c < kMaxIngestCols; c++) { in->run_lo[c].store(0); in->run_hi[c].store(0); in->run_set[c].store(0); }
  in->ops = pack_ops_for(getenv("BGPU_INGEST_NO_AVX512") ? 1 : 0);
  if (getenv("BGPU_INGEST_NO_AVX2")) in->ops = &pack_base::ops;
  BGPU_TRY(used_columns(g, &in->cols));
  const size_t n_cols = in->cols.size();
  in->data_cap = env_size("BGPU_INGEST_CHUNK_MB", 32) << 20;
  in->max_rows = (uint32_t)std::min<size_t>(env_size("BGPU_INGEST_CHUNK_ROWS", 4u << 20), 1u << 30);
  in->max_segs = (uint32_t)env_size("BGPU_INGEST_CHUNK_SEGS", 1024);
  in->rec_bytes = (uint32_t)(sizeof(WireSeg) + n_cols * sizeof(WireCol));
  in->header_bytes = align16((size_t)in->max_segs * in->rec_bytes);
  BGPU_CUDA(ctx, cudaStreamCreateWithFlags(&in->copy_stream, cudaStreamNonBlocking));
  for (Chunk& c : in->chunks) {
    BGPU_TRY(ctx_pinned_get(ctx, in->header_bytes + in->data_cap, (void**)&c.pinned));
    c.pinned_bytes = in->header_bytes + in->data_cap;
    BGPU_CUDA(ctx, cudaEventCreateWithFlags(&c.h2d_done, cudaEventDisableTiming));
    c.str_bytes.assign(n_cols, 0);
    c.has_nulls.assign(n_cols, 0);
  }
  BGPU_CUDA(ctx, cudaHostAlloc((void**)&in->h_status, kSets * 32, cudaHostAllocDefault));
  for (int s = 0; s < kSets; s++) {
    DeviceSet& ds = in->sets[s];
    BGPU_TRY(dev_alloc(ctx, in->data_cap, &ds.wire));
    BGPU_TRY(dev_alloc(ctx, in->header_bytes, &ds.header));
    BGPU_TRY(dev_alloc(ctx, 32, &ds.status));
    ds.values.resize(n_cols); ds.offsets.resize(n_cols); ds.data.resize(n_cols); ds.validity.resize(n_cols);
    for (size_t c = 0; c < n_cols; c++) {
      const IngestCol& ic = in->cols[c];
      const size_t bitmap = ((size_t)in->max_rows + 31) / 32 * 4 + 16;
      BGPU_TRY(dev_alloc(ctx, bitmap, &ds.validity[c]));
      if (ic.kind == K_UTF8 || ic.kind == K_BINARY) {
        BGPU_TRY(dev_alloc(ctx, ((size_t)in->max_rows + 1) * 4, &ds.offsets[c]));
        BGPU_TRY(dev_alloc(ctx, in->data_cap + 64, &ds.data[c]));  // 32 bytes of slack either side: keys are read as aligned words
      } else if (ic.kind == K_BOOL) {
        BGPU_TRY(dev_alloc(ctx, bitmap, &ds.values[c]));
      } else {
        BGPU_TRY(dev_alloc(ctx, (size_t)in->max_rows * (size_t)ic.width + 16, &ds.values[c]));
      }
    }
    BGPU_CUDA(ctx, cudaEventCreateWithFlags(&ds.widen_done, cudaEventDisableTiming));
    BGPU_CUDA(ctx, cudaEventCreate(&ds.req.k0));
    BGPU_CUDA(ctx, cudaEventCreate(&ds.req.k1));
    BGPU_CUDA(ctx, cudaEventCreateWithFlags(&ds.req.done, cudaEventDisableTiming));
    ds.req.d_status = (uint32_t*)ds.status->p;
    ds.req.h_status = in->h_status + s * 8;
  }
  g->ingest = in;
  return BGPU_OK;
}

// Take the outcome of the launch over set `s` (if any); replay it when it did not commit. Under submit_mu.
int poll_set(Ingest* in, int s) {
  DeviceSet& ds = in->sets[s];
  if (!ds.pending) return BGPU_OK;
  ds.pending = false;
  bool replay = false;
  INGEST_TRACE("poll set %d", s);
  BGPU_TRY(agg_async_result(in->g, &ds.req, &replay));
  INGEST_TRACE("set %d: %s", s, replay ? "replay" : "committed");
  if (replay) {
    in->st_replays++;
    BGPU_TRY(agg_update_replay(in->g, ds.batch.get()));
  }
  return BGPU_OK;
}
int poll_all(Ingest* in, uint64_t next_seq) {
  // oldest first: a launch that did not commit left its flag up, so every later one has to be replayed after it
  for (int k = kSets; k >= 1; k--)
    if (next_seq >= (uint64_t)k) BGPU_TRY(poll_set(in, (int)((next_seq - k) % kSets)));
  return BGPU_OK;
}

// Upload + widen + aggregate one sealed chunk. Under submit_mu; the chunk's cursors are final.
int submit_chunk(Ingest* in, Chunk* ch) {
  bgpu_ctx* ctx = in->ctx;
  bgpu_agg* g = in->g;
  BGPU_CUDA(ctx, cudaSetDevice(ctx->device));
  const int s = (int)(in->seq % kSets);
  DeviceSet& ds = in->sets[s];
  INGEST_TRACE("submit slot %d as seq %llu (rows %u) into set %d", (int)(ch - in->chunks), (unsigned long long)in->seq, ch->rows, s);
  // this set's previous launch (the oldest one pending) must be known good, or replayed from the set's buffers, before
  // they are overwritten
  BGPU_TRY(poll_set(in, s));
  const size_t n_cols = in->cols.size();
  const uint32_t rows = ch->rows, n_segs = ch->n_segs;
  if (ds.used) BGPU_CUDA(ctx, cudaStreamWaitEvent(in->copy_stream, ds.widen_done, 0));
  {
    GpuTimeScope timed_copy(ctx, T_H2D, in->copy_stream);
    BGPU_CUDA(ctx, cudaMemcpyAsync(ds.header->p, ch->pinned, (size_t)n_segs * in->rec_bytes, cudaMemcpyHostToDevice, in->copy_stream));
    if (ch->bytes)
      BGPU_CUDA(ctx, cudaMemcpyAsync(ds.wire->p, ch->pinned + in->header_bytes, ch->bytes, cudaMemcpyHostToDevice, in->copy_stream));
  }
  BGPU_CUDA(ctx, cudaEventRecord(ch->h2d_done, in->copy_stream));
  GpuTimeScope timed(ctx, T_KERNEL);  // widen + the aggregate kernels of this chunk
  BGPU_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, ch->h2d_done, 0));
  in->st_chunks++;
  in->st_wire_bytes += ch->bytes + (size_t)n_segs * in->rec_bytes;

  WidenArgs wa{};
  wa.wire = (const uint8_t*)ds.wire->p;
  wa.header = (const uint8_t*)ds.header->p;
  wa.rec_bytes = in->rec_bytes;
  std::unique_ptr<bgpu_batch> batch(new bgpu_batch(ctx));
  batch->n_rows = rows;
  batch->cols.resize(g->fields.size());
  for (size_t i = 0; i < g->fields.size(); i++) {  // columns the aggregate does not read stay empty shells
    Column& c = batch->cols[i];
    const Column& f = g->fields[i];
    c.name = f.name; c.format = f.format; c.flags = f.flags; c.kind = f.kind; c.width = f.width; c.length = rows;
  }
  const size_t bitmap_bytes = ((size_t)rows + 31) / 32 * 4;
  bool any_nulls = false;
  for (size_t c = 0; c < n_cols; c++) {
    const IngestCol& ic = in->cols[c];
    WidenCol& wc = wa.c[c];
    Column& col = batch->cols[(size_t)ic.batch_col];
    wc.kind = ic.kind; wc.width = ic.width;
    if (ch->has_nulls[c]) {
      BGPU_CUDA(ctx, cudaMemsetAsync(ds.validity[c]->p, 0, bitmap_bytes, ctx->stream));
      wc.validity = (uint32_t*)ds.validity[c]->p;
      col.validity = (uint8_t*)ds.validity[c]->p;
      col.null_count = -1;
      any_nulls = true;
    }
    if (ic.kind == K_UTF8 || ic.kind == K_BINARY) {
      wc.offsets = (int32_t*)ds.offsets[c]->p;
      wc.data = (uint8_t*)ds.data[c]->p + 32;
      col.offsets = wc.offsets; col.values = wc.data; col.data_bytes = (int64_t)ch->str_bytes[c];
    } else {
      wc.values = ds.values[c]->p;
      col.values = wc.values;
      if (ic.kind == K_BOOL) BGPU_CUDA(ctx, cudaMemsetAsync(wc.values, 0, bitmap_bytes, ctx->stream));
    }
  }
  if (rows) {
    widen_kernel<<<dim3(n_segs, (unsigned)n_cols), 256, 0, ctx->stream>>>(wa);
    ctx->launches++;
    BGPU_CUDA(ctx, cudaGetLastError());
  }
  BGPU_CUDA(ctx, cudaEventRecord(ds.widen_done, ctx->stream));
  ds.used = true;
  // the pinned chunk is free again once its copies have completed (the producer that takes the slot waits on h2d_done)
  {
    std::lock_guard<std::mutex> lk(in->mu);
    ch->state = Chunk::IN_FLIGHT;
  }
  in->cv.notify_all();
  // the first batch with nulls sets up the per-group "seen" masks from the host's group count: it must be current
  if (any_nulls && !g->seen_tracking) BGPU_TRY(poll_all(in, in->seq));
  ds.batch = std::move(batch);
  const bool sync_only = getenv("BGPU_INGEST_SYNC") != nullptr;
  BGPU_TRY(agg_update_impl(g, ds.batch.get(), sync_only ? nullptr : &ds.req));
  ds.pending = ds.req.launched;
  in->st_async += ds.pending;
  in->seq++;
  INGEST_TRACE("submitted seq %llu (async %d)", (unsigned long long)in->seq - 1, (int)ds.pending);
  return BGPU_OK;
}

void fail_sticky(Ingest* in, int rc) {
  std::lock_guard<std::mutex> lk(in->mu);
  if (in->sticky_error == BGPU_OK) {
    in->sticky_error = rc;
    in->sticky_msg = in->ctx->err;
  }
}

// Seal + submit what is being filled, then look at every launch still pending. Called by emit / update / flush.
int ingest_drain(bgpu_agg* g) {
  Ingest* in = (Ingest*)g->ingest.get();
  if (!in) return BGPU_OK;
  bgpu_ctx* ctx = g->ctx;
  Chunk* ch = nullptr;
  {
    std::unique_lock<std::mutex> lk(in->mu);
    if (in->sticky_error != BGPU_OK) return ctx->fail(in->sticky_error, in->sticky_msg);
    if (in->cur >= 0) {
      Chunk& c = in->chunks[in->cur];
      in->cv.wait(lk, [&] { return c.writers == 0; });  // pushes still packing (the caller should have joined them)
      c.sealed = true;
      in->cur = -1;
      if (c.state == Chunk::FILLING) ch = &c;
    }
  }
  std::lock_guard<std::mutex> sl(in->submit_mu);
  if (ch && ch->rows) BGPU_TRY(submit_chunk(in, ch));
  else if (ch) {
    std::lock_guard<std::mutex> lk(in->mu);
    ch->state = Chunk::FREE;
  }
  BGPU_TRY(poll_all(in, in->seq));
  return BGPU_OK;
}

void ingest_reset(bgpu_agg* g) {
  Ingest* in = (Ingest*)g->ingest.get();
  if (!in) return;
  std::lock_guard<std::mutex> lk(in->mu);
  in->sticky_error = BGPU_OK;
  in->sticky_msg.clear();
  for (DeviceSet& s : in->sets) s.pending = false;
  in->st_rows = in->st_batches = in->st_wire_bytes = in->st_raw_bytes = in->st_chunks = in->st_replays = in->st_async = 0;
  in->st_one_pass.store(0);
  in->st_two_pass.store(0);
}

struct SegPlan {  // what pass 1 found for one column of one segment
  bgpu_pack::FixedStats fs{0, 0, 0, 0, 0};
  const uint8_t* packed = nullptr;  // the column already packed by the one-pass route (thread-local scratch)
  size_t bytes = 0, bytes2 = 0, valid_bytes = 0;  // wire bytes: packed values / lengths, string bytes, validity bits
  int32_t str_first = 0;
  bool has_validity = false;
};

int push_segment(Ingest* in, const struct ArrowArray* batch, int64_t lo, uint32_t n) {
  bgpu_ctx* ctx = in->ctx;
  const size_t n_cols = in->cols.size();
  SegPlan plan[kMaxIngestCols];
  size_t need = 0;
  static const bool one_pass = getenv("BGPU_INGEST_TWO_PASS") == nullptr;  // (measurement knob: always range + pack)
  static thread_local std::vector<uint8_t> scratch;  // one-pass packed columns of this segment, n x 4 bytes each
  if (one_pass && scratch.size() < n_cols * (size_t)n * 4) scratch.resize(n_cols * (size_t)n * 4);
  for (size_t c = 0; c < n_cols; c++) {  // pass 1: ranges -> wire widths
    const IngestCol& ic = in->cols[c];
    const struct ArrowArray* a = batch->children[ic.batch_col];
    SegPlan& sp = plan[c];
    const int64_t off = a->offset + lo;
    sp.has_validity = a->null_count != 0 && a->n_buffers > 0 && a->buffers[0] != nullptr;
    if (sp.has_validity) sp.valid_bytes = align16((size_t)((n + 7) / 8));
    if (ic.kind == K_UTF8 || ic.kind == K_BINARY) {
      if (a->n_buffers < 3 || !a->buffers[1]) return ctx->fail(BGPU_EINVAL, "utf8 column needs 3 buffers");
      const int32_t* offs = (const int32_t*)a->buffers[1] + off;
      int32_t mn = 0, mx = 0;
      int mono = 1;
      in->ops->utf8_stats(offs, n, &mn, &mx, &mono);
      if (!mono || offs[0] < 0) return ctx->fail(BGPU_EINVAL, "utf8 offsets are not monotonic");
      sp.str_first = offs[0];
      sp.bytes2 = (size_t)(offs[n] - offs[0]);
      sp.fs.code = mn == mx ? 0 : mx < 256 ? 1 : mx < 65536 ? 2 : 4;
      sp.fs.base = mn;
      sp.bytes = align16((size_t)n * sp.fs.code);
    } else if (ic.kind == K_BOOL) {
      sp.bytes = align16((size_t)((n + 7) / 8));
    } else {
      if (a->n_buffers < 2 || !a->buffers[1]) return ctx->fail(BGPU_EINVAL, "primitive column needs 2 buffers");
      const uint8_t* v = (const uint8_t*)a->buffers[1] + (size_t)off * ic.width;
      // one pass when the column's running range says how it will pack: pack into scratch against that base and width
      // and verify on the way (the input is read from DRAM once instead of once + once from cache)
      if (one_pass && !ic.is_float && in->run_set[c].load(std::memory_order_acquire) == 1) {
        const int64_t rlo = in->run_lo[c].load(std::memory_order_relaxed), rhi = in->run_hi[c].load(std::memory_order_relaxed);
        const uint64_t r = (uint64_t)rhi - (uint64_t)rlo;
        const uint32_t code = rhi < rlo ? 8 : r < (1ULL << 8) ? 1 : r < (1ULL << 16) ? 2 : r < (1ULL << 32) ? 4 : 8;
        uint8_t* dst = scratch.data() + c * (size_t)n * 4;
        if (code <= 4 && (int)code < ic.width && in->ops->fixed_pack_checked(v, n, ic.width, ic.is_signed, ic.is_decimal, rlo, code, dst)) {
          sp.fs.base = rlo;
          sp.fs.code = code;
          sp.packed = dst;
          in->st_one_pass.fetch_add(1, std::memory_order_relaxed);
        }
      }
      if (!sp.packed) {
        in->ops->fixed_stats(v, n, ic.width, ic.is_signed, ic.is_decimal, ic.is_float, &sp.fs);
        in->st_two_pass.fetch_add(1, std::memory_order_relaxed);
        if (sp.fs.has_range) {  // widen the running range (lock-free min / max)
          if (!in->run_set[c].load(std::memory_order_acquire)) {
            int32_t expect = 0;
            if (in->run_set[c].compare_exchange_strong(expect, 2)) {  // first writer initialises, then publishes
              in->run_lo[c].store(sp.fs.lo, std::memory_order_relaxed);
              in->run_hi[c].store(sp.fs.hi, std::memory_order_relaxed);
              in->run_set[c].store(1, std::memory_order_release);
            }
          }
          if (in->run_set[c].load(std::memory_order_acquire) == 1) {
            int64_t cur = in->run_lo[c].load(std::memory_order_relaxed);
            while (sp.fs.lo < cur && !in->run_lo[c].compare_exchange_weak(cur, sp.fs.lo, std::memory_order_relaxed)) {}
            cur = in->run_hi[c].load(std::memory_order_relaxed);
            while (sp.fs.hi > cur && !in->run_hi[c].compare_exchange_weak(cur, sp.fs.hi, std::memory_order_relaxed)) {}
          }
        }
      }
      sp.bytes = align16((size_t)n * (sp.fs.code == bgpu_pack::kRaw ? (size_t)ic.width : sp.fs.code));
    }
    need += sp.bytes + align16(sp.bytes2) + sp.valid_bytes;
  }
  if (need > in->data_cap) return ctx->fail(BGPU_EUNSUPPORTED, "bgpu_agg_push_host: one segment exceeds the staging chunk (use bgpu_batch_upload)");

  // reserve rows + bytes in the chunk being filled; a full chunk is sealed and the next slot taken
  Chunk* ch = nullptr;
  Chunk* to_submit = nullptr;
  uint32_t seg_idx = 0, row_start = 0;
  size_t byte_start = 0;
  int64_t str_start[kMaxIngestCols];
  {
    std::unique_lock<std::mutex> lk(in->mu);
    for (;;) {
      if (in->sticky_error != BGPU_OK) return ctx->fail(in->sticky_error, in->sticky_msg);
      if (in->cur >= 0) {
        Chunk& c = in->chunks[in->cur];
        if (c.rows + n <= in->max_rows && c.n_segs < in->max_segs && c.bytes + need <= in->data_cap) { ch = &c; break; }
        c.sealed = true;
        INGEST_TRACE("seal slot %d (rows %u, writers %d)", in->cur, c.rows, c.writers);
        if (c.writers == 0) to_submit = &c;  // otherwise the last writer to finish submits it
        in->cur = -1;
        if (to_submit) break;  // submit first (outside the lock), then come back for a slot
      }
      int free_slot = -1;
      for (int i = 0; i < kSlots && free_slot < 0; i++)
        if (in->chunks[i].state == Chunk::FREE) free_slot = i;
      for (int i = 0; i < kSlots && free_slot < 0; i++)
        if (in->chunks[i].state == Chunk::IN_FLIGHT) free_slot = i;
      if (free_slot < 0) {
        INGEST_TRACE("no free slot: wait");
        in->cv.wait(lk);
        continue;
      }
      Chunk& c = in->chunks[free_slot];
      if (c.state == Chunk::IN_FLIGHT) {  // its copies must have left the pinned memory
        lk.unlock();
        INGEST_TRACE("slot %d in flight: wait for its copies", free_slot);
        const cudaError_t e = cudaEventSynchronize(c.h2d_done);
        INGEST_TRACE("slot %d copies done", free_slot);
        lk.lock();
        if (e != cudaSuccess) return ctx->fail(BGPU_ECUDA, std::string("cudaEventSynchronize: ") + cudaGetErrorString(e));
        if (c.state != Chunk::IN_FLIGHT || in->cur >= 0) continue;  // somebody else took it meanwhile
      }
      INGEST_TRACE("open slot %d", free_slot);
      c.state = Chunk::FILLING;
      c.rows = 0; c.n_segs = 0; c.bytes = 0; c.writers = 0; c.sealed = false;
      std::fill(c.str_bytes.begin(), c.str_bytes.end(), 0);
      std::fill(c.has_nulls.begin(), c.has_nulls.end(), 0);
      in->cur = free_slot;
    }
    if (ch) {
      seg_idx = ch->n_segs++;
      row_start = ch->rows;
      ch->rows += n;
      byte_start = ch->bytes;
      ch->bytes += need;
      for (size_t c = 0; c < n_cols; c++) {
        str_start[c] = (int64_t)ch->str_bytes[c];
        ch->str_bytes[c] += plan[c].bytes2;
        if (plan[c].has_validity) ch->has_nulls[c] = 1;
      }
      ch->writers++;
      in->st_rows += n;
    }
  }
  if (!ch) {  // the chunk this thread sealed had no writers left: submit it, then retry this segment
    int rc;
    {  // (the retry below may have to submit again: submit_mu must be free by then)
      std::lock_guard<std::mutex> sl(in->submit_mu);
      rc = submit_chunk(in, to_submit);
    }
    if (rc != BGPU_OK) { fail_sticky(in, rc); return rc; }
    return push_segment(in, batch, lo, n);
  }

  // pass 2: pack (no lock held)
  uint8_t* data = ch->pinned + in->header_bytes;
  uint8_t* rec = ch->pinned + (size_t)seg_idx * in->rec_bytes;
  WireSeg ws{row_start, n};
  memcpy(rec, &ws, sizeof(ws));
  size_t cursor = byte_start;
  size_t raw = 0;
  for (size_t c = 0; c < n_cols; c++) {
    const IngestCol& ic = in->cols[c];
    const struct ArrowArray* a = batch->children[ic.batch_col];
    const SegPlan& sp = plan[c];
    const int64_t off = a->offset + lo;
    WireCol wc{};
    wc.valid_off = kNoValidity;
    wc.code = sp.fs.code;
    wc.base = sp.fs.base;
    wc.off = cursor;
    if (ic.kind == K_UTF8 || ic.kind == K_BINARY) {
      const int32_t* offs = (const int32_t*)a->buffers[1] + off;
      in->ops->utf8_pack_lens(offs, n, sp.fs.code, data + cursor);
      cursor += sp.bytes;
      wc.off2 = cursor;
      if (sp.bytes2) memcpy(data + cursor, (const uint8_t*)a->buffers[2] + sp.str_first, sp.bytes2);
      cursor += align16(sp.bytes2);
      wc.aux = sp.fs.code == 0 ? (uint32_t)sp.fs.base : (uint32_t)sp.bytes2;
      wc.base = str_start[c];
      raw += (size_t)(n + 1) * 4 + sp.bytes2;
    } else if (ic.kind == K_BOOL) {
      in->ops->bits_realign((const uint8_t*)a->buffers[1], off, n, data + cursor);
      cursor += sp.bytes;
      raw += (size_t)((n + 7) / 8);
    } else {
      const uint8_t* v = (const uint8_t*)a->buffers[1] + (size_t)off * ic.width;
      if (sp.packed) memcpy(data + cursor, sp.packed, (size_t)n * sp.fs.code);
      else in->ops->fixed_pack(v, n, ic.width, ic.is_signed, ic.is_decimal, sp.fs.base, sp.fs.code, data + cursor);
      cursor += sp.bytes;
      raw += (size_t)n * ic.width;
    }
    if (sp.has_validity) {
      wc.valid_off = cursor;
      in->ops->bits_realign((const uint8_t*)a->buffers[0], off, n, data + cursor);
      cursor += sp.valid_bytes;
      raw += (size_t)((n + 7) / 8);
    }
    memcpy(rec + sizeof(WireSeg) + c * sizeof(WireCol), &wc, sizeof(wc));
  }

  bool submit = false;
  {
    std::lock_guard<std::mutex> lk(in->mu);
    in->st_raw_bytes += raw;
    ch->writers--;
    submit = ch->sealed && ch->writers == 0 && ch->state == Chunk::FILLING;
  }
  in->cv.notify_all();
  if (submit) {
    std::lock_guard<std::mutex> sl(in->submit_mu);
    const int rc = submit_chunk(in, ch);
    if (rc != BGPU_OK) { fail_sticky(in, rc); return rc; }
  }
  return BGPU_OK;
}
}  // namespace

extern "C" int bgpu_agg_push_host(bgpu_agg* g, const struct ArrowArray* batch, const struct ArrowSchema* schema) try {
  if (!g || !batch) return BGPU_EINVAL;
  bgpu_ctx* ctx = g->ctx;
  if (g->emitted) return ctx->fail(BGPU_EINVAL, "aggregate already emitted");
  if (!ctx->has_device) return ctx->fail(BGPU_ENODEVICE, "context has no device (libbgpu has no CPU fallback)");
  BGPU_CUDA(ctx, cudaSetDevice(ctx->device));
  if (!g->ingest) {
    static std::mutex create_mu;
    std::lock_guard<std::mutex> lk(create_mu);
    if (!g->ingest) {
      BGPU_TRY(ingest_create(g));
      g->ingest_drain = ingest_drain;
      g->ingest_reset = ingest_reset;
    }
  }
  Ingest* in = (Ingest*)g->ingest.get();
  if (batch->n_children < (int64_t)g->fields.size()) return ctx->fail(BGPU_EINVAL, "batch has too few columns");
  if (batch->offset != 0) return ctx->fail(BGPU_EINVAL, "struct-level offset is not supported");
  if (schema) {  // optional: check the columns the aggregate reads against input_schema
    if (!schema->format || strcmp(schema->format, "+s") != 0 || schema->n_children != batch->n_children)
      return ctx->fail(BGPU_EINVAL, "expected a struct-typed (RecordBatch) ArrowSchema matching the batch");
    for (const IngestCol& ic : in->cols) {
      const struct ArrowSchema* cs = schema->children[ic.batch_col];
      const Column& f = g->fields[(size_t)ic.batch_col];
      if (cs->dictionary || !cs->format || f.format != cs->format)
        return ctx->fail(BGPU_EINVAL, "column " + f.name + " does not match input_schema");
    }
  }
  for (const IngestCol& ic : in->cols)
    if (batch->children[ic.batch_col]->length != batch->length) return ctx->fail(BGPU_EINVAL, "column length mismatch");
  {
    std::lock_guard<std::mutex> lk(in->mu);
    in->st_batches++;
  }
  for (int64_t lo = 0; lo < batch->length; lo += kSegRows)
    BGPU_TRY(push_segment(in, batch, lo, (uint32_t)std::min<int64_t>(kSegRows, batch->length - lo)));
  return BGPU_OK;
} catch (const std::exception& e) {
  return bgpu::abi_guard(g ? g->ctx : nullptr, e);  // nothing is thrown across the C ABI
}

extern "C" int bgpu_agg_flush(bgpu_agg* g) try {
  if (!g) return BGPU_EINVAL;
  BGPU_CUDA(g->ctx, cudaSetDevice(g->ctx->device));
  return g->ingest_drain ? g->ingest_drain(g) : BGPU_OK;
} catch (const std::exception& e) {
  return bgpu::abi_guard(g ? g->ctx : nullptr, e);  // nothing is thrown across the C ABI
}

extern "C" int bgpu_agg_ingest_stats(const bgpu_agg* g, uint64_t* out8) {
  if (!g || !out8) return BGPU_EINVAL;
  memset(out8, 0, 8 * sizeof(uint64_t));
  Ingest* in = (Ingest*)g->ingest.get();
  if (!in) return BGPU_OK;
  std::lock_guard<std::mutex> lk(in->mu);
  out8[0] = in->st_rows; out8[1] = in->st_batches; out8[2] = in->st_wire_bytes; out8[3] = in->st_raw_bytes;
  out8[4] = in->st_chunks; out8[5] = in->st_async; out8[6] = in->st_replays; out8[7] = in->st_one_pass.load();
  return BGPU_OK;
}

// test hook (host only, no device): the wire form of one fixed-width column segment, as push packs it. format = the
// column's Arrow C format string; out_packed has room for n * 16 bytes. which = 0: the AVX2 build, 1: the plain one,
// 2: the AVX-512 build (BGPU_EUNSUPPORTED when the CPU cannot run the one asked for).
// test hook (host only): the one-pass route of the packer - pack `values` against a GIVEN base and code (1, 2 or 4) and
// say whether every value fitted. Returns 1 (fits, out_packed valid), 0 (does not fit), or a negative error.
extern "C" int bgpu_debug_pack_checked_host(const char* format, const void* values, int64_t n, int32_t which, int64_t base,
                                            uint32_t code, uint8_t* out_packed) {
  if (!format || !values || !out_packed || n < 0) return -BGPU_EINVAL;
  if (which == 0 && !__builtin_cpu_supports("avx2")) return -BGPU_EUNSUPPORTED;
  if (which == 2 && pack_ops_for(0) != &pack_avx512::ops) return -BGPU_EUNSUPPORTED;
  const bgpu_pack::Ops* ops = which == 2 ? &pack_avx512::ops : which == 0 ? &pack_avx2::ops : &pack_base::ops;
  const char c0 = format[0];
  int width = 0;
  switch (c0) {
    case 'c': case 'C': width = 1; break;
    case 's': case 'S': width = 2; break;
    case 'i': case 'I': width = 4; break;
    case 'l': case 'L': width = 8; break;
    case 'd': width = 16; break;
    case 't': width = (format[1] == 'd' && format[2] == 'D') ? 4 : 8; break;
    default: return -BGPU_EUNSUPPORTED;
  }
  const bool is_signed = !(c0 == 'C' || c0 == 'S' || c0 == 'I' || c0 == 'L');
  return ops->fixed_pack_checked(values, n, width, is_signed, c0 == 'd', base, code, out_packed) ? 1 : 0;
}

extern "C" int bgpu_debug_pack_fixed_host(const char* format, const void* values, int64_t n, int32_t which, int64_t* out_base,
                                          uint32_t* out_code, uint8_t* out_packed) {
  if (!format || !values || !out_base || !out_code || !out_packed || n < 0) return BGPU_EINVAL;
  if (which == 0 && !__builtin_cpu_supports("avx2")) return BGPU_EUNSUPPORTED;
  if (which == 2 && pack_ops_for(0) != &pack_avx512::ops) return BGPU_EUNSUPPORTED;
  const bgpu_pack::Ops* ops = which == 2 ? &pack_avx512::ops : which == 0 ? &pack_avx2::ops : &pack_base::ops;
  const char c0 = format[0];
  int width = 0;
  switch (c0) {
    case 'c': case 'C': width = 1; break;
    case 's': case 'S': width = 2; break;
    case 'i': case 'I': case 'f': width = 4; break;
    case 'l': case 'L': case 'g': width = 8; break;
    case 'd': width = 16; break;
    case 't': width = (format[1] == 'd' && format[2] == 'D') ? 4 : 8; break;
    default: return BGPU_EUNSUPPORTED;
  }
  const bool is_float = c0 == 'f' || c0 == 'g', is_dec = c0 == 'd';
  const bool is_signed = !(c0 == 'C' || c0 == 'S' || c0 == 'I' || c0 == 'L');
  bgpu_pack::FixedStats fs{0, 0, 0, 0, 0};
  ops->fixed_stats(values, n, width, is_signed, is_dec, is_float, &fs);
  ops->fixed_pack(values, n, width, is_signed, is_dec, fs.base, fs.code, out_packed);
  *out_base = fs.base;
  *out_code = fs.code;
  return BGPU_OK;
}
