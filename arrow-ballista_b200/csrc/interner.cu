// Value interning on the device: byte strings (utf8 / binary values, or fixed-width values read as bytes) -> dense Int32 ids.
//
// DataFusion's GroupValuesRows keys its table by the arrow-row encoding of ALL group columns, whatever their width
// (plan shape /root/reference/ballista/scheduler/src/planner.rs:378-380). The aggregate kernels here keep a
// packed key of at most 16 bytes inline in the table slot (one sector per probe). Group columns that do not fit - long
// strings, many wide columns - are first replaced by dense ids from an Interner (hash-and-verify table, key bytes in a side
// arena) and enter the packed key as 4 bytes, like dictionary indices; emit looks the values up again. The same primitive
// folds the dictionaries of dictionary-encoded group columns into one canonical dictionary without leaving the device.
#include <cub/device/device_scan.cuh>

#include "hash.cuh"
#include "interner.h"

namespace bgpu {
namespace {
constexpr int kThreads = 256;
constexpr uint32_t kDead = 0xffffffffu;
constexpr uint32_t kMaxProbes = 512;   // at load <= 1/2 a probe sequence this long does not happen; it means "full"
struct Slot { unsigned long long tag; uint32_t idp1; uint32_t pad; };   // tag 0 = empty; idp1 0 = being published
struct Span { unsigned long long off; uint32_t len; uint32_t pad; };
struct Counters { uint32_t n_ids; uint32_t flags; unsigned long long data_bytes; };  // flags: 1 ids/table full, 2 arena full

struct InternArgs {
  const uint8_t* data;     // utf8 bytes, or fixed-width values
  const int32_t* offsets;  // nullptr: fixed width
  const uint8_t* validity; // nullptr: no nulls
  int32_t width;           // fixed width
  int64_t n;
  Slot* slots;
  uint32_t mask, cap_ids;
  Span* spans;
  uint8_t* arena;
  unsigned long long cap_data;
  Counters* cnt;
  int32_t* out_ids;        // -1 null, -2 not interned (capacity: the host grows and runs the batch again)
  int32_t seed;            // 1: row r gets id r (the values are known to be distinct: a dictionary that becomes canonical)
};

__device__ __forceinline__ bool bytes_equal(const uint8_t* a, const uint8_t* b, uint32_t n) {
  for (uint32_t i = 0; i < n; i++)
    if (a[i] != b[i]) return false;
  return true;
}

__global__ void __launch_bounds__(kThreads) intern_kernel(const __grid_constant__ InternArgs a) {
  for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < a.n; r += (int64_t)gridDim.x * blockDim.x) {
    if (a.validity && !bit_is_set(a.validity, r)) { a.out_ids[r] = -1; continue; }
    const uint8_t* p;
    uint32_t len;
    if (a.offsets) { p = a.data + a.offsets[r]; len = (uint32_t)(a.offsets[r + 1] - a.offsets[r]); }
    else { p = a.data + (size_t)r * a.width; len = (uint32_t)a.width; }
    const uint64_t h = hash_bytes(p, len);
    const unsigned long long tag = h | 1ULL;
    uint32_t i = (uint32_t)(h >> 17) & a.mask;
    int32_t result = -2;
    // once the table or the arena is known to be full the launch is void (the host grows and runs the batch again): later rows do not
    // probe at all, and no row walks more than kMaxProbes slots of a nearly full table
    if (*(volatile uint32_t*)&a.cnt->flags) { a.out_ids[r] = -2; continue; }
    const uint32_t max_probes = a.mask < kMaxProbes ? a.mask : kMaxProbes;
    for (uint32_t probes = 0; probes <= max_probes; probes++, i = (i + 1) & a.mask) {
      Slot* s = a.slots + i;
      unsigned long long t = *(volatile unsigned long long*)&s->tag;
      if (t == 0) {
        t = atomicCAS(&s->tag, 0ULL, tag);
        if (t == 0) {  // this thread owns the slot: arena bytes first, then an id (so that every id handed out below the
                       // capacity has its span), copy, publish
          const unsigned long long off = atomicAdd(&a.cnt->data_bytes, (unsigned long long)len);
          if (off + len > a.cap_data) { atomicOr(&a.cnt->flags, 2u); *(volatile uint32_t*)&s->idp1 = kDead; break; }
          const uint32_t id = a.seed ? (uint32_t)r : atomicAdd(&a.cnt->n_ids, 1u);
          if (id >= a.cap_ids) { atomicOr(&a.cnt->flags, 1u); *(volatile uint32_t*)&s->idp1 = kDead; break; }
          for (uint32_t k = 0; k < len; k++) a.arena[off + k] = p[k];
          a.spans[id] = Span{off, len, 0};
          __threadfence();
          *(volatile uint32_t*)&s->idp1 = id + 1;
          result = (int32_t)id;
          break;
        }
      }
      if (t == tag) {
        uint32_t idp1;
        while ((idp1 = *(volatile uint32_t*)&s->idp1) == 0) __nanosleep(40);  // the owner is publishing
        if (idp1 == kDead) break;  // its owner ran out of room: the flag is up, the batch runs again
        __threadfence();
        const Span sp = a.spans[idp1 - 1];
        if (sp.len == len && bytes_equal(a.arena + sp.off, p, len)) { result = (int32_t)(idp1 - 1); break; }
      }
    }
    if (result == -2) atomicOr(&a.cnt->flags, 1u);
    a.out_ids[r] = result;
  }
}

__global__ void __launch_bounds__(kThreads) rehash_kernel(const Slot* __restrict__ old, uint32_t old_cap, Slot* slots, uint32_t mask) {
  for (uint32_t s = blockIdx.x * blockDim.x + threadIdx.x; s < old_cap; s += gridDim.x * blockDim.x) {
    const Slot o = old[s];
    if (o.tag == 0 || o.idp1 == 0 || o.idp1 == kDead) continue;
    uint32_t i = (uint32_t)(o.tag >> 17) & mask;
    while (atomicCAS(&slots[i].tag, 0ULL, o.tag) != 0ULL) i = (i + 1) & mask;
    slots[i].idp1 = o.idp1;
  }
}

// values of ids -> lengths (+ validity by ballot); then bytes
__global__ void __launch_bounds__(kThreads)
    take_lens_kernel(const int32_t* __restrict__ ids, const uint32_t* __restrict__ in_valid, const Span* __restrict__ spans, int64_t n,
                     int32_t* __restrict__ lens, uint32_t* __restrict__ out_valid) {
  const int64_t n_pad = (n + 31) & ~31LL;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_pad; i += (int64_t)gridDim.x * blockDim.x) {
    bool valid = false;
    if (i < n) {
      valid = (!in_valid || ((in_valid[i >> 5] >> (i & 31)) & 1)) && ids[i] >= 0;
      if (lens) lens[i] = valid ? (int32_t)spans[ids[i]].len : 0;
    }
    if (out_valid) {
      const uint32_t vb = __ballot_sync(0xffffffffu, valid);
      if ((threadIdx.x & 31) == 0) out_valid[i >> 5] = vb;
    }
  }
}
__global__ void __launch_bounds__(kThreads)
    take_bytes_kernel(const int32_t* __restrict__ ids, const Span* __restrict__ spans, const uint8_t* __restrict__ arena,
                      const int32_t* __restrict__ out_offs, int32_t width, uint8_t* __restrict__ out, int64_t n) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int32_t b = out_offs ? out_offs[i] : (int32_t)0;
    const int32_t len = out_offs ? out_offs[i + 1] - b : width;
    uint8_t* dst = out_offs ? out + b : out + (size_t)i * width;
    if (ids[i] < 0 || len <= 0) {
      if (!out_offs) for (int32_t j = 0; j < width; j++) dst[j] = 0;
      continue;
    }
    const uint8_t* src = arena + spans[ids[i]].off;
    for (int32_t j = 0; j < len; j++) dst[j] = src[j];
  }
}
__global__ void __launch_bounds__(kThreads) iota_kernel(int32_t* out, int64_t n) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) out[i] = (int32_t)i;
}
int grid_for(bgpu_ctx* ctx, int64_t n) {
  return (int)std::max<int64_t>(1, std::min<int64_t>(ceil_div64(n, kThreads), (int64_t)ctx->sm_count * 8));
}
}  // namespace

struct Interner {
  bgpu_ctx* ctx = nullptr;
  DevAllocPtr slots, spans, arena, cnt;
  uint32_t cap_slots = 0, cap_ids = 0;
  uint64_t cap_data = 0;
  uint32_t n_ids = 0;
};

int interner_create(bgpu_ctx* ctx, std::shared_ptr<Interner>* out) {
  auto in = std::make_shared<Interner>();
  in->ctx = ctx;
  in->cap_slots = 1u << 16;
  in->cap_ids = in->cap_slots / 2;
  in->cap_data = 1u << 20;
  BGPU_TRY(dev_alloc(ctx, (size_t)in->cap_slots * sizeof(Slot), &in->slots));
  BGPU_TRY(dev_alloc(ctx, (size_t)in->cap_ids * sizeof(Span), &in->spans));
  BGPU_TRY(dev_alloc(ctx, in->cap_data, &in->arena));
  BGPU_TRY(dev_alloc(ctx, sizeof(Counters), &in->cnt));
  BGPU_CUDA(ctx, cudaMemsetAsync(in->slots->p, 0, (size_t)in->cap_slots * sizeof(Slot), ctx->stream));
  BGPU_CUDA(ctx, cudaMemsetAsync(in->cnt->p, 0, sizeof(Counters), ctx->stream));
  *out = in;
  return BGPU_OK;
}

uint32_t interner_size(const Interner* in) { return in->n_ids; }

// Forget every value, keep the capacity (a new stream through the same operator usually needs the same room again).
int interner_clear(Interner* in) {
  bgpu_ctx* ctx = in->ctx;
  BGPU_CUDA(ctx, cudaMemsetAsync(in->slots->p, 0, (size_t)in->cap_slots * sizeof(Slot), ctx->stream));
  BGPU_CUDA(ctx, cudaMemsetAsync(in->cnt->p, 0, sizeof(Counters), ctx->stream));
  in->n_ids = 0;
  return BGPU_OK;
}

// Make room: a table for `want_ids` ids (never smaller than now) and an arena of `want_data` bytes. The table is ALWAYS
// rebuilt: slots whose owner ran out of room (kDead) must not survive into the next attempt.
static int grow(Interner* in, uint64_t want_ids, uint64_t want_data) {
  bgpu_ctx* ctx = in->ctx;
  uint32_t cap = in->cap_slots;
  while ((uint64_t)cap / 2 < want_ids) {
    if (cap >= (1u << 30)) return ctx->fail(BGPU_EUNSUPPORTED, "interner: more than 2^29 distinct values");
    cap <<= 1;
  }
  DevAllocPtr slots, spans = in->spans;
  BGPU_TRY(dev_alloc(ctx, (size_t)cap * sizeof(Slot), &slots));
  BGPU_CUDA(ctx, cudaMemsetAsync(slots->p, 0, (size_t)cap * sizeof(Slot), ctx->stream));
  if (cap != in->cap_slots) {
    BGPU_TRY(dev_alloc(ctx, (size_t)(cap / 2) * sizeof(Span), &spans));
    BGPU_CUDA(ctx, cudaMemcpyAsync(spans->p, in->spans->p, (size_t)in->cap_ids * sizeof(Span), cudaMemcpyDeviceToDevice, ctx->stream));
  }
  rehash_kernel<<<grid_for(ctx, in->cap_slots), kThreads, 0, ctx->stream>>>((const Slot*)in->slots->p, in->cap_slots, (Slot*)slots->p, cap - 1);
  ctx->launches++;
  BGPU_CUDA(ctx, cudaGetLastError());
  in->slots = slots; in->spans = spans; in->cap_slots = cap; in->cap_ids = cap / 2;
  if (want_data > in->cap_data) {
    uint64_t dcap = in->cap_data;
    while (dcap < want_data) dcap <<= 1;
    DevAllocPtr arena;
    BGPU_TRY(dev_alloc(ctx, dcap, &arena));
    BGPU_CUDA(ctx, cudaMemcpyAsync(arena->p, in->arena->p, in->cap_data, cudaMemcpyDeviceToDevice, ctx->stream));
    in->arena = arena; in->cap_data = dcap;
  }
  return BGPU_OK;
}

// ids of the n values of `col` (utf8 / binary, or fixed-width read as bytes); null rows get -1. seed: the values are known
// to be pairwise distinct and the interner is empty: row r gets id r.
int interner_intern(Interner* in, const Column& col, int64_t n, int32_t* d_ids, bool seed) {
  bgpu_ctx* ctx = in->ctx;
  if (n == 0) return BGPU_OK;
  const bool var = col.kind == K_UTF8 || col.kind == K_BINARY;
  if (!var && col.kind != K_FIXED) return ctx->fail(BGPU_EUNSUPPORTED, "interner: column kind");
  if (seed) {
    if (in->n_ids) return ctx->fail(BGPU_EINVAL, "internal: seeding a non-empty interner");
    const uint64_t bytes = var ? (col.data_bytes > 0 ? (uint64_t)col.data_bytes : 0) : (uint64_t)n * (uint64_t)col.width;
    if ((uint64_t)n > in->cap_ids || bytes > in->cap_data) BGPU_TRY(grow(in, (uint64_t)n, bytes));
  }
  Counters* pin = (Counters*)((uint8_t*)ctx->pinned + 1024);
  for (int attempt = 0; attempt < 40; attempt++) {
    InternArgs a{};
    a.data = (const uint8_t*)col.values; a.offsets = var ? col.offsets : nullptr; a.validity = col.validity;
    a.width = col.width; a.n = n;
    a.slots = (Slot*)in->slots->p; a.mask = in->cap_slots - 1; a.cap_ids = in->cap_ids;
    a.spans = (Span*)in->spans->p; a.arena = (uint8_t*)in->arena->p; a.cap_data = in->cap_data;
    a.cnt = (Counters*)in->cnt->p; a.out_ids = d_ids; a.seed = seed ? 1 : 0;
    intern_kernel<<<grid_for(ctx, n), kThreads, 0, ctx->stream>>>(a);
    ctx->launches++;
    BGPU_CUDA(ctx, cudaGetLastError());
    BGPU_CUDA(ctx, cudaMemcpyAsync(pin, in->cnt->p, sizeof(Counters), cudaMemcpyDeviceToHost, ctx->stream));
    BGPU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    const Counters c = *pin;
    if (!c.flags) {
      in->n_ids = seed ? (uint32_t)n : c.n_ids;
      if (seed) {  // ids were not drawn from the counter
        Counters fix = c;
        fix.n_ids = (uint32_t)n;
        *pin = fix;
        BGPU_CUDA(ctx, cudaMemcpyAsync(in->cnt->p, pin, sizeof(Counters), cudaMemcpyHostToDevice, ctx->stream));
        BGPU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
      }
      return BGPU_OK;
    }
    // out of ids / slots / arena: grow for what this batch can still add and run it again (values already interned are
    // found again; ids handed out beyond the capacity were never published)
    Counters reset = c;
    reset.flags = 0;
    reset.n_ids = seed ? 0u : std::min(c.n_ids, in->cap_ids);
    reset.data_bytes = std::min<unsigned long long>(c.data_bytes, in->cap_data);
    *pin = reset;
    BGPU_CUDA(ctx, cudaMemcpyAsync(in->cnt->p, pin, sizeof(Counters), cudaMemcpyHostToDevice, ctx->stream));
    BGPU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    uint64_t want_ids = in->cap_ids;
    if (c.flags & 1)  // x4, or an eighth of the batch if that is more; never more than the batch can add
      want_ids = std::min<uint64_t>(std::max<uint64_t>((uint64_t)in->cap_ids * 4, (uint64_t)reset.n_ids + (uint64_t)n / 8),
                                    (uint64_t)reset.n_ids + (uint64_t)n + 1024);
    const uint64_t want_data = (c.flags & 2) ? std::max<uint64_t>(in->cap_data * 4, c.data_bytes + (1u << 20)) : 0;
    BGPU_TRY(grow(in, want_ids, want_data));
  }
  return ctx->fail(BGPU_EUNSUPPORTED, "interner: could not make room for the batch's values");
}

// The values of ids[0 .. n) as a column shaped like `proto` (utf8 / binary / fixed). ids < 0 and rows whose bit in
// `valid_words` (may be nullptr) is clear are null.
int interner_take(Interner* in, const int32_t* d_ids, const uint32_t* valid_words, int64_t n, const Column& proto, Column* out) {
  bgpu_ctx* ctx = in->ctx;
  cudaStream_t st = ctx->stream;
  *out = Column();
  out->name = proto.name; out->format = proto.format; out->flags = proto.flags | ARROW_FLAG_NULLABLE; out->kind = proto.kind;
  out->width = proto.width; out->length = n; out->null_count = -1;
  const int64_t n1 = std::max<int64_t>(n, 1);
  DevAllocPtr valid;
  BGPU_TRY(dev_alloc(ctx, (size_t)ceil_div64(n1, 32) * 4, &valid));
  out->validity = (uint8_t*)valid->p;
  out->owned.push_back(valid);
  const Span* spans = (const Span*)in->spans->p;
  if (proto.kind == K_UTF8 || proto.kind == K_BINARY) {
    DevAllocPtr lens, offs, data;
    BGPU_TRY(dev_alloc(ctx, (size_t)n1 * 4, &lens));
    BGPU_TRY(dev_alloc(ctx, (size_t)(n + 1) * 4, &offs));
    BGPU_CUDA(ctx, cudaMemsetAsync(offs->p, 0, 4, st));
    int32_t total = 0;
    if (n) {
      take_lens_kernel<<<grid_for(ctx, n), kThreads, 0, st>>>(d_ids, valid_words, spans, n, (int32_t*)lens->p, (uint32_t*)valid->p);
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
      take_bytes_kernel<<<grid_for(ctx, n), kThreads, 0, st>>>(d_ids, spans, (const uint8_t*)in->arena->p, (const int32_t*)offs->p, 0,
                                                             (uint8_t*)data->p, n);
      ctx->launches++;
    }
    out->offsets = (int32_t*)offs->p; out->values = data->p; out->data_bytes = total;
    out->owned.push_back(offs); out->owned.push_back(data);
  } else {
    DevAllocPtr vals;
    BGPU_TRY(dev_alloc(ctx, (size_t)n1 * proto.width, &vals));
    if (n) {
      take_lens_kernel<<<grid_for(ctx, n), kThreads, 0, st>>>(d_ids, valid_words, spans, n, nullptr, (uint32_t*)valid->p);
      take_bytes_kernel<<<grid_for(ctx, n), kThreads, 0, st>>>(d_ids, spans, (const uint8_t*)in->arena->p, nullptr, proto.width,
                                                             (uint8_t*)vals->p, n);
      ctx->launches += 2;
    }
    out->values = vals->p;
    out->owned.push_back(vals);
  }
  BGPU_CUDA(ctx, cudaGetLastError());
  return BGPU_OK;
}

// Every interned value in id order, as a column shaped like `proto` (the canonical dictionary of a dictionary group column).
int interner_export(Interner* in, const Column& proto, Column* out) {
  bgpu_ctx* ctx = in->ctx;
  const int64_t n = in->n_ids;
  DevAllocPtr ids;
  BGPU_TRY(dev_alloc(ctx, (size_t)std::max<int64_t>(n, 1) * 4, &ids));
  if (n) {
    iota_kernel<<<grid_for(ctx, n), kThreads, 0, ctx->stream>>>((int32_t*)ids->p, n);
    ctx->launches++;
  }
  BGPU_TRY(interner_take(in, (const int32_t*)ids->p, nullptr, n, proto, out));
  out->validity = nullptr;  // every interned value exists
  out->null_count = 0;
  return BGPU_OK;
}
}  // namespace bgpu
