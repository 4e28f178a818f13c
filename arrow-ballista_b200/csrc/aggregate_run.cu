// K4/K5 host side, part 2: binding a batch to the kernel plans, choosing the path, update and emit.
//
// The reference's stream does, per batch: evaluate group and argument expressions, intern keys -> dense group ids,
// then one update per accumulator. Here (kernels in agg_kernels.cuh):
//   fused path     one kernel does all of that per row with thread-private shared-memory accumulators; used while
//                  the number of groups fits the private capacity (a ladder of three configurations), committed by a
//                  second tiny kernel. Aborts cleanly (nothing committed) when a CTA meets too many groups.
//   two-pass path  intern_kernel (idempotent; re-run after growing the table on overflow) then accumulate_kernel
//                  with replicated shared-memory atomics or global atomics.
// Int64/Decimal128 sums are wrapping two's-complement adds, so any order is bit-exact; Float64 sums carry the north
// star's 1e-9 tolerance (the fused path is reproducible run to run, the atomic path is not).
#include <stdio.h>
#include <stdlib.h>
#include <math.h>
#include <string.h>

#include <algorithm>
#include <chrono>
#include <mutex>
#include <tuple>
#include <unordered_map>

#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>

#include "agg_host.h"
#include "agg_kernels.cuh"
#include "interner.h"
#include "jit.h"

using namespace bgpu;

namespace {
uint32_t next_pow2(uint64_t x) {
  uint64_t p = 1;
  while (p < x) p <<= 1;
  return (uint32_t)std::min<uint64_t>(p, 1ULL << 31);
}
int vk_width(int vk) {
  switch (vk) {
    case V_I128: return 16;
    case V_I64: case V_U64: case V_F64: return 8;
    case V_I32: case V_U32: case V_F32: return 4;
    case V_I16: case V_U16: return 2;
    default: return 1;
  }
}
bool same_shape(const Column& a, const Column& b) { return a.kind == b.kind && a.width == b.width; }
}  // namespace

int agg_resize(bgpu_agg* g, uint32_t new_max, uint32_t keep) {
  bgpu_ctx* ctx = g->ctx;
  const size_t slot_bytes = g->key_words == 1 ? 16 : 32;
  DevAllocPtr slots, acc, seen;
  BGPU_TRY(dev_alloc(ctx, (size_t)new_max * 2 * slot_bytes, &slots));
  BGPU_CUDA(ctx, cudaMemsetAsync(slots->p, 0, (size_t)new_max * 2 * slot_bytes, ctx->stream));
  const size_t rec = (size_t)g->acc_stride * 8, keep_bytes = (size_t)keep * rec;
  BGPU_TRY(dev_alloc(ctx, (size_t)new_max * rec, &acc));
  if (keep) BGPU_CUDA(ctx, cudaMemcpyAsync(acc->p, g->acc->p, keep_bytes, cudaMemcpyDeviceToDevice, ctx->stream));
  BGPU_CUDA(ctx, cudaMemsetAsync((char*)acc->p + keep_bytes, 0, (size_t)new_max * rec - keep_bytes, ctx->stream));
  if (g->seen_tracking) {
    BGPU_TRY(dev_alloc(ctx, (size_t)new_max * 4, &seen));
    if (keep) BGPU_CUDA(ctx, cudaMemcpyAsync(seen->p, g->seen->p, (size_t)keep * 4, cudaMemcpyDeviceToDevice, ctx->stream));
    BGPU_CUDA(ctx, cudaMemsetAsync((char*)seen->p + (size_t)keep * 4, 0, (size_t)(new_max - keep) * 4, ctx->stream));
    g->seen = seen;
  }
  DevAllocPtr old_slots = g->slots;
  const uint32_t old_capacity = g->max_groups * 2;
  g->slots = slots; g->acc = acc;
  g->max_groups = new_max;
  if (keep) {
    rehash_kernel<<<agg_launch_grid(ctx, old_capacity, kAggThreads), kAggThreads, 0, ctx->stream>>>(
        (const uint64_t*)old_slots->p, old_capacity, keep, g->table());
    ctx->launches++;
    BGPU_CUDA(ctx, cudaGetLastError());
  }
  return BGPU_OK;
}

__global__ void adjacent_equal_kernel(const uint64_t* __restrict__ sorted, int64_t n, uint32_t* __restrict__ out) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i + 1 < n; i += (int64_t)gridDim.x * blockDim.x)
    if (sorted[i] == sorted[i + 1]) atomicAdd(out, 1u);
}

// Are the values of dictionary `d` pairwise distinct? Decided on the device from their 64-bit hashes (create_hashes,
// radix sort, adjacent compare): distinct hashes prove distinct values; equal hashes (a duplicate or a collision)
// answer "no", which only sends the caller down the exact host path.
static int dict_values_unique(bgpu_agg* g, const std::shared_ptr<Column>& dict, bool* unique) {
  bgpu_ctx* ctx = g->ctx;
  *unique = false;
  const int64_t n = dict->length;
  if (dict->validity) return BGPU_OK;  // null dictionary values: host path
  if (n <= 1) { *unique = true; return BGPU_OK; }
  if (n > 0x7fffffffLL) return BGPU_OK;
  bgpu_batch tmp(ctx);
  tmp.n_rows = n;
  tmp.cols.push_back(*dict);
  DevAllocPtr h0, h1, cnt, scratch;
  BGPU_TRY(dev_alloc(ctx, (size_t)n * 8, &h0));
  BGPU_TRY(dev_alloc(ctx, (size_t)n * 8, &h1));
  BGPU_TRY(dev_alloc(ctx, 16, &cnt));
  const int32_t key0 = 0;
  BGPU_TRY(bgpu_create_hashes(ctx, &tmp, &key0, 1, (uint64_t*)h0->p));
  size_t bytes = 0;
  cub::DeviceRadixSort::SortKeys(nullptr, bytes, (const uint64_t*)h0->p, (uint64_t*)h1->p, (int)n, 0, 64, ctx->stream);
  BGPU_TRY(dev_alloc(ctx, bytes, &scratch));
  BGPU_CUDA(ctx, cub::DeviceRadixSort::SortKeys(scratch->p, bytes, (const uint64_t*)h0->p, (uint64_t*)h1->p, (int)n, 0, 64, ctx->stream));
  BGPU_CUDA(ctx, cudaMemsetAsync(cnt->p, 0, 16, ctx->stream));
  adjacent_equal_kernel<<<agg_launch_grid(ctx, n, kAggThreads), kAggThreads, 0, ctx->stream>>>((const uint64_t*)h1->p, n, (uint32_t*)cnt->p);
  ctx->launches += 3;
  uint32_t* pin = (uint32_t*)ctx->pinned + 32;
  BGPU_CUDA(ctx, cudaMemcpyAsync(pin, cnt->p, 4, cudaMemcpyDeviceToHost, ctx->stream));
  BGPU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  *unique = *pin == 0;
  return BGPU_OK;
}

// A library-owned device copy of a (borrowed) utf8 dictionary column.
static int own_dictionary(bgpu_agg* g, const Column& d, std::shared_ptr<Column>* out) {
  bgpu_ctx* ctx = g->ctx;
  auto c = std::make_shared<Column>(d);
  c->owned.clear();
  const int64_t n = d.length;
  int32_t ends[2] = {0, 0};
  if (n) {
    BGPU_CUDA(ctx, cudaMemcpyAsync(&ends[0], d.offsets, 4, cudaMemcpyDeviceToHost, ctx->stream));
    BGPU_CUDA(ctx, cudaMemcpyAsync(&ends[1], d.offsets + n, 4, cudaMemcpyDeviceToHost, ctx->stream));
    BGPU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  }
  DevAllocPtr o, b;
  BGPU_TRY(dev_alloc(ctx, (size_t)(n + 1) * 4, &o));
  BGPU_TRY(dev_alloc(ctx, (size_t)ends[1], &b));
  if (n) BGPU_CUDA(ctx, cudaMemcpyAsync(o->p, d.offsets, (size_t)(n + 1) * 4, cudaMemcpyDeviceToDevice, ctx->stream));
  else BGPU_CUDA(ctx, cudaMemsetAsync(o->p, 0, 4, ctx->stream));
  if (ends[1]) BGPU_CUDA(ctx, cudaMemcpyAsync(b->p, d.values, (size_t)ends[1], cudaMemcpyDeviceToDevice, ctx->stream));
  c->offsets = (int32_t*)o->p; c->values = b->p; c->data_bytes = ends[1];
  c->owned.push_back(o); c->owned.push_back(b);
  *out = c;
  return BGPU_OK;
}

// Map the batch's dictionary onto the aggregate's canonical dictionary of group column i (group keys are compared by
// VALUE, as the reference does). *remap = device array or nullptr for the identity.
//   fast path: the first dictionary an aggregate sees, when its values are pairwise distinct and non-null, IS the
//              canonical dictionary (identity map, nothing leaves the device) — the common case of one dictionary
//              shared by every batch of a scan (db-benchmark id3 has 1e5..1e7 values);
//   exact path: a different dictionary arrives (or duplicates / nulls): every dictionary is folded, on the device, into an
//              Interner keyed by value (round 1 did this on the host through a std::unordered_map).
static int canonicalize_dictionary(bgpu_agg* g, int32_t i, const Column& c, const int32_t** remap) {
  bgpu_ctx* ctx = g->ctx;
  bgpu_agg::DictState& ds = g->dicts[(size_t)i];
  *remap = nullptr;
  if (ds.last && ds.last == c.dict) {  // same Column object (kept alive by ds.last, so the identity cannot be recycled)
    *remap = ds.last_remap ? (const int32_t*)ds.last_remap->p : nullptr;
    return BGPU_OK;
  }
  const Column& d = *c.dict;
  if (d.kind != K_UTF8 && d.kind != K_BINARY)
    return ctx->fail(BGPU_EUNSUPPORTED, "group column " + c.name + ": only Dictionary<Int32, Utf8|Binary> keys are supported");
  ds.value_format = d.format;
  ds.value_kind = d.kind;
  if (!ds.canon_dev && !ds.interner) {
    bool unique = false;
    BGPU_TRY(dict_values_unique(g, c.dict, &unique));
    if (unique) {
      ds.canon_dev = c.dict;
      if (c.dict->owned.empty()) BGPU_TRY(own_dictionary(g, d, &ds.canon_dev));  // borrowed buffers may die with the batch
      ds.last = c.dict;
      ds.last_remap = nullptr;
      return BGPU_OK;
    }
  }
  // exact path, on the device: every dictionary is folded into the column's Interner (interner.cu); remap[j] = the
  // canonical id of value j (-1 for a null value). A canonical dictionary from the fast path seeds it with ids = indices.
  if (!ds.interner) {
    if (ds.spare) std::swap(ds.interner, ds.spare);  // emptied by bgpu_agg_reset, capacity kept
    else BGPU_TRY(interner_create(ctx, &ds.interner));
    if (ds.canon_dev) {
      std::shared_ptr<Column> first = ds.canon_dev;
      ds.canon_dev = nullptr;
      DevAllocPtr seed_ids;
      BGPU_TRY(dev_alloc(ctx, (size_t)std::max<int64_t>(first->length, 1) * 4, &seed_ids));
      BGPU_TRY(interner_intern(ds.interner.get(), *first, first->length, (int32_t*)seed_ids->p, true));
    }
  }
  const int64_t n = d.length;
  BGPU_TRY(dev_alloc(ctx, (size_t)std::max<int64_t>(n, 1) * 4, &ds.last_remap));
  BGPU_TRY(interner_intern(ds.interner.get(), d, n, (int32_t*)ds.last_remap->p, false));
  ds.last = c.dict;
  *remap = (const int32_t*)ds.last_remap->p;
  return BGPU_OK;
}

// The canonical dictionary of group column i as a device column (emit).
static int upload_canonical_dictionary(bgpu_agg* g, size_t i, std::shared_ptr<Column>* out) {
  bgpu_ctx* ctx = g->ctx;
  const bgpu_agg::DictState& ds = g->dicts[i];
  if (ds.canon_dev) {  // fast path: the (only) dictionary seen is the canonical one and already lives on the device
    *out = ds.canon_dev;
    return BGPU_OK;
  }
  if (ds.interner) {  // every value folded in, in id order
    Column proto;
    proto.format = ds.value_format; proto.kind = ds.value_kind; proto.flags = ARROW_FLAG_NULLABLE;
    auto d = std::make_shared<Column>();
    BGPU_TRY(interner_export(ds.interner.get(), proto, d.get()));
    *out = d;
    return BGPU_OK;
  }
  auto d = std::make_shared<Column>();
  d->name = ""; d->format = ds.value_format; d->kind = ds.value_kind; d->flags = ARROW_FLAG_NULLABLE;
  d->length = (int64_t)ds.values.size(); d->null_count = 0;
  std::vector<int32_t> offs(ds.values.size() + 1, 0);
  std::string bytes;
  for (size_t j = 0; j < ds.values.size(); j++) {
    bytes += ds.values[j];
    offs[j + 1] = (int32_t)bytes.size();
  }
  DevAllocPtr o, b;
  BGPU_TRY(dev_alloc(ctx, offs.size() * 4, &o));
  BGPU_TRY(dev_alloc(ctx, bytes.size(), &b));
  BGPU_CUDA(ctx, cudaMemcpyAsync(o->p, offs.data(), offs.size() * 4, cudaMemcpyHostToDevice, ctx->stream));
  if (!bytes.empty()) BGPU_CUDA(ctx, cudaMemcpyAsync(b->p, bytes.data(), bytes.size(), cudaMemcpyHostToDevice, ctx->stream));
  BGPU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  d->offsets = (int32_t*)o->p; d->values = b->p; d->data_bytes = (int64_t)bytes.size();
  d->owned.push_back(o); d->owned.push_back(b);
  *out = d;
  return BGPU_OK;
}

// ---- binding ----------------------------------------------------------------------------------------------------
// One batch bound to the aggregate: the shape the kernels are specialised for (JitShape), their argument struct
// (JitArgs) and the interpreted accumulator plan the small AOT commit kernel still uses (AccPlan).
struct Bound {
  JitShape sh;
  JitArgs args{};
  AccPlan ap{};
  const JitModule* mod = nullptr;
  std::vector<DevAllocPtr> ids;  // interned group columns: the batch's ids (freed on the stream once the kernels have run)
};
enum JitKernel : int { JK_FUSED = 0, JK_INTERN = 0, JK_ACC_SHARED = 1, JK_ACC_GLOBAL = 2, JK_SKETCH = 3, JK_BUCKET = 4 };  // index into the module of its kind

static int bind_value_atom(bgpu_ctx* ctx, const bgpu_batch* batch, const HAtom& h, Bound* b, bool* nullable) {
  if (!h.is_col) return BGPU_OK;
  if (h.col < 0 || h.col >= (int32_t)batch->cols.size()) return ctx->fail(BGPU_EINVAL, "batch has too few columns");
  const Column& c = batch->cols[(size_t)h.col];
  if (c.kind != K_FIXED || c.width != vk_width(h.vk))
    return ctx->fail(BGPU_EINVAL, "aggregate input column " + c.name + " has an unexpected type");
  const int32_t s = jit_slot(&b->sh, h.col);
  if (s < 0) return ctx->fail(BGPU_EUNSUPPORTED, "aggregate reads more than 24 distinct columns");
  JitSlot& sl = b->sh.slots[(size_t)s];
  if (sl.key_use && sl.kind != c.kind) return ctx->fail(BGPU_EUNSUPPORTED, "a wide (interned) group column that is also an aggregate argument");
  sl.kind = c.kind; sl.width = c.width; sl.vk = h.vk; sl.value_use = true;
  sl.nullable = c.validity != nullptr;
  b->args.col[s] = c.values;
  b->args.valid[s] = c.validity;
  if (c.validity) *nullable = true;
  return BGPU_OK;
}

static int bind_batch(bgpu_agg* g, const bgpu_batch* batch, Bound* b) {
  bgpu_ctx* ctx = g->ctx;
  const int32_t n_group = (int32_t)g->group_cols.size();
  for (int32_t i = 0; i < n_group; i++) {
    const int32_t bc = g->final_mode ? i : g->group_cols[(size_t)i];
    if (bc >= (int32_t)batch->cols.size()) return ctx->fail(BGPU_EINVAL, "batch has too few columns");
    const Column& c = batch->cols[(size_t)bc];
    const Column& f = g->fields[(size_t)g->group_cols[(size_t)i]];
    if (!same_shape(c, f)) return ctx->fail(BGPU_EINVAL, "group column " + f.name + " does not match input_schema");
    if (g->key_interned[(size_t)i]) {
      // a wide group column: its values become dense ids (interner.cu) and enter the packed key like dictionary
      // indices: 4 bytes, -1 = null
      if (!g->interners[(size_t)i]) BGPU_TRY(interner_create(ctx, &g->interners[(size_t)i]));
      DevAllocPtr ids;
      BGPU_TRY(dev_alloc(ctx, (size_t)std::max<int64_t>(batch->n_rows, 1) * 4, &ids));
      BGPU_TRY(interner_intern(g->interners[(size_t)i].get(), c, batch->n_rows, (int32_t*)ids->p, false));
      b->ids.push_back(ids);
      // the ids are a column of their own (slot keyed beyond any batch column): the same batch column may also be an
      // aggregate argument or a filter operand, read by value through its own slot
      const int32_t s = jit_slot(&b->sh, 100000 + i);
      if (s < 0) return ctx->fail(BGPU_EUNSUPPORTED, "aggregate reads more than 24 distinct columns");
      JitSlot& sl = b->sh.slots[(size_t)s];
      sl.kind = K_DICT32; sl.width = 4; sl.key_use = true; sl.nullable = false; sl.remap = false;
      b->args.col[s] = ids->p;
      b->args.valid[s] = nullptr;
      b->args.aux[s] = nullptr;
      b->sh.key_slot.push_back(s);
      continue;
    }
    const int32_t* remap = nullptr;
    if (c.kind == K_DICT32) BGPU_TRY(canonicalize_dictionary(g, i, c, &remap));
    const int32_t s = jit_slot(&b->sh, bc);
    if (s < 0) return ctx->fail(BGPU_EUNSUPPORTED, "aggregate reads more than 24 distinct columns");
    JitSlot& sl = b->sh.slots[(size_t)s];
    sl.kind = c.kind; sl.width = c.width; sl.key_use = true;
    sl.nullable = c.validity != nullptr;
    sl.remap = remap != nullptr;
    b->args.col[s] = c.values;
    b->args.valid[s] = c.validity;
    b->args.aux[s] = c.kind == K_DICT32 ? remap : c.offsets;
    b->sh.key_slot.push_back(s);
  }
  for (size_t i = 0; i < g->filter.size(); i++) {
    const HFilterTerm& h = g->filter[i];
    if (h.col >= (int32_t)batch->cols.size()) return ctx->fail(BGPU_EINVAL, "batch has too few columns");
    const Column& c = batch->cols[(size_t)h.col];
    if (c.kind != K_FIXED || c.width != vk_width(h.vk))
      return ctx->fail(BGPU_EINVAL, "filter column " + c.name + " has an unexpected type");
    const int32_t s = jit_slot(&b->sh, h.col);
    if (s < 0) return ctx->fail(BGPU_EUNSUPPORTED, "aggregate reads more than 24 distinct columns");
    JitSlot& sl = b->sh.slots[(size_t)s];
    if (sl.key_use && sl.kind != c.kind) return ctx->fail(BGPU_EUNSUPPORTED, "a wide (interned) group column that is also a filter column");
    sl.kind = c.kind; sl.width = c.width; sl.vk = h.vk; sl.filter_use = true;
    sl.nullable = c.validity != nullptr;
    b->args.col[s] = c.values;
    b->args.valid[s] = c.validity;
    b->sh.filter_slot.push_back(s);
  }
  b->sh.value_nullable.assign(g->values.size(), false);
  for (size_t vi = 0; vi < g->values.size(); vi++) {
    const HValue& h = g->values[vi];
    bool nullable = g->value_from_prev[vi] ? (bool)b->sh.value_nullable[vi - 1] : false;
    for (const HFactor& hf : h.f) {
      BGPU_TRY(bind_value_atom(ctx, batch, hf.x, b, &nullable));
      if (hf.op) BGPU_TRY(bind_value_atom(ctx, batch, hf.y, b, &nullable));
    }
    b->sh.value_nullable[vi] = nullable;
    b->sh.any_validity |= nullable;
  }
  // accumulator words. Counts over null-free values all equal the row count of the group, so only the first is
  // kept live and the others alias it (the flush adds the alias' partial; the global-atomic tier adds all of them).
  b->ap.n = (int32_t)g->phys.size();
  int32_t first_count = -1, w = 0;
  for (int32_t i = 0; i < b->ap.n; i++) {
    const PhysAcc& p = g->phys[(size_t)i];
    AccK& a = b->ap.a[i];
    a = AccK{p.out_off, p.op, p.words, 0, i, -1};
    if (p.op == OP_COUNT && !b->sh.value_nullable[(size_t)p.value]) {
      if (first_count < 0) first_count = i;
      else a.alias = first_count;
    }
    if (a.alias < 0) {
      a.word0 = w;
      w += a.words;
    }
    b->sh.acc_alias.push_back(a.alias);
    b->sh.acc_word0.push_back(a.word0);
  }
  b->ap.total_words = w;
  b->sh.W = w;
  return BGPU_OK;
}

// accumulator / seen arrays and the table move when the table grows
static void refresh_pointers(bgpu_agg* g, Bound* b) {
  b->ap.seen = g->seen_tracking ? (uint32_t*)g->seen->p : nullptr;
  b->ap.acc = (uint64_t*)g->acc->p;
  b->ap.stride = g->acc_stride;
  b->args.seen = b->ap.seen;
  b->args.t = g->table();
}

// Everything of a JitShape the generated source depends on, as a compact key.
static void shape_key(const JitShape& sh, std::string* k) {
  auto put = [&](int32_t v) { k->append((const char*)&v, 4); };
  put((int32_t)sh.slots.size());
  for (const JitSlot& sl : sh.slots) {
    put(sl.batch_col); put(sl.kind); put(sl.width); put(sl.vk);
    put((sl.nullable ? 1 : 0) | (sl.remap ? 2 : 0) | (sl.key_use ? 4 : 0) | (sl.value_use ? 8 : 0) | (sl.filter_use ? 16 : 0));
  }
  put((int32_t)sh.key_slot.size());
  for (int32_t v : sh.key_slot) put(v);
  put((int32_t)sh.filter_slot.size());
  for (int32_t v : sh.filter_slot) put(v);
  put((int32_t)sh.value_nullable.size());
  for (bool v : sh.value_nullable) put(v ? 1 : 0);
  put((int32_t)sh.acc_alias.size());
  for (int32_t v : sh.acc_alias) put(v);
  for (int32_t v : sh.acc_word0) put(v);
  put(sh.W); put(sh.has_seen ? 1 : 0); put(sh.T); put(sh.R); put(sh.min_ctas); put(sh.kernels);
  put((sh.any_validity ? 1 : 0) | (sh.row32 ? 2 : 0));
}

// Generate + compile (cached) the fused kernel (kernels = 1 private, 4 shared-atomic) or the two-pass kernels (2) for this batch
// shape at the given launch geometry. b->mod->fn is indexed by JitKernel. The process-wide cache is keyed by the plan (the
// arguments bgpu_agg_create_expr was given, bgpu_agg::plan_key) and the shape, so that a stage that runs the same plan
// again - a new aggregate per task - neither generates the source nor looks a multi-kilobyte string up.
static int specialise(bgpu_agg* g, Bound* b, int kernels, int T, int R, int min_ctas) {
  b->sh.T = T; b->sh.R = R; b->sh.min_ctas = min_ctas; b->sh.kernels = kernels;
  b->sh.has_seen = g->seen_tracking;
  static std::mutex mu;
  static std::unordered_map<std::string, const JitModule*> cache;
  std::string key;
  key.reserve(g->plan_key.size() + 256);
  key.append((const char*)&g->ctx->device, sizeof(int));
  key += g->plan_key;
  key.push_back(g->interned_mode ? 'I' : 'i');  // the key layout the source is generated for
  key.append((const char*)g->key_interned.data(), g->key_interned.size());
  shape_key(b->sh, &key);
  {
    std::lock_guard<std::mutex> lk(mu);
    auto it = cache.find(key);
    if (it != cache.end()) { b->mod = it->second; return BGPU_OK; }
  }
  std::string src;
  BGPU_TRY(agg_generate_source(g, b->sh, &src));
  static const std::vector<std::string> fused{"bgpu_fused"}, fused_shared{"bgpu_fused_shared"},
      two_pass{"bgpu_intern", "bgpu_acc_shared", "bgpu_acc_global", "bgpu_sketch", "bgpu_bucket"};
  BGPU_TRY(jit_compile(g->ctx, agg_jit_prelude(), src, kernels == 1 ? fused : kernels == 4 ? fused_shared : two_pass, &b->mod));
  std::lock_guard<std::mutex> lk(mu);
  cache.emplace(std::move(key), b->mod);
  return BGPU_OK;
}

// attributes of a compiled kernel that never change: asked once
static int cached_local_bytes(bgpu_ctx* ctx, CUfunction fn, int* out) {
  static std::mutex mu;
  static std::map<CUfunction, int> cache;
  {
    std::lock_guard<std::mutex> lk(mu);
    auto it = cache.find(fn);
    if (it != cache.end()) { *out = it->second; return BGPU_OK; }
  }
  BGPU_TRY(jit_local_bytes(ctx, fn, out));
  std::lock_guard<std::mutex> lk(mu);
  cache[fn] = *out;
  return BGPU_OK;
}
static int cached_occupancy(bgpu_ctx* ctx, CUfunction fn, int block, size_t smem, int* out) {
  static std::mutex mu;
  static std::map<std::tuple<CUfunction, int, size_t>, int> cache;
  const auto key = std::make_tuple(fn, block, smem);
  {
    std::lock_guard<std::mutex> lk(mu);
    auto it = cache.find(key);
    if (it != cache.end()) { *out = it->second; return BGPU_OK; }
  }
  BGPU_TRY(jit_occupancy(ctx, fn, block, smem, out));
  std::lock_guard<std::mutex> lk(mu);
  cache[key] = *out;
  return BGPU_OK;
}

static int read_counters(bgpu_agg* g, uint32_t* h) {
  bgpu_ctx* ctx = g->ctx;
  uint32_t* pin = (uint32_t*)ctx->pinned;
  BGPU_CUDA(ctx, cudaMemcpyAsync(pin, g->counters->p, 32, cudaMemcpyDeviceToHost, ctx->stream));
  BGPU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  memcpy(h, pin, 32);
  return BGPU_OK;
}
static int write_counters(bgpu_agg* g, const uint32_t* h) {
  bgpu_ctx* ctx = g->ctx;
  uint32_t* pin = (uint32_t*)ctx->pinned + 16;
  memcpy(pin, h, 32);
  BGPU_CUDA(ctx, cudaMemcpyAsync(g->counters->p, pin, 32, cudaMemcpyHostToDevice, ctx->stream));
  BGPU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return BGPU_OK;
}

// counters layout: [0] n_groups, [1] table overflow, [2] string key too long, [3] arithmetic overflow, [4] fused abort
static int check_common_flags(bgpu_agg* g, const uint32_t* h) {
  bgpu_ctx* ctx = g->ctx;
  if (h[2]) g->key_too_long = true;
  if (h[2]) return ctx->fail(BGPU_EUNSUPPORTED, "string group key longer than the packed-key budget (" +
                                                    std::to_string(16 / std::max<size_t>(1, g->group_cols.size()) - 1) + " bytes)");
  if (h[3]) return ctx->fail(BGPU_EOVERFLOW, "Arithmetic overflow in an aggregate argument (Decimal128)");
  return BGPU_OK;
}

// Table overflowed during a pass: groups [0, max_groups) are intact. Interning stops as soon as the overflow flag is
// up, so `counter` barely exceeds the old capacity and says little about the true cardinality; every retry is a full
// pass over the batch, so grow geometrically fast (x16, bounded by what this batch can still add) rather than x2.
static int grow_after_overflow(bgpu_agg* g, uint32_t counter, int64_t batch_rows) {
  const uint32_t keep = g->max_groups;
  if (counter > 0x70000000u) return g->ctx->fail(BGPU_EUNSUPPORTED, "too many groups");
  const uint64_t bound = (uint64_t)keep + (uint64_t)batch_rows;  // no batch can create more groups than it has rows
  const uint64_t want = std::max<uint64_t>((uint64_t)counter + counter / 4, std::min<uint64_t>((uint64_t)keep * 16, bound));
  const uint32_t new_max = next_pow2(std::max<uint64_t>(want, (uint64_t)keep * 2));
  const uint32_t reset[8] = {keep, 0, 0, 0, 0, 0, 0, 0};
  BGPU_TRY(write_counters(g, reset));
  return agg_resize(g, new_max, keep);
}

// The single-pass ("fused") ladder. mode 0 = thread-private accumulators, mode 1 = replicated shared-memory
// accumulators with 32-bit shared atomics.
// (alt_*: the geometry to fall back to when the kernel spills registers at the first one)
struct Rung { int mode; int threads; int rows; int min_ctas; const char* name; int alt_threads; int alt_rows; int alt_ctas; };

// Try the fused ladder. *done = true when the batch has been committed.
static int try_fused(bgpu_agg* g, Bound* b, int64_t n_rows, bool* done, AggAsyncReq* req) {
  bgpu_ctx* ctx = g->ctx;
  *done = false;
  const int W = b->sh.W;
  if (W == 0) { g->fused_rung = kFusedRungs; return BGPU_OK; }
  const size_t smem_max = ctx->smem_optin - 1024;
  // rung 0: thread-private accumulators cost 8 bytes per live accumulator (+4 per i128 sum) per group and THREAD, so
  //   shared memory and the register file together bound the threads per SM: 640 threads x 4 groups of Q1 fill one SM
  //   at <= 96 registers. One row per thread and tile through the three-stage load pipeline measured best on B200
  //   (Q1, 300M rows: 640x1 pipelined 4.76 ms, 640x2 4.96 ms, 512x2 6.6 ms, 768x2 (spills) 6.2 ms).
  // rungs 1, 2: `replicas` copies of [G][W] per CTA (one per warp while they fit), native 32-bit shared atomics; the
  //   accumulators no longer scale with the thread count, so hundreds to thousands of groups run at full residency.
  // rung 3: as rung 2 without the per-CTA key cache (group ids straight from the global table, which is L2 resident at
  //   this cardinality): all of shared memory holds accumulators, ~10^4 groups.
  // Shared-atomic rungs: one row per thread through the load pipeline at the highest residency the registers allow
  //   (db-benchmark q2, 5e8 rows: 1024x1 7.9 ms vs 512x2 11.2 ms).
  Rung rungs[kFusedRungs] = {{0, 640, 1, 1, "fused_private", 512, 1, 1}, {1, 512, 1, 2, "fused_shared2", 256, 4, 3},
                             {1, 1024, 1, 1, "fused_shared1", 512, 2, 1}, {2, 1024, 1, 1, "fused_shared0", 512, 2, 1}};
  // tuning knobs: "threads,rows,ctas_per_sm" for the first rung (BGPU_FUSED_GEOMETRY) or rung i (BGPU_FUSED_RUNG<i>)
  for (int i = 0; i < kFusedRungs; i++) {
    const std::string name = "BGPU_FUSED_RUNG" + std::to_string(i);
    const char* e = getenv(name.c_str());
    if (!e && i == 0) e = getenv("BGPU_FUSED_GEOMETRY");
    int t = 0, r = 0, c = 0;
    if (e && sscanf(e, "%d,%d,%d", &t, &r, &c) == 3 && t >= 32 && t <= 1024 && t % 32 == 0 && r >= 1 && r <= 16 && c >= 1 && c <= 8) {
      rungs[i].threads = t; rungs[i].rows = r; rungs[i].min_ctas = c; rungs[i].alt_threads = 0;
    }
  }
  while (g->fused_rung < kFusedRungs) {
    Rung rg = rungs[g->fused_rung];
    int waves = 8;
    if (const char* e = getenv("BGPU_FUSED_WAVES")) waves = std::max(1, atoi(e));
    b->sh.row32 = !getenv("BGPU_ROW64") && n_rows < (1LL << 31) && (int64_t)ctx->sm_count * 32 * waves * 1024 * 16 < (1LL << 31);
    if (rg.alt_threads) {  // does the kernel fit the register budget of the preferred geometry?
      BGPU_TRY(specialise(g, b, rg.mode == 0 ? 1 : 4, rg.threads, rg.rows, rg.min_ctas));
      int local = 0;
      BGPU_TRY(cached_local_bytes(ctx, b->mod->fn[JK_FUSED], &local));
      if (local > 0) { rg.threads = rg.alt_threads; rg.rows = rg.alt_rows; rg.min_ctas = rg.alt_ctas; }
    }
    const int T = rg.threads;
    const size_t cap = smem_max / (size_t)rg.min_ctas - 1024;
    // capacity G: largest with accumulators + seen + key cache (2G slots rounded up, 24 B each) <= cap
    uint32_t G, replicas = 1;
    // thread-private layout (PrivateSink): one u64 per live accumulator + one i32 per i128 sum, per thread; one u64
    // spill word per i128 sum, per CTA
    int n128 = 0;
    for (int32_t i = 0; i < b->ap.n; i++) n128 += b->ap.a[i].alias < 0 && b->ap.a[i].words == 2;
    const size_t priv_bytes = ((size_t)(W - n128) + (size_t)(n128 + 1) / 2) * 8 * T + (size_t)n128 * 8;
    bool g_can_grow = false;
    if (rg.mode == 0) {
      // Shared memory not taken by accumulators stays L1 cache, which the overlapping utf8 window and offset loads
      // live on (Q1: capacity 5 instead of 4 groups costs 30%): start small and grow on demand.
      const uint32_t g_max = (uint32_t)(cap / (priv_bytes + 4 + 4 * 24));
      G = std::min<uint32_t>(g_max, std::max<uint32_t>(g->private_gcap, g->n_groups));
      if (const char* e = getenv("BGPU_FUSED_GCAP")) G = std::min<uint32_t>(g_max, std::max(1, atoi(e)));
      g_can_grow = G < g_max;
    } else if (rg.mode == 1) {
      G = (uint32_t)(cap / ((size_t)W * 8 + 4 + 4 * 24));
      if (rg.min_ctas > 1) {
        // the multi-CTA rung: capacity by demand. Shared memory the groups do not need becomes copies of the
        // accumulators, and with one copy per warp the kernel needs no atomics at all (PlainSharedSink).
        const uint32_t g_max = std::min<uint32_t>(G, 1024);
        G = std::min<uint32_t>(g_max, std::max<uint32_t>(g->shared_gcap, g->n_groups));
        if (const char* e = getenv("BGPU_SHARED_GCAP")) G = std::min<uint32_t>(g_max, std::max(1, atoi(e)));
        g_can_grow = G < g_max;
      }
    } else {
      G = (uint32_t)((cap - 64) / ((size_t)W * 8 + 4));
    }
    if (G > 32768) G = 32768;
    if (G < 1 || G < g->n_groups) { g->fused_rung++; continue; }
    const uint32_t slots = rg.mode == 2 ? 0u : std::max<uint32_t>(16, next_pow2((uint64_t)G * 2));
    // seen masks + key cache (+ below: one claim byte per group and warp when every warp gets its own copy)
    size_t fixed = (size_t)((G + 1) & ~1u) * 4 + (size_t)slots * 24;
    if (rg.mode != 0) {
      if (fixed + (size_t)G * W * 8 > cap) { g->fused_rung++; continue; }
      replicas = (uint32_t)std::min<size_t>((size_t)T / 32, (cap - fixed) / ((size_t)G * W * 8));
      if (replicas == (uint32_t)(T / 32)) {  // one copy per warp = the atomic-free path, if its claim bytes fit as well
        const size_t claims = (size_t)(T / 32) * ((G + 15) & ~15u);
        // Measured (5e8 rows, k = 100): four accumulator words per row (q4: avg x 3, one of them Float64): 5.6 ms
        // atomic-free against 9.5 ms with shared atomics; ONE word per row (q1): 3.1 ms against 2.3 ms - a single
        // fire-and-forget ATOMS costs less than the claim round trip. So: atomic-free from two words per row up.
        const bool want = W >= 2 && !getenv("BGPU_NO_WARP_PRIVATE");
        if (want && fixed + claims + (size_t)G * W * 8 * replicas <= cap) fixed += claims;
        else replicas--;  // fewer copies than warps: shared atomics
      }
      if (replicas < 1) { g->fused_rung++; continue; }
    }
    const size_t smem = (rg.mode == 0 ? (size_t)G * priv_bytes : (size_t)G * W * 8 * replicas) + fixed;
    if (smem > smem_max) { g->fused_rung++; continue; }
    BGPU_TRY(specialise(g, b, rg.mode == 0 ? 1 : 4, T, rg.rows, rg.min_ctas));
    const CUfunction fn = b->mod->fn[JK_FUSED];
    b->args.replicas = replicas;
    BGPU_TRY(jit_set_max_smem(ctx, fn, smem));
    int occ = 1;
    BGPU_TRY(cached_occupancy(ctx, fn, T, smem, &occ));
    occ = std::max(1, occ);
    // several CTAs per SM slot: the hardware block scheduler then evens out the per-SM speed differences that a
    // one-wave persistent grid leaves as a 20% idle tail (ncu: SM active 77% of elapsed), at ~10 us of prologue +
    // flush per extra CTA
    // (the private i32 high words need at most 65535 rows per thread)
    const int grid = (int)std::max<int64_t>(std::max<int64_t>(1, ceil_div64(n_rows, (int64_t)T * 65535)),
                                            std::min<int64_t>(ceil_div64(n_rows, (int64_t)T * rg.rows), (int64_t)ctx->sm_count * occ * waves));
    const size_t scratch_bytes = (size_t)grid * fused_slab_words(G, W) * 8;
    if (!g->scratch || g->scratch->bytes < scratch_bytes) BGPU_TRY(dev_alloc(ctx, scratch_bytes, &g->scratch));
    b->args.scratch = (uint64_t*)g->scratch->p;
    b->args.g_cap = G;
    b->args.cache_slots = slots;
    bool retry_rung = false;
    for (int attempt = 0;; attempt++) {
      refresh_pointers(g, b);
      if (req && g->async_ready) {
        // streamed input: launch, let the device decide whether to commit, look at the outcome later
        if (b->sh.any_validity && !g->seen_tracking) return ctx->fail(BGPU_EINVAL, "internal: seen tracking not set up");
        BGPU_CUDA(ctx, cudaEventRecord(req->k0, ctx->stream));
        BGPU_TRY(jit_launch(ctx, fn, grid, T, smem, &b->args));
        BGPU_CUDA(ctx, cudaEventRecord(req->k1, ctx->stream));
        commit_cond_kernel<<<agg_launch_grid(ctx, (int64_t)std::max<uint32_t>(1, G * (uint32_t)b->ap.n) * 32, kAggThreads), kAggThreads, 0,
                             ctx->stream>>>(b->ap, (const uint64_t*)g->scratch->p, (uint32_t)grid, G, (const uint32_t*)g->counters->p,
                                            req->d_status);
        ctx->launches++;
        BGPU_CUDA(ctx, cudaGetLastError());
        BGPU_CUDA(ctx, cudaMemcpyAsync((void*)req->h_status, req->d_status, 32, cudaMemcpyDeviceToHost, ctx->stream));
        BGPU_CUDA(ctx, cudaEventRecord(req->done, ctx->stream));
        req->launched = true;
        g->last_path = rg.name;
        *done = true;
        return BGPU_OK;
      }
      BGPU_CUDA(ctx, cudaEventRecord(g->ev[0], ctx->stream));
      BGPU_TRY(jit_launch(ctx, fn, grid, T, smem, &b->args));
      BGPU_CUDA(ctx, cudaEventRecord(g->ev[1], ctx->stream));
      uint32_t h[8];
      BGPU_TRY(read_counters(g, h));  // synchronises the stream
      float ms = 0.f;
      if (cudaEventElapsedTime(&ms, g->ev[0], g->ev[1]) == cudaSuccess) g->hot_kernel_ms += ms;
      BGPU_TRY(check_common_flags(g, h));
      if (h[1]) {  // global table overflow: grow and run the same rung again
        if (attempt >= 8) return ctx->fail(BGPU_EUNSUPPORTED, "too many groups");
        BGPU_TRY(grow_after_overflow(g, h[0], n_rows));
        continue;
      }
      if (h[4]) {  // some CTA met a group id >= G: nothing was committed; the groups interned so far stay
        h[4] = 0;
        BGPU_TRY(write_counters(g, h));
        g->n_groups = h[0];
        if (g_can_grow) {  // same rung, more groups per copy
          if (rg.mode == 0) g->private_gcap = std::max<uint32_t>(2 * G, g->n_groups);
          else g->shared_gcap = std::max<uint32_t>(2 * G, g->n_groups);
          retry_rung = true;
        }
        break;
      }
      g->n_groups = h[0];
      if (b->sh.any_validity && !g->seen_tracking) return ctx->fail(BGPU_EINVAL, "internal: seen tracking not set up");
      commit_kernel<<<agg_launch_grid(ctx, (int64_t)std::max<uint32_t>(1, g->n_groups * (uint32_t)b->ap.n) * 32, kAggThreads), kAggThreads, 0,
                      ctx->stream>>>(b->ap, (const uint64_t*)g->scratch->p, (uint32_t)grid, G, g->n_groups);
      ctx->launches++;
      BGPU_CUDA(ctx, cudaGetLastError());
      g->last_path = rg.name;
      g->any_committed = true;
      g->async_ready = true;  // this rung has just committed a batch: the next one may run without a host round trip
      *done = true;
      return BGPU_OK;
    }
    if (!retry_rung) g->fused_rung++;
  }
  return BGPU_OK;
}

// BGPU_AGG_TRACE=1: per-phase wall times of the two-pass path and of emit on stderr (each phase synchronises)
struct PhaseTrace {
  bgpu_ctx* ctx;
  bool on;
  std::chrono::steady_clock::time_point t0;
  explicit PhaseTrace(bgpu_ctx* c) : ctx(c), on(getenv("BGPU_AGG_TRACE") != nullptr) {
    if (on) { cudaStreamSynchronize(ctx->stream); t0 = std::chrono::steady_clock::now(); }
  }
  void mark(const char* what) {
    if (!on) return;
    cudaStreamSynchronize(ctx->stream);
    const auto t1 = std::chrono::steady_clock::now();
    fprintf(stderr, "[bgpu agg] %-14s %8.3f ms\n", what, std::chrono::duration<double, std::milli>(t1 - t0).count());
    t0 = t1;
  }
};

// ---- pass-through (skip a Partial aggregation that does not reduce) ---------------------------------------------------------
// Eligible: Partial mode, no filter, non-null fixed-width group columns, every aggregate argument a plain non-null column
// whose bits ARE its state (Decimal128 -> Decimal128 sum, Int64 -> Int64 sum, Float64 -> Float64 sum) or COUNT.
static bool passthrough_eligible(const bgpu_agg* g, const bgpu_batch* batch) {
  if (g->mode != BGPU_AGG_MODE_PARTIAL || !g->filter.empty() || getenv("BGPU_NO_PASSTHROUGH")) return false;
  for (int32_t c : g->group_cols) {
    const Column& col = batch->cols[(size_t)c];
    if (col.kind != K_FIXED || col.validity) return false;
  }
  for (const PhysAcc& p : g->phys) {
    const HValue& v = g->values[(size_t)p.value];
    if (p.op == OP_COUNT) {
      if (v.f.empty()) continue;  // COUNT(*)
      if (!v.plain_column() || batch->cols[(size_t)v.f[0].x.col].validity) return false;
      continue;
    }
    if (!v.plain_column()) return false;
    const Column& col = batch->cols[(size_t)v.f[0].x.col];
    if (col.validity || col.kind != K_FIXED) return false;
    const int32_t vk = v.f[0].x.vk;
    const bool same_bits = (p.op == OP_ADD_I128 && vk == V_I128) || (p.op == OP_ADD_I64 && (vk == V_I64 || vk == V_U64)) ||
                           (p.op == OP_ADD_F64 && vk == V_F64);
    if (!same_bits) return false;
  }
  return true;
}

// a column of the chunk that holds the values of input column `src`: shared when the batch owns its buffers, copied
// when they are borrowed (they are only promised for the call)
static int pass_column(bgpu_ctx* ctx, const Column& src, int64_t n, Column* out) {
  out->kind = K_FIXED; out->width = src.width; out->length = n; out->null_count = 0; out->flags = ARROW_FLAG_NULLABLE;
  if (!src.owned.empty()) {
    out->values = src.values;
    out->owned = src.owned;
    return BGPU_OK;
  }
  DevAllocPtr a;
  BGPU_TRY(dev_alloc(ctx, (size_t)n * src.width, &a));
  if (n) BGPU_CUDA(ctx, cudaMemcpyAsync(a->p, src.values, (size_t)n * src.width, cudaMemcpyDeviceToDevice, ctx->stream));
  out->values = a->p;
  out->owned.push_back(a);
  return BGPU_OK;
}

static int passthrough_update(bgpu_agg* g, const bgpu_batch* batch) {
  bgpu_ctx* ctx = g->ctx;
  const int64_t n = batch->n_rows;
  std::unique_ptr<bgpu_batch> out(new bgpu_batch(ctx));
  out->n_rows = n;
  for (int32_t c : g->group_cols) {
    const Column& f = g->fields[(size_t)c];
    Column col;
    BGPU_TRY(pass_column(ctx, batch->cols[(size_t)c], n, &col));
    col.name = f.name; col.format = f.format; col.flags = f.flags | ARROW_FLAG_NULLABLE;
    out->cols.push_back(std::move(col));
  }
  DevAllocPtr ones;  // one buffer of 1s behind every count column
  auto count_column = [&](const std::string& name, const char* format) -> int {
    if (!ones) {
      BGPU_TRY(dev_alloc(ctx, (size_t)std::max<int64_t>(n, 1) * 8, &ones));
      if (n) {
        fill_u64_kernel<<<agg_launch_grid(ctx, n, kAggThreads), kAggThreads, 0, ctx->stream>>>((uint64_t*)ones->p, n, 1ULL);
        ctx->launches++;
        BGPU_CUDA(ctx, cudaGetLastError());
      }
    }
    Column col;
    col.name = name; col.format = format; col.flags = ARROW_FLAG_NULLABLE; col.kind = K_FIXED; col.width = 8;
    col.length = n; col.null_count = 0; col.values = ones->p;
    col.owned.push_back(ones);
    out->cols.push_back(std::move(col));
    return BGPU_OK;
  };
  auto sum_column = [&](int phys_idx, const std::string& name, const std::string& format) -> int {
    const HValue& v = g->values[(size_t)g->phys[(size_t)phys_idx].value];
    Column col;
    BGPU_TRY(pass_column(ctx, batch->cols[(size_t)v.f[0].x.col], n, &col));
    col.name = name; col.format = format;
    out->cols.push_back(std::move(col));
    return BGPU_OK;
  };
  for (const LogicalAgg& a : g->aggs) {
    if (a.fn == BGPU_AGG_COUNT) {
      BGPU_TRY(count_column(a.name + "[count]", "l"));
    } else if (a.fn == BGPU_AGG_SUM) {
      BGPU_TRY(sum_column(a.phys[0], a.name + "[sum]", a.sum_format));
    } else {
      BGPU_TRY(count_column(a.name + "[count]", "L"));
      BGPU_TRY(sum_column(a.phys[1], a.name + "[sum]", a.sum_format));
    }
  }
  g->pass_chunks.push_back(std::move(out));
  g->last_path = "passthrough";
  return BGPU_OK;
}

// Radix pre-pass of the two-pass path (VERDICT r1 #4b). With 10^7..10^8 groups the table and the accumulator records are
// gigabytes: every row is a random sector of the table plus a random record, and both passes run at DRAM-latency speed.
// When the groups of a batch reach far beyond the L2, the rows the kernels read (the bound columns only) are first moved
// into BUCKET order - bucket = top bits of the row's first slot index, so one bucket touches one contiguous slice of the
// table and, ids being handed out in arrival order, one contiguous run of records - by bgpu_bucket (ids + histogram) and
// partition.cu's scan + scatter; pass 1 and 2 then walk the reordered rows tile by tile across the whole grid, i.e. all
// SMs work inside the same few L2-resident slices at any time. Sums do not depend on row order; results are unchanged.
static int radix_reorder(bgpu_agg* g, Bound* b, int64_t n_rows, double est_groups, int T, std::unique_ptr<bgpu_batch>* keep) {
  bgpu_ctx* ctx = g->ctx;
  // Measured (profiles/r2_orderkey.md): the accumulate pass gains (6.8 -> 3.8 ms per 1e8 rows x 2.5e7 groups) but the intern
  // pass does not (it is issue-bound at 126 registers, not DRAM-latency-bound: 6.8 -> 7.2 ms) and the reorder costs 3.4 ms
  // (n = 64 buckets) to 28 ms (6e8 rows, 512 buckets): no net gain, so the pre-pass is OFF unless BGPU_RADIX=1 asks for it.
  const char* on = getenv("BGPU_RADIX");
  if (!on || atoi(on) == 0) return BGPU_OK;
  double slice_mb = 32.0, min_mb = 160.0;  // bucket working set to aim for; below min_mb the L2 holds enough of it anyway
  if (const char* e = getenv("BGPU_RADIX_SLICE_MB")) slice_mb = std::max(1.0, atof(e));
  if (const char* e = getenv("BGPU_RADIX_MIN_MB")) min_mb = atof(e);
  const double touched = est_groups * (32.0 + (double)g->acc_stride * 8.0);  // one sector of the table + the record, per group
  int64_t min_rows = 4 << 20;
  if (const char* e = getenv("BGPU_RADIX_MIN_ROWS")) min_rows = atoll(e);
  if (n_rows < min_rows || touched < min_mb * 1048576.0) return BGPU_OK;
  const uint32_t cap = g->max_groups * 2;
  uint32_t nb = next_pow2((uint64_t)(touched / (slice_mb * 1048576.0)) + 1);
  nb = std::min<uint32_t>(std::max<uint32_t>(nb, 16), 512);
  if (const char* e = getenv("BGPU_RADIX_BUCKETS")) nb = next_pow2((uint64_t)std::max(2, atoi(e)));
  if (nb > 2048 || cap < nb * 64) return BGPU_OK;
  const size_t n_slots = b->sh.slots.size();
  bgpu_batch tmp(ctx);
  tmp.n_rows = n_rows;
  tmp.cols.resize(n_slots);
  for (size_t s = 0; s < n_slots; s++) {
    const JitSlot& sl = b->sh.slots[s];
    Column& c = tmp.cols[s];
    if (sl.kind == K_FIXED && (sl.width == 1 || sl.width == 2 || sl.width == 4 || sl.width == 8 || sl.width == 16)) {
      c.kind = K_FIXED; c.width = sl.width;
    } else if (sl.kind == K_DICT32) {
      c.kind = K_FIXED; c.width = 4;  // the indices move; the dictionary (and its remap) belongs to no row
    } else if (sl.kind == K_BOOL) {
      c.kind = K_BOOL; c.width = 0;
    } else {
      return BGPU_OK;  // utf8 / binary columns stay on the plain path
    }
    c.length = n_rows;
    c.values = const_cast<void*>(b->args.col[s]);
    c.validity = sl.nullable ? const_cast<uint8_t*>(b->args.valid[s]) : nullptr;
    if (!c.values) return BGPU_OK;
  }
  int shift = 0;
  while ((cap >> shift) > nb) shift++;
  const CUfunction fn = b->mod->fn[JK_BUCKET];
  bgpu_batch* out = nullptr;
  BGPU_TRY(partition_rows(ctx, &tmp, nb,
                          [&](int64_t grid, int64_t chunk_rows, uint32_t* hist, uint16_t* pids) -> int {
                            const uint32_t share = (uint32_t)std::max<int64_t>(1, ((int64_t)ctx->sm_count * 8) / grid);
                            BGPU_CUDA(ctx, cudaMemsetAsync(hist, 0, (size_t)nb * (size_t)grid * 4, ctx->stream));
                            refresh_pointers(g, b);
                            b->args.bucket_ids = pids;
                            b->args.bucket_hist = hist;
                            b->args.bucket_shift = (uint32_t)shift;
                            b->args.bucket_chunk = (uint32_t)chunk_rows;
                            b->args.n_groups = nb;
                            b->args.replicas = share;
                            return jit_launch(ctx, fn, (int)(grid * share), T, (size_t)nb * 4, &b->args);
                          },
                          &out, nullptr));
  keep->reset(out);
  for (size_t s = 0; s < n_slots; s++) {
    b->args.col[s] = out->cols[s].values;
    if (b->sh.slots[s].nullable) b->args.valid[s] = out->cols[s].validity;
  }
  g->radix_batches++;
  return BGPU_OK;
}

static int run_two_pass(bgpu_agg* g, Bound* b, int64_t n_rows, const bgpu_batch* batch) {
  bgpu_ctx* ctx = g->ctx;
  PhaseTrace trace(ctx);
  int T = 256, R = 8, min_ctas = 2;  // 8 rows per lane in flight: the probes are DRAM round trips (R = 4: 7.1 ms intern, R = 8: 5.8 ms per 1e8 rows)
  if (const char* e = getenv("BGPU_TWOPASS_GEOMETRY")) {  // tuning knob: "threads,rows,ctas_per_sm"
    int t = 0, r = 0, c = 0;
    if (sscanf(e, "%d,%d,%d", &t, &r, &c) == 3 && t >= 32 && t <= 1024 && t % 32 == 0 && r >= 1 && r <= 16 && c >= 1 && c <= 8) {
      T = t; R = r; min_ctas = c;
    }
  }
  BGPU_TRY(specialise(g, b, 2, T, R, min_ctas));
  const int64_t tiles = ceil_div64(n_rows, (int64_t)T * R);
  // A batch with more rows than the table has room for may overflow it, and every overflow costs a full extra pass:
  // estimate the batch's distinct keys first (linear counting over a bitmap, one cheap pass over the key columns)
  // and size the table once.
  double est_groups = -1.0;  // distinct keys of this batch, when the sketch ran
  if ((uint64_t)n_rows + g->n_groups > g->max_groups) {
    const uint32_t words = next_pow2(std::max<uint64_t>(1024, std::min<uint64_t>((uint64_t)n_rows, 1ULL << 31) / 32));
    DevAllocPtr bitmap, cnt;
    BGPU_TRY(dev_alloc(ctx, (size_t)words * 4, &bitmap));
    BGPU_TRY(dev_alloc(ctx, 16, &cnt));
    BGPU_CUDA(ctx, cudaMemsetAsync(bitmap->p, 0, (size_t)words * 4, ctx->stream));
    BGPU_CUDA(ctx, cudaMemsetAsync(cnt->p, 0, 16, ctx->stream));
    refresh_pointers(g, b);
    b->args.gids = (uint32_t*)bitmap->p;
    b->args.n_groups = words;
    BGPU_TRY(jit_launch(ctx, b->mod->fn[JK_SKETCH], (int)std::min<int64_t>(tiles, (int64_t)ctx->sm_count * 8), T, 0, &b->args));
    popcount_kernel<<<agg_launch_grid(ctx, words, kAggThreads), kAggThreads, 0, ctx->stream>>>((const uint32_t*)bitmap->p, words,
                                                                                         (unsigned long long*)cnt->p);
    ctx->launches++;
    unsigned long long* pin = (unsigned long long*)ctx->pinned + 8;
    BGPU_CUDA(ctx, cudaMemcpyAsync(pin, cnt->p, 8, cudaMemcpyDeviceToHost, ctx->stream));
    BGPU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    const double m = (double)words * 32.0, set = (double)*pin;
    // linear counting: distinct ~ -m ln(zero fraction); a saturated bitmap means "at least as many as rows"
    const double est = set >= m ? (double)n_rows : std::min<double>((double)n_rows, -m * log(1.0 - set / m));
    // a first batch that is mostly distinct keys: the Partial pass would be all cost and no reduction
    // break-even: the Partial pass costs ~0.22 ns per input row, everything downstream ~0.33 ns per state row (2 GPUs,
    // config 4b), so passing rows through wins once distinct / rows exceeds ~1/3; measured at 0.43: 4.5-5.5 -> 7.9 G rows/s
    double ratio = 0.4;
    if (const char* e = getenv("BGPU_PASSTHROUGH_RATIO")) ratio = atof(e);
    if (!g->any_committed && g->pass_chunks.empty() && n_rows >= (1 << 20) && est > ratio * (double)n_rows &&
        passthrough_eligible(g, batch)) {
      g->passthrough = true;
      return passthrough_update(g, batch);
    }
    est_groups = est;
    const uint64_t need = (uint64_t)g->n_groups + (uint64_t)(est * 1.10) + 1024;
    if (need > g->max_groups) {
      if (need > 0x70000000ULL) return ctx->fail(BGPU_EUNSUPPORTED, "too many groups");
      trace.mark("sketch");
      BGPU_TRY(agg_resize(g, next_pow2(need), g->n_groups));
      trace.mark("resize");
    }
  }
  std::unique_ptr<bgpu_batch> reordered;  // the bound columns in bucket order (freed on the stream after the kernels)
  BGPU_TRY(radix_reorder(g, b, n_rows, est_groups < 0 ? (double)std::min<uint64_t>((uint64_t)n_rows, g->max_groups) : est_groups, T,
                         &reordered));
  if (reordered) trace.mark("radix reorder");
  DevAllocPtr gids;
  BGPU_TRY(dev_alloc(ctx, (size_t)n_rows * 4, &gids));
  b->args.gids = (uint32_t*)gids->p;
  for (int attempt = 0;; attempt++) {
    refresh_pointers(g, b);
    const int grid = (int)std::min<int64_t>(tiles, (int64_t)ctx->sm_count * 8);
    BGPU_TRY(jit_launch(ctx, b->mod->fn[JK_INTERN], grid, T, 0, &b->args));
    uint32_t h[8];
    BGPU_TRY(read_counters(g, h));
    BGPU_TRY(check_common_flags(g, h));
    if (!h[1]) {
      g->n_groups = h[0];
      break;
    }
    if (attempt >= 8) return ctx->fail(BGPU_EUNSUPPORTED, "too many groups");
    BGPU_TRY(grow_after_overflow(g, h[0], n_rows));
  }
  trace.mark("intern");
  g->any_committed = true;
  refresh_pointers(g, b);
  if (b->ap.n == 0) { g->last_path = "intern"; return BGPU_OK; }
  const uint32_t ng = g->n_groups;
  b->args.n_groups = ng;
  const size_t limit = std::min<size_t>(ctx->smem_optin, 200 * 1024);
  const size_t group_bytes = (size_t)b->sh.W * 8;
  if ((size_t)ng * group_bytes + (size_t)ng * 4 <= limit / 2) {
    int replicas = (int)std::min<size_t>(32, (limit / 2 - (size_t)ng * 4) / std::max<size_t>(1, (size_t)ng * group_bytes));
    if (replicas < 1) replicas = 1;
    const size_t smem = (size_t)ng * group_bytes * replicas + (size_t)ng * 4;
    const CUfunction fn = b->mod->fn[JK_ACC_SHARED];
    BGPU_TRY(jit_set_max_smem(ctx, fn, smem));
    int occ = 1;
    BGPU_TRY(cached_occupancy(ctx, fn, T, smem, &occ));
    const int grid = (int)std::min<int64_t>(tiles, (int64_t)ctx->sm_count * std::max(1, occ));
    b->args.replicas = (uint32_t)replicas;
    BGPU_TRY(jit_launch(ctx, fn, grid, T, smem, &b->args));
    g->last_path = "shared";
  } else {
    // hot-group cache in shared memory (skewed keys): the largest power of two of slots that fits ~40 KB, so that the
    // occupancy of the global tier stays what it was; BGPU_HOT_CACHE=0 switches it off (measurement)
    uint32_t cslots = 1024;
    const size_t slot_bytes = group_bytes + 8;
    while (cslots > 64 && cslots * slot_bytes > 40 * 1024) cslots >>= 1;
    if (const char* e = getenv("BGPU_HOT_CACHE")) cslots = (uint32_t)atoi(e);
    if (cslots & (cslots - 1)) cslots = 0;
    const size_t csmem = (size_t)cslots * slot_bytes;
    const int grid = (int)std::min<int64_t>(tiles, (int64_t)ctx->sm_count * 8);
    b->args.replicas = 1;
    b->args.cache_slots = cslots;
    BGPU_TRY(jit_launch(ctx, b->mod->fn[JK_ACC_GLOBAL], grid, T, csmem, &b->args));
    g->last_path = reordered ? "global_radix" : "global";
  }
  trace.mark("accumulate");
  return BGPU_OK;
}

static int agg_update_body(bgpu_agg* g, const bgpu_batch* batch, AggAsyncReq* req);

int agg_update_impl(bgpu_agg* g, const bgpu_batch* batch, AggAsyncReq* req) {
  g->key_too_long = false;
  const int rc = agg_update_body(g, batch, req);
  if (rc == BGPU_OK || !g->key_too_long) return rc;
  // A string key did not fit its inline field. Nothing of this stream has been aggregated yet: switch the key layout to
  // interned ids (keys of any width) and run the batch again. Later in a stream the error stands.
  g->key_too_long = false;
  if (g->any_committed || g->interned_mode || !g->pass_chunks.empty() || agg_key_layout(g, true) != BGPU_OK) return rc;
  const uint32_t zero[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  BGPU_TRY(write_counters(g, zero));
  g->n_groups = 0;
  g->fused_rung = 0;
  g->async_ready = false;
  BGPU_TRY(agg_resize(g, g->max_groups, 0));  // the slot size follows the key width
  return agg_update_body(g, batch, req);
}

static int agg_update_body(bgpu_agg* g, const bgpu_batch* batch, AggAsyncReq* req) {
  bgpu_ctx* ctx = g->ctx;
  if (req) req->launched = false;
  const int64_t n_rows = batch->n_rows;
  if (n_rows > 0xfffffff0LL) return ctx->fail(BGPU_EUNSUPPORTED, "more than 2^32 rows in one batch");
  Bound b;
  BGPU_TRY(bind_batch(g, batch, &b));
  if (n_rows == 0) return BGPU_OK;
  if (g->passthrough) {
    if (passthrough_eligible(g, batch)) return passthrough_update(g, batch);
    return ctx->fail(BGPU_EUNSUPPORTED, "a batch with nulls follows batches the Partial aggregate passed through");
  }
  if (b.sh.any_validity && !g->seen_tracking) {
    // first batch with nulls in an aggregate input: groups created by earlier (null-free) batches have seen
    // every accumulator; groups new in this batch start at 0
    BGPU_TRY(dev_alloc(ctx, (size_t)g->max_groups * 4, &g->seen));
    BGPU_CUDA(ctx, cudaMemsetAsync(g->seen->p, 0, (size_t)g->max_groups * 4, ctx->stream));
    if (g->n_groups) {
      fill_u32_kernel<<<agg_launch_grid(ctx, g->n_groups, kAggThreads), kAggThreads, 0, ctx->stream>>>(
          (uint32_t*)g->seen->p, g->n_groups, 0xffffffffu);
      ctx->launches++;
    }
    g->seen_tracking = true;
  }
  b.args.n_rows = n_rows;
  bool done = false;
  PhaseTrace trace(ctx);
  if (g->fused_rung < kFusedRungs) BGPU_TRY(try_fused(g, &b, n_rows, &done, req));
  trace.mark(done ? "fused" : "fused (void)");
  if (!done) {
    g->async_ready = false;
    BGPU_TRY(run_two_pass(g, &b, n_rows, batch));
  }
  return BGPU_OK;
}

int agg_async_result(bgpu_agg* g, AggAsyncReq* req, bool* replay) {
  bgpu_ctx* ctx = g->ctx;
  *replay = false;
  if (!req->launched) return BGPU_OK;
  BGPU_CUDA(ctx, cudaEventSynchronize(req->done));
  uint32_t h[8];
  for (int i = 0; i < 8; i++) h[i] = req->h_status[i];
  float ms = 0.f;
  if (cudaEventElapsedTime(&ms, req->k0, req->k1) == cudaSuccess) g->hot_kernel_ms += ms;
  req->launched = false;
  // h = {committed, n_groups, counters[1..6]}: counters[2] string key too long, counters[3] arithmetic overflow
  // a launch that did not commit (capacity, table growth, a key too long for the inline layout, an arithmetic overflow)
  // is run again synchronously: that path grows / switches / reports
  if (h[0]) {
    g->n_groups = h[1];
    g->any_committed = true;
    return BGPU_OK;
  }
  *replay = true;
  return BGPU_OK;
}

int agg_update_replay(bgpu_agg* g, const bgpu_batch* batch) {
  // the skipped launch left its abort flag up (sticky, so that later launches skip too): clear it, then take the
  // synchronous path, which grows the capacity / the table or changes rung as the batch needs
  uint32_t h[8];
  BGPU_TRY(read_counters(g, h));
  if (h[4]) {
    h[4] = 0;
    BGPU_TRY(write_counters(g, h));
  }
  g->async_ready = false;
  return agg_update_impl(g, batch, nullptr);
}

extern "C" int bgpu_agg_update(bgpu_agg* g, const bgpu_batch* batch) try {
  if (!g || !batch) return BGPU_EINVAL;
  bgpu_ctx* ctx = g->ctx;
  if (g->emitted) return ctx->fail(BGPU_EINVAL, "aggregate already emitted");
  BGPU_CUDA(ctx, cudaSetDevice(ctx->device));
  GpuTimeScope timed(ctx, T_KERNEL);
  if (g->ingest_drain) BGPU_TRY(g->ingest_drain(g));  // streamed input still in flight goes first
  return agg_update_impl(g, batch, nullptr);
} catch (const std::exception& e) {
  return bgpu::abi_guard(g ? g->ctx : nullptr, e);  // nothing is thrown across the C ABI
}

extern "C" const char* bgpu_agg_last_path(const bgpu_agg* g) { return g ? g->last_path.c_str() : ""; }
extern "C" double bgpu_agg_hot_kernel_ms(const bgpu_agg* g) { return g ? g->hot_kernel_ms : 0.0; }

// ---- emit ---------------------------------------------------------------------------------------------------------
// Two launches and one host round trip whatever the plan: densify_keys_kernel (table slots -> dense key arrays) and
// emit_all_kernel (every key column, accumulator column, AVG evaluation and validity bitmap); the counters and the OR of
// the key null masks come back together at the end. utf8 key columns of more groups than one CTA covers add a device scan
// and one unpack kernel, still without a round trip: their data buffer is sized by the packed key's width.
extern "C" int bgpu_agg_emit(bgpu_agg* g, bgpu_batch** out_batch) try {
  if (!g || !out_batch) return BGPU_EINVAL;
  *out_batch = nullptr;
  bgpu_ctx* ctx = g->ctx;
  if (g->emitted) return ctx->fail(BGPU_EINVAL, "aggregate already emitted");
  BGPU_CUDA(ctx, cudaSetDevice(ctx->device));
  GpuTimeScope timed(ctx, T_KERNEL);
  if (g->ingest_drain) BGPU_TRY(g->ingest_drain(g));  // end of input: everything pushed so far is aggregated first
  if (g->passthrough) {  // the state rows are the chunks themselves
    if (g->pass_chunks.size() == 1) {
      *out_batch = g->pass_chunks[0].release();
    } else {
      std::vector<const bgpu_batch*> parts;
      for (auto& c : g->pass_chunks) parts.push_back(c.get());
      BGPU_TRY(bgpu_batch_concat(ctx, parts.data(), (int32_t)parts.size(), out_batch));
    }
    g->pass_chunks.clear();
    g->emitted = true;
    return BGPU_OK;
  }
  const uint32_t n = g->n_groups, n1 = std::max<uint32_t>(n, 1);
  std::unique_ptr<bgpu_batch> out(new bgpu_batch(ctx));
  out->n_rows = n;
  PhaseTrace trace(ctx);
  const size_t bitmap_bytes = (size_t)ceil_div64(n1, 32) * 4;
  // one allocation for everything that is scratch or small: dense keys, metas, bitmaps come out of `pool`
  EmitArgs ea{};
  DevAllocPtr d_key0, d_key1, d_kmeta;
  BGPU_TRY(dev_alloc(ctx, (size_t)n1 * 8, &d_key0));
  if (g->key_words > 1) BGPU_TRY(dev_alloc(ctx, (size_t)n1 * 8, &d_key1));
  BGPU_TRY(dev_alloc(ctx, (size_t)n1 * 4, &d_kmeta));
  if (n) {
    bgpu::GroupTable t = g->table();
    t.max_groups = n;  // ids handed out by a void (overflowed) pass are not groups
    densify_keys_kernel<<<agg_launch_grid(ctx, (int64_t)g->max_groups * 2, kAggThreads), kAggThreads, 0, ctx->stream>>>(
        t, (uint64_t*)d_key0->p, d_key1 ? (uint64_t*)d_key1->p : nullptr, (uint32_t*)d_kmeta->p);
    ctx->launches++;
    BGPU_CUDA(ctx, cudaGetLastError());
  }
  ea.key0 = (const uint64_t*)d_key0->p;
  ea.key1 = d_key1 ? (const uint64_t*)d_key1->p : nullptr;
  ea.kmeta = (const uint32_t*)d_kmeta->p;
  ea.acc = (const uint64_t*)g->acc->p;
  ea.seen = g->seen_tracking ? (const uint32_t*)g->seen->p : nullptr;
  ea.flags = g->d_flags();
  ea.key_null_or = (uint32_t*)g->counters->p + 5;
  ea.stride = g->acc_stride;
  const bool single = n <= 4 * kEmitThreads;
  // group key columns
  struct Utf8Post { size_t col; DevAllocPtr lens; };
  std::vector<Utf8Post> utf8;
  for (size_t i = 0; i < g->group_cols.size(); i++) {
    const Column& f = g->fields[(size_t)g->group_cols[i]];
    Column c;
    c.name = f.name; c.format = f.format; c.flags = f.flags | ARROW_FLAG_NULLABLE; c.kind = f.kind; c.width = f.width;
    c.length = n; c.null_count = -1;
    EmitKeyCol& kc = ea.k[i];
    kc.kind = f.kind; kc.width = f.width; kc.byte_off = g->key_off[i];
    DevAllocPtr valid;
    BGPU_TRY(dev_alloc(ctx, bitmap_bytes, &valid));
    c.owned.push_back(valid);
    c.validity = (uint8_t*)valid->p;
    kc.valid = (uint32_t*)valid->p;
    if (g->key_interned[i]) {
      // the packed key holds the value's id: emitted as Int32, looked up in the interner after the kernel
      DevAllocPtr ids;
      BGPU_TRY(dev_alloc(ctx, (size_t)n1 * 4, &ids));
      kc.kind = K_FIXED; kc.width = 4;
      kc.values = ids->p;
      c.values = ids->p;
      c.owned.push_back(ids);
    } else if (f.kind == K_UTF8 || f.kind == K_BINARY) {
      DevAllocPtr offs, data;
      const size_t cap = (size_t)n1 * (size_t)std::max(1, g->key_width[i]) + 16;  // the packed key bounds every string
      BGPU_TRY(dev_alloc(ctx, (size_t)(n + 1) * 4, &offs));
      BGPU_TRY(dev_alloc(ctx, cap, &data));
      kc.offsets = (int32_t*)offs->p; kc.data = (uint8_t*)data->p;
      if (!single) {
        Utf8Post up; up.col = i;
        BGPU_TRY(dev_alloc(ctx, (size_t)n1 * 4, &up.lens));
        kc.lens = (int32_t*)up.lens->p;
        utf8.push_back(up);
      } else if (!n) {
        BGPU_CUDA(ctx, cudaMemsetAsync(offs->p, 0, 4, ctx->stream));
      }
      c.offsets = (int32_t*)offs->p; c.values = data->p; c.data_bytes = (int64_t)cap - 16;
      c.owned.push_back(offs); c.owned.push_back(data);
    } else {
      DevAllocPtr vals;
      BGPU_TRY(dev_alloc(ctx, f.kind == K_BOOL ? bitmap_bytes : (size_t)n1 * f.width, &vals));
      kc.values = vals->p;
      c.values = vals->p;
      c.owned.push_back(vals);
      if (f.kind == K_DICT32) BGPU_TRY(upload_canonical_dictionary(g, i, &c.dict));
    }
    out->cols.push_back(std::move(c));
  }
  ea.n_key = (int32_t)g->group_cols.size();
  // aggregate columns
  int n_acc = 0;
  auto acc_column = [&](int type, int phys_idx, int count_phys, const LogicalAgg* la, const std::string& name,
                        const std::string& format, bool nullable_by_seen) -> int {
    if (n_acc >= kMaxEmitAcc) return ctx->fail(BGPU_EUNSUPPORTED, "too many output columns");
    const PhysAcc& p = g->phys[(size_t)phys_idx];
    Column c;
    c.name = name; c.format = format; c.flags = ARROW_FLAG_NULLABLE; c.kind = K_FIXED; c.length = n; c.null_count = 0;
    EmitAccCol& ac = ea.a[n_acc++];
    ac.type = type; ac.words = p.words; ac.off = p.out_off; ac.seen_bit = phys_idx;
    c.width = type == 0 ? p.words * 8 : (type == 2 ? 16 : 8);
    DevAllocPtr vals;
    BGPU_TRY(dev_alloc(ctx, (size_t)n1 * c.width, &vals));
    ac.out = (uint64_t*)vals->p;
    c.values = vals->p;
    c.owned.push_back(vals);
    if (type != 0) {
      ac.count_off = g->phys[(size_t)count_phys].out_off;
      ac.scale_up = la->scale_up;
    }
    if (type != 0 || (nullable_by_seen && g->seen_tracking)) {
      DevAllocPtr valid;
      BGPU_TRY(dev_alloc(ctx, bitmap_bytes, &valid));
      ac.valid = (uint32_t*)valid->p;
      c.validity = (uint8_t*)valid->p;
      c.null_count = -1;
      c.owned.push_back(valid);
    }
    out->cols.push_back(std::move(c));
    return BGPU_OK;
  };
  for (const LogicalAgg& a : g->aggs) {
    if (!g->final_mode && g->mode != BGPU_AGG_MODE_SINGLE) {  // Partial: state columns
      if (a.fn == BGPU_AGG_COUNT) {
        BGPU_TRY(acc_column(0, a.phys[0], -1, &a, a.name + "[count]", "l", false));
      } else if (a.fn == BGPU_AGG_SUM) {
        BGPU_TRY(acc_column(0, a.phys[0], -1, &a, a.name + "[sum]", a.sum_format, true));
      } else {
        BGPU_TRY(acc_column(0, a.phys[0], -1, &a, a.name + "[count]", "L", false));
        BGPU_TRY(acc_column(0, a.phys[1], -1, &a, a.name + "[sum]", a.sum_format, true));
      }
      continue;
    }
    if (a.fn == BGPU_AGG_COUNT) BGPU_TRY(acc_column(0, a.phys[0], -1, &a, a.name, "l", false));
    else if (a.fn == BGPU_AGG_SUM) BGPU_TRY(acc_column(0, a.phys[0], -1, &a, a.name, a.result_format, true));
    else BGPU_TRY(acc_column(a.decimal ? 2 : 1, a.phys[1], a.phys[0], &a, a.name, a.result_format, false));  // Avg: evaluate
  }
  ea.n_acc = n_acc;
  if (n) {
    if (single) emit_all_kernel<true><<<1, kEmitThreads, 0, ctx->stream>>>(ea, n);
    else emit_all_kernel<false><<<agg_launch_grid(ctx, n, kEmitThreads), kEmitThreads, 0, ctx->stream>>>(ea, n);
    ctx->launches++;
    BGPU_CUDA(ctx, cudaGetLastError());
    for (const Utf8Post& up : utf8) {  // lengths -> offsets -> bytes, all on the stream
      Column& c = out->cols[up.col];
      BGPU_CUDA(ctx, cudaMemsetAsync(c.offsets, 0, 4, ctx->stream));
      size_t tmp_bytes = 0;
      cub::DeviceScan::InclusiveSum(nullptr, tmp_bytes, (const int32_t*)up.lens->p, c.offsets + 1, (int)n, ctx->stream);
      DevAllocPtr tmp;
      BGPU_TRY(dev_alloc(ctx, tmp_bytes, &tmp));
      BGPU_CUDA(ctx, cub::DeviceScan::InclusiveSum(tmp->p, tmp_bytes, (const int32_t*)up.lens->p, c.offsets + 1, (int)n, ctx->stream));
      unpack_utf8_kernel<<<agg_launch_grid(ctx, n, kAggThreads), kAggThreads, 0, ctx->stream>>>(
          ea.key0, ea.key1, ea.kmeta, n, (int)up.col, g->key_off[up.col], 1, c.offsets, (uint8_t*)c.values, nullptr);
      ctx->launches += 2;
      BGPU_CUDA(ctx, cudaGetLastError());
    }
  }
  for (size_t i = 0; i < g->group_cols.size(); i++) {  // interned key columns: ids -> values
    if (!g->key_interned[i]) continue;
    const Column& f = g->fields[(size_t)g->group_cols[i]];
    Column proto;
    proto.name = f.name; proto.format = f.format; proto.flags = f.flags; proto.kind = f.kind; proto.width = f.width;
    Column looked_up;
    if (!g->interners[i]) BGPU_TRY(interner_create(ctx, &g->interners[i]));  // a stream without rows
    const Column& idc = out->cols[i];
    BGPU_TRY(interner_take(g->interners[i].get(), (const int32_t*)idc.values, (const uint32_t*)idc.validity, n, proto, &looked_up));
    looked_up.owned.insert(looked_up.owned.end(), idc.owned.begin(), idc.owned.end());
    out->cols[i] = std::move(looked_up);
  }
  trace.mark("emit kernels");
  uint32_t h[8];
  BGPU_TRY(read_counters(g, h));  // the one host round trip of emit
  if (h[3]) return ctx->fail(BGPU_EOVERFLOW, "Arithmetic overflow in a Decimal128 aggregate argument or AVG");
  // key columns no group has a null in are emitted without a validity bitmap: the state batch of the common null-free
  // case is then plain fixed-width data, which is what the fused repartition + exchange kernel takes
  const uint32_t key_null_mask = n ? h[5] : 0;
  for (size_t i = 0; i < g->group_cols.size(); i++)
    if (!((key_null_mask >> i) & 1)) {
      out->cols[i].validity = nullptr;
      out->cols[i].null_count = 0;
    }
  trace.mark("emit readback");
  g->emitted = true;
  *out_batch = out.release();
  return BGPU_OK;
} catch (const std::exception& e) {
  return bgpu::abi_guard(g ? g->ctx : nullptr, e);  // nothing is thrown across the C ABI
}
