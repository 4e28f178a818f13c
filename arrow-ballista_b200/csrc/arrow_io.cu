// RecordBatch <-> device: Arrow C-Data import (upload / wrap) and export (download / device export).
// The types accepted are those on the aggregation+shuffle path: fixed-width primitives,
// Decimal128, Boolean, Utf8/Binary, Dictionary<Int32, Utf8>.
#include <stdlib.h>
#include <string.h>

#include "internal.h"

int parse_field(bgpu_ctx* ctx, const struct ArrowSchema* s, Column* c) {
  const char* f = s->format ? s->format : "";
  c->name = s->name ? s->name : "";
  c->format = f;
  c->flags = s->flags;
  c->dict = nullptr;
  auto fixed = [&](int w) {
    c->kind = K_FIXED;
    c->width = w;
    return BGPU_OK;
  };
  if (s->dictionary) {
    if (strcmp(f, "i") != 0) return ctx->fail(BGPU_EUNSUPPORTED, "dictionary index type must be Int32, got " + c->format);
    c->kind = K_DICT32;
    c->width = 4;
    c->dict = std::make_shared<Column>();
    return parse_field(ctx, s->dictionary, c->dict.get());
  }
  switch (f[0]) {
    case 'c': case 'C': return fixed(1);
    case 's': case 'S': return fixed(2);
    case 'i': case 'I': case 'f': return fixed(4);
    case 'l': case 'L': case 'g': return fixed(8);
    case 'e': return fixed(2);
    case 'b': c->kind = K_BOOL; c->width = 0; return BGPU_OK;
    case 'u': c->kind = K_UTF8; c->width = 0; return BGPU_OK;
    case 'z': c->kind = K_BINARY; c->width = 0; return BGPU_OK;
    case 'd': {  // d:precision,scale[,bitwidth]
      int p = 0, sc = 0, bw = 128;
      int n = sscanf(f, "d:%d,%d,%d", &p, &sc, &bw);
      if (n < 2 || bw != 128) return ctx->fail(BGPU_EUNSUPPORTED, "only Decimal128 is supported, got " + c->format);
      return fixed(16);
    }
    case 't':
      switch (f[1]) {
        case 'd': return fixed(f[2] == 'D' ? 4 : 8);                 // date32 / date64
        case 't': return fixed((f[2] == 's' || f[2] == 'm') ? 4 : 8);  // time32 / time64
        case 's': case 'D': return fixed(8);                         // timestamp / duration
        default: break;
      }
      break;
    default: break;
  }
  return ctx->fail(BGPU_EUNSUPPORTED, "unsupported Arrow type format '" + c->format + "' for column " + c->name);
}

static size_t values_bytes(const Column& c, int64_t n) {
  if (c.kind == K_BOOL) return (size_t)((n + 7) / 8);
  return (size_t)n * (size_t)c.width;
}

// Copy a bitmap of n bits starting at bit `off` into a fresh host buffer aligned at bit 0.
static std::vector<uint8_t> realign_bits(const uint8_t* src, int64_t off, int64_t n) {
  std::vector<uint8_t> out((size_t)((n + 7) / 8), 0);
  if ((off & 7) == 0) {
    memcpy(out.data(), src + (off >> 3), out.size());
  } else {
    for (int64_t i = 0; i < n; i++)
      if ((src[(off + i) >> 3] >> ((off + i) & 7)) & 1) out[(size_t)(i >> 3)] |= (uint8_t)(1u << (i & 7));
  }
  return out;
}

static int h2d(bgpu_ctx* ctx, const void* src, size_t bytes, Column* c, void** dst) {
  DevAllocPtr a;
  BGPU_TRY(dev_alloc(ctx, bytes, &a));
  if (bytes) BGPU_CUDA(ctx, cudaMemcpyAsync(a->p, src, bytes, cudaMemcpyHostToDevice, ctx->stream));
  c->owned.push_back(a);
  *dst = a->p;
  return BGPU_OK;
}

// Import one column. device=true: buffers are device pointers and are borrowed.
static int import_column(bgpu_ctx* ctx, const struct ArrowArray* a, const struct ArrowSchema* s, bool device,
                         Column* c) {
  BGPU_TRY(parse_field(ctx, s, c));
  c->length = a->length;
  c->null_count = a->null_count;
  const int64_t off = a->offset, n = a->length;
  const uint8_t* validity = (a->n_buffers > 0) ? (const uint8_t*)a->buffers[0] : nullptr;
  if (a->null_count == 0) validity = nullptr;
  if (device) {
    if (off != 0) return ctx->fail(BGPU_EINVAL, "bgpu_batch_wrap_device requires array offset 0");
    c->validity = (uint8_t*)validity;
    if (c->kind == K_UTF8 || c->kind == K_BINARY) {
      if (a->n_buffers < 3) return ctx->fail(BGPU_EINVAL, "utf8 column needs 3 buffers");
      c->offsets = (int32_t*)a->buffers[1];
      c->values = (void*)a->buffers[2];
      c->data_bytes = -1;  // unknown without a device read; filled lazily by users that need it
    } else {
      if (a->n_buffers < 2) return ctx->fail(BGPU_EINVAL, "primitive column needs 2 buffers");
      c->values = (void*)a->buffers[1];
    }
  } else {
    // pageable or pinned host memory; synchronous staging copies must complete before we return because the
    // temporaries die here -> use the realigned vector only for bitmaps and sync below.
    std::vector<std::vector<uint8_t>> temps;
    if (validity) {
      temps.push_back(realign_bits(validity, off, n));
      BGPU_TRY(h2d(ctx, temps.back().data(), temps.back().size(), c, (void**)&c->validity));
    }
    if (c->kind == K_UTF8 || c->kind == K_BINARY) {
      if (a->n_buffers < 3) return ctx->fail(BGPU_EINVAL, "utf8 column needs 3 buffers");
      const int32_t* offs = (const int32_t*)a->buffers[1] + off;
      const int32_t first = n ? offs[0] : 0, last = n ? offs[n] : 0;
      BGPU_TRY(h2d(ctx, offs, (size_t)(n + 1) * 4, c, (void**)&c->offsets));
      // keep offsets un-rebased: point `values` so that values[offsets[i]] is row i's first byte
      void* data = nullptr;
      BGPU_TRY(h2d(ctx, (const uint8_t*)a->buffers[2] + first, (size_t)(last - first), c, &data));
      c->values = (uint8_t*)data - first;
      c->data_bytes = last;
      c->data_first = first;
    } else if (c->kind == K_BOOL) {
      temps.push_back(realign_bits((const uint8_t*)a->buffers[1], off, n));
      BGPU_TRY(h2d(ctx, temps.back().data(), temps.back().size(), c, &c->values));
    } else {
      BGPU_TRY(h2d(ctx, (const uint8_t*)a->buffers[1] + (size_t)off * c->width, values_bytes(*c, n), c, &c->values));
    }
    if (!temps.empty()) BGPU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  }
  if (c->kind == K_DICT32) {
    if (!a->dictionary || !s->dictionary) return ctx->fail(BGPU_EINVAL, "dictionary column without dictionary");
    BGPU_TRY(import_column(ctx, a->dictionary, s->dictionary, device, c->dict.get()));
  }
  return BGPU_OK;
}

static int import_batch(bgpu_ctx* ctx, const struct ArrowArray* a, const struct ArrowSchema* s, bool device,
                        bgpu_batch** out) {
  if (!ctx || !a || !s || !out) return BGPU_EINVAL;
  *out = nullptr;
  if (!ctx->has_device) return ctx->fail(BGPU_ENODEVICE, "context has no device (libbgpu has no CPU fallback)");
  if (!s->format || strcmp(s->format, "+s") != 0)
    return ctx->fail(BGPU_EINVAL, "expected a struct-typed (RecordBatch) ArrowArray, format '+s'");
  if (a->n_children != s->n_children) return ctx->fail(BGPU_EINVAL, "array/schema children mismatch");
  if (a->offset != 0) return ctx->fail(BGPU_EINVAL, "struct-level offset is not supported");
  BGPU_CUDA(ctx, cudaSetDevice(ctx->device));
  std::unique_ptr<bgpu_batch> b(new bgpu_batch(ctx));
  b->n_rows = a->length;
  b->cols.resize((size_t)a->n_children);
  for (int64_t i = 0; i < a->n_children; i++) {
    if (a->children[i]->length != a->length) return ctx->fail(BGPU_EINVAL, "column length mismatch");
    BGPU_TRY(import_column(ctx, a->children[i], s->children[i], device, &b->cols[(size_t)i]));
  }
  // Host buffers are borrowed "for the call" (include/bgpu.h): every copy out of them must have completed before the
  // caller may release or reuse them. Pageable memory is staged synchronously by the driver, pinned memory is not.
  if (!device) BGPU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  *out = b.release();
  return BGPU_OK;
}

extern "C" int bgpu_batch_upload(bgpu_ctx* ctx, const struct ArrowArray* a, const struct ArrowSchema* s,
                                 bgpu_batch** out) try {
  if (ctx && ctx->has_device) cudaSetDevice(ctx->device);
  GpuTimeScope timed(ctx, T_H2D);
  return import_batch(ctx, a, s, false, out);
} catch (const std::exception& e) {
  return bgpu::abi_guard(ctx, e);  // nothing is thrown across the C ABI
}
extern "C" int bgpu_batch_wrap_device(bgpu_ctx* ctx, const struct ArrowArray* a, const struct ArrowSchema* s,
                                      bgpu_batch** out) try {
  return import_batch(ctx, a, s, true, out);
} catch (const std::exception& e) {
  return bgpu::abi_guard(ctx, e);  // nothing is thrown across the C ABI
}
extern "C" int64_t bgpu_batch_num_rows(const bgpu_batch* b) { return b ? b->n_rows : -1; }
extern "C" int32_t bgpu_batch_num_columns(const bgpu_batch* b) { return b ? (int32_t)b->cols.size() : -1; }
static int64_t column_memory_size(const Column& c, int64_t n) {
  int64_t b = c.validity ? (n + 7) / 8 : 0;
  if (c.kind == K_UTF8 || c.kind == K_BINARY) b += (n + 1) * 4 + (c.data_bytes > 0 ? c.data_bytes - c.data_first : 0);
  else if (c.kind == K_BOOL) b += (n + 7) / 8;
  else b += n * c.width;
  if (c.kind == K_DICT32 && c.dict) b += column_memory_size(*c.dict, c.dict->length);
  return b;
}
extern "C" int64_t bgpu_batch_memory_size(const bgpu_batch* b) {
  if (!b) return -1;
  int64_t total = 0;
  for (const Column& c : b->cols) total += column_memory_size(c, b->n_rows);
  return total;
}
extern "C" void bgpu_batch_destroy(bgpu_batch* b) {
  if (!b) return;
  if (b->ctx) cudaSetDevice(b->ctx->device);
  delete b;
}

// ---- export ------------------------------------------------------------------------------------------
namespace {
struct SchemaPriv {
  std::string format, name;
  std::vector<ArrowSchema*> children;
  ArrowSchema* dict = nullptr;
};
void release_schema(ArrowSchema* s) {
  if (!s || !s->release) return;
  SchemaPriv* p = (SchemaPriv*)s->private_data;
  for (ArrowSchema* c : p->children) {
    if (c->release) c->release(c);
    delete c;
  }
  if (p->dict) {
    if (p->dict->release) p->dict->release(p->dict);
    delete p->dict;
  }
  delete p;
  s->release = nullptr;
}
void fill_schema(ArrowSchema* s, const std::string& format, const std::string& name, int64_t flags) {
  SchemaPriv* p = new SchemaPriv();
  p->format = format;
  p->name = name;
  memset(s, 0, sizeof(*s));
  s->format = p->format.c_str();
  s->name = p->name.c_str();
  s->flags = flags;
  s->release = release_schema;
  s->private_data = p;
}
void column_schema(const Column& c, ArrowSchema* s) {
  fill_schema(s, c.format, c.name, c.flags);
  if (c.kind == K_DICT32) {
    SchemaPriv* p = (SchemaPriv*)s->private_data;
    p->dict = new ArrowSchema();
    column_schema(*c.dict, p->dict);
    s->dictionary = p->dict;
  }
}
void batch_schema(const bgpu_batch* b, ArrowSchema* s) {
  fill_schema(s, "+s", "", 0);
  SchemaPriv* p = (SchemaPriv*)s->private_data;
  for (const Column& c : b->cols) {
    ArrowSchema* cs = new ArrowSchema();
    column_schema(c, cs);
    p->children.push_back(cs);
  }
  s->n_children = (int64_t)p->children.size();
  s->children = p->children.data();
}

struct ArrayPriv {
  std::vector<const void*> buffers;
  std::vector<void*> pinned;     // host buffers to cudaFreeHost on release
  std::vector<DevAllocPtr> keep; // device buffers kept alive (device export)
  std::vector<ArrowArray*> children;
  ArrowArray* dict = nullptr;
  // fast download: every buffer of the batch lives in one pinned block of the context's pool (top-level array only)
  bgpu_ctx* pool_ctx = nullptr;
  void* pool_block = nullptr;
  size_t pool_bytes = 0;
};
void release_array(ArrowArray* a) {
  if (!a || !a->release) return;
  ArrayPriv* p = (ArrayPriv*)a->private_data;
  for (ArrowArray* c : p->children) {
    if (c->release) c->release(c);
    delete c;
  }
  if (p->dict) {
    if (p->dict->release) p->dict->release(p->dict);
    delete p->dict;
  }
  for (void* h : p->pinned) cudaFreeHost(h);
  if (p->pool_ctx) {
    ctx_pinned_put(p->pool_ctx, p->pool_block, p->pool_bytes);
    ctx_release(p->pool_ctx);
  }
  delete p;
  a->release = nullptr;
}
ArrayPriv* fill_array(ArrowArray* a, int64_t length, int64_t null_count, int64_t offset) {
  ArrayPriv* p = new ArrayPriv();
  memset(a, 0, sizeof(*a));
  a->length = length;
  a->null_count = null_count;
  a->offset = offset;
  a->release = release_array;
  a->private_data = p;
  return p;
}
void finish_array(ArrowArray* a) {
  ArrayPriv* p = (ArrayPriv*)a->private_data;
  a->n_buffers = (int64_t)p->buffers.size();
  a->buffers = p->buffers.data();
  a->n_children = (int64_t)p->children.size();
  a->children = p->children.data();
  a->dictionary = p->dict;
}

int d2h(bgpu_ctx* ctx, ArrayPriv* p, const void* src, size_t bytes, void** out) {
  void* h = nullptr;
  BGPU_CUDA(ctx, cudaMallocHost(&h, bytes ? bytes : 8));
  p->pinned.push_back(h);
  if (bytes) BGPU_CUDA(ctx, cudaMemcpyAsync(h, src, bytes, cudaMemcpyDeviceToHost, ctx->stream));
  *out = h;
  return BGPU_OK;
}

// Download rows [row0, row0+n) of column c. The exported array carries offset o = row0 % 8 so that validity and
// boolean bitmaps can be copied bytewise.
int download_column(bgpu_ctx* ctx, const Column& c, int64_t row0, int64_t n, ArrowArray* a) {
  const int64_t o = row0 & 7, start = row0 - o, cnt = n + o;
  ArrayPriv* p = fill_array(a, n, c.validity ? -1 : 0, o);
  void* h = nullptr;
  if (c.validity) {
    BGPU_TRY(d2h(ctx, p, c.validity + (start >> 3), (size_t)((cnt + 7) / 8), &h));
    p->buffers.push_back(h);
  } else {
    p->buffers.push_back(nullptr);
  }
  if (c.kind == K_UTF8 || c.kind == K_BINARY) {
    int32_t* offs = nullptr;
    BGPU_TRY(d2h(ctx, p, c.offsets + start, (size_t)(cnt + 1) * 4, (void**)&offs));
    BGPU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    const int32_t first = offs[0], last = offs[cnt];
    BGPU_TRY(d2h(ctx, p, (const uint8_t*)c.values + first, (size_t)(last - first), &h));
    if (first)
      for (int64_t i = 0; i <= cnt; i++) offs[i] -= first;
    p->buffers.push_back(offs);
    p->buffers.push_back(h);
  } else if (c.kind == K_BOOL) {
    BGPU_TRY(d2h(ctx, p, (const uint8_t*)c.values + (start >> 3), (size_t)((cnt + 7) / 8), &h));
    p->buffers.push_back(h);
  } else {
    BGPU_TRY(d2h(ctx, p, (const uint8_t*)c.values + (size_t)start * c.width, (size_t)cnt * c.width, &h));
    p->buffers.push_back(h);
  }
  if (c.kind == K_DICT32) {
    p->dict = new ArrowArray();
    BGPU_TRY(download_column(ctx, *c.dict, 0, c.dict->length, p->dict));
  }
  finish_array(a);
  return BGPU_OK;
}

void device_column(const Column& c, ArrowArray* a) {
  ArrayPriv* p = fill_array(a, c.length, c.validity ? c.null_count : 0, 0);
  p->keep = c.owned;
  p->buffers.push_back(c.validity);
  if (c.kind == K_UTF8 || c.kind == K_BINARY) p->buffers.push_back(c.offsets);
  p->buffers.push_back(c.values);
  if (c.kind == K_DICT32) {
    p->dict = new ArrowArray();
    device_column(*c.dict, p->dict);
  }
  finish_array(a);
}
}  // namespace

// schema of a batch as an Arrow C schema (used by the IPC reader for streams without batches)
void export_batch_schema(const bgpu_batch* b, struct ArrowSchema* out) { batch_schema(b, out); }

// ---- fast download of small batches --------------------------------------------------------------------------------------
// Results and aggregate states are a few rows by a dozen columns: pinning a host buffer per Arrow buffer and one copy per
// buffer cost more than the data. Here every buffer of the batch goes into ONE pinned block of the context's pool:
// a gather kernel packs them into a device staging buffer, one copy brings them over, one synchronise.
namespace {
constexpr int kMaxGatherJobs = 96;
constexpr size_t kFastDownloadBytes = 1u << 20;
struct GatherJobs {
  const uint8_t* src[kMaxGatherJobs];
  uint32_t dst[kMaxGatherJobs], bytes[kMaxGatherJobs];
};
__global__ void __launch_bounds__(256) gather_kernel(const __grid_constant__ GatherJobs j, uint8_t* __restrict__ out) {
  const uint8_t* s = j.src[blockIdx.x];
  uint8_t* d = out + j.dst[blockIdx.x];
  for (uint32_t i = threadIdx.x; i < j.bytes[blockIdx.x]; i += 256) d[i] = s[i];
}
struct FastPlan {
  GatherJobs jobs;
  int n = 0;
  size_t total = 0;
  // returns the offset of the buffer inside the block, or -1 when the plan is full
  int64_t add(const void* src, size_t bytes) {
    if (n >= kMaxGatherJobs || total + bytes > kFastDownloadBytes) return -1;
    jobs.src[n] = (const uint8_t*)src;
    jobs.dst[n] = (uint32_t)total;
    jobs.bytes[n] = (uint32_t)bytes;
    n++;
    const int64_t off = (int64_t)total;
    total = (total + bytes + 63) & ~(size_t)63;
    return off;
  }
};
// Plan column c (whole column: rows [0, c.length)); buffer offsets into `offs` in Arrow buffer order (-1 = absent).
bool plan_column(const Column& c, FastPlan* fp, std::vector<int64_t>* offs) {
  const int64_t n = c.length;
  int64_t o = -1;
  if (c.validity && (o = fp->add(c.validity, (size_t)((n + 7) / 8))) < 0) return false;
  offs->push_back(c.validity ? o : -1);
  if (c.kind == K_UTF8 || c.kind == K_BINARY) {
    if (c.data_bytes < 0 || c.data_first != 0) return false;  // unknown extent (borrowed device buffers): the general path reads the offsets first
    if ((o = fp->add(c.offsets, (size_t)(n + 1) * 4)) < 0) return false;
    offs->push_back(o);
    if ((o = fp->add(c.values, (size_t)c.data_bytes)) < 0) return false;
    offs->push_back(o);
  } else {
    const size_t bytes = c.kind == K_BOOL ? (size_t)((n + 7) / 8) : (size_t)n * c.width;
    if ((o = fp->add(c.values, bytes)) < 0) return false;
    offs->push_back(o);
  }
  if (c.kind == K_DICT32) return plan_column(*c.dict, fp, offs);
  return true;
}
void build_column(const Column& c, const uint8_t* block, const std::vector<int64_t>& offs, size_t* k, ArrowArray* a) {
  ArrayPriv* p = fill_array(a, c.length, c.validity ? -1 : 0, 0);
  const int n_buf = (c.kind == K_UTF8 || c.kind == K_BINARY) ? 3 : 2;
  for (int i = 0; i < n_buf; i++) {
    const int64_t o = offs[(*k)++];
    p->buffers.push_back(o < 0 ? nullptr : block + o);
  }
  if (c.kind == K_DICT32) {
    p->dict = new ArrowArray();
    build_column(*c.dict, block, offs, k, p->dict);
  }
  finish_array(a);
}
// *done = false: not eligible (too large, sliced, unknown utf8 extents) - take the general path
int download_fast(bgpu_ctx* ctx, const bgpu_batch* b, struct ArrowArray* out, bool* done) {
  *done = false;
  FastPlan fp;
  std::vector<int64_t> offs;
  for (const Column& c : b->cols)
    if (!plan_column(c, &fp, &offs)) return BGPU_OK;
  size_t block_bytes = 64 << 10;
  while (block_bytes < fp.total) block_bytes <<= 1;
  void* block = nullptr;
  BGPU_TRY(ctx_pinned_get(ctx, block_bytes, &block));
  int rc = BGPU_OK;
  if (fp.n) {
    DevAllocPtr stage;
    rc = dev_alloc(ctx, fp.total, &stage);
    if (rc == BGPU_OK) {
      gather_kernel<<<fp.n, 256, 0, ctx->stream>>>(fp.jobs, (uint8_t*)stage->p);
      ctx->launches++;
      cudaError_t e = cudaGetLastError();
      if (e == cudaSuccess) e = cudaMemcpyAsync(block, stage->p, fp.total, cudaMemcpyDeviceToHost, ctx->stream);
      if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
      if (e != cudaSuccess) rc = ctx->fail(BGPU_ECUDA, std::string("download: ") + cudaGetErrorString(e));
    }
  }
  if (rc != BGPU_OK) {
    ctx_pinned_put(ctx, block, block_bytes);
    return rc;
  }
  ArrayPriv* p = fill_array(out, b->n_rows, 0, 0);
  p->buffers.push_back(nullptr);
  p->pool_ctx = ctx; p->pool_block = block; p->pool_bytes = block_bytes;
  ctx_retain(ctx);
  size_t k = 0;
  for (const Column& c : b->cols) {
    ArrowArray* ca = new ArrowArray();
    build_column(c, (const uint8_t*)block, offs, &k, ca);
    p->children.push_back(ca);
  }
  finish_array(out);
  *done = true;
  return BGPU_OK;
}
}  // namespace

extern "C" int bgpu_batch_download(bgpu_ctx* ctx, const bgpu_batch* b, int64_t row_offset, int64_t num_rows,
                                   struct ArrowArray* out, struct ArrowSchema* out_schema) try {
  if (!ctx || !b || !out || !out_schema) return BGPU_EINVAL;
  if (row_offset < 0 || num_rows < 0 || row_offset + num_rows > b->n_rows)
    return ctx->fail(BGPU_EINVAL, "row range out of bounds");
  BGPU_CUDA(ctx, cudaSetDevice(ctx->device));
  GpuTimeScope timed(ctx, T_D2H);
  if (row_offset == 0 && num_rows == b->n_rows && !getenv("BGPU_NO_FAST_DOWNLOAD")) {
    bool done = false;
    BGPU_TRY(download_fast(ctx, b, out, &done));
    if (done) {
      batch_schema(b, out_schema);
      return BGPU_OK;
    }
  }
  ArrayPriv* p = fill_array(out, num_rows, 0, row_offset & 7);
  p->buffers.push_back(nullptr);
  int rc = BGPU_OK;
  for (const Column& c : b->cols) {
    ArrowArray* ca = new ArrowArray();
    memset(ca, 0, sizeof(*ca));
    p->children.push_back(ca);
    rc = download_column(ctx, c, row_offset, num_rows, ca);
    if (rc != BGPU_OK) break;
  }
  finish_array(out);
  // struct-level offset applies on top of children in Arrow semantics; children already carry it -> reset.
  out->offset = 0;
  if (rc == BGPU_OK) {
    cudaError_t e = cudaStreamSynchronize(ctx->stream);
    if (e != cudaSuccess) rc = ctx->fail(BGPU_ECUDA, std::string("download sync: ") + cudaGetErrorString(e));
  }
  if (rc != BGPU_OK) {
    release_array(out);
    return rc;
  }
  batch_schema(b, out_schema);
  return BGPU_OK;
} catch (const std::exception& e) {
  return bgpu::abi_guard(ctx, e);  // nothing is thrown across the C ABI
}

extern "C" int bgpu_batch_export_device(bgpu_ctx* ctx, const bgpu_batch* b, struct ArrowArray* out,
                                        struct ArrowSchema* out_schema) try {
  if (!ctx || !b || !out || !out_schema) return BGPU_EINVAL;
  ArrayPriv* p = fill_array(out, b->n_rows, 0, 0);
  p->buffers.push_back(nullptr);
  for (const Column& c : b->cols) {
    ArrowArray* ca = new ArrowArray();
    device_column(c, ca);
    p->children.push_back(ca);
  }
  finish_array(out);
  batch_schema(b, out_schema);
  return BGPU_OK;
} catch (const std::exception& e) {
  return bgpu::abi_guard(ctx, e);  // nothing is thrown across the C ABI
}
