// bgpu_batch_concat: rows of several device batches of one schema, in order, as one batch.
//
// The reference plans CoalesceBatchesExec(target_batch_size = 8192) above every ShuffleReaderExec
// (/root/reference/ballista/scheduler/src/planner.rs:386-387, 409-413); its `concat_batches` is arrow-select's
// `concat` per column. Here one kernel runs a job table: RAW byte ranges (fixed-width values, utf8 bytes) to arbitrary
// byte positions, utf8 OFFSETS rebased onto the running byte total, and BITS (validity, booleans) shifted to the running
// row total (edge bytes shared by neighbouring inputs are OR-ed atomically into the zeroed output).
#include <algorithm>
#include <string>
#include <vector>

#include "internal.h"

namespace bgpu {
namespace {

enum CKind : int32_t { C_RAW = 0, C_OFFSETS = 1, C_BITS = 2, C_ONES = 3 };
struct CJob {
  const uint8_t* src;  // C_RAW: first byte; C_OFFSETS: first offset; C_BITS: bitmap (bit 0 = row 0 of the input); C_ONES: unused
  uint8_t* dst;        // C_RAW: first byte; C_OFFSETS: first output offset; C_BITS / C_ONES: the output bitmap
  uint64_t n;          // C_RAW: bytes; C_OFFSETS: entries; C_BITS / C_ONES: bits
  uint64_t dbit0;      // C_BITS / C_ONES: first output bit; C_OFFSETS: value added after subtracting src[0]
  int32_t kind;
  int32_t via_offsets;  // C_RAW: dbit0 is a device pointer to the input's utf8 offsets; the bytes start at src + offsets[0]
};
constexpr int kThreads = 256;

__device__ __forceinline__ void or_byte(uint8_t* p, uint32_t v) {
  if (!v) return;
  const uint64_t a = (uint64_t)p;
  atomicOr((unsigned int*)(a & ~3ULL), v << (8 * (a & 3)));
}

__global__ void __launch_bounds__(kThreads) concat_kernel(const CJob* __restrict__ jobs) {
  const CJob j = jobs[blockIdx.y];
  const uint64_t tid = (uint64_t)blockIdx.x * kThreads + threadIdx.x, stride = (uint64_t)gridDim.x * kThreads;
  if (j.kind == C_RAW) {
    // destination-aligned 8-byte words fed by funnel-shifted aligned source words; ragged edges by bytes
    const uint8_t* src = j.via_offsets ? j.src + ((const int32_t*)j.dbit0)[0] : j.src;
    const uint64_t d = (uint64_t)j.dst, head = std::min<uint64_t>(j.n, (8 - (d & 7)) & 7);
    const uint64_t words = (j.n - head) >> 3, tail0 = head + (words << 3);
    if (blockIdx.x == 0) {
      if (threadIdx.x < head) j.dst[threadIdx.x] = src[threadIdx.x];
      if (threadIdx.x >= 32 && threadIdx.x - 32 < j.n - tail0) j.dst[tail0 + threadIdx.x - 32] = src[tail0 + threadIdx.x - 32];
    }
    const uint64_t s = (uint64_t)(src + head), sh = (s & 7) * 8;
    const uint64_t* w = (const uint64_t*)(s & ~7ULL);
    uint64_t* o = (uint64_t*)(j.dst + head);
    if (sh == 0) {
      for (uint64_t i = tid; i < words; i += stride) o[i] = w[i];
    } else {
      for (uint64_t i = tid; i < words; i += stride) o[i] = (w[i] >> sh) | (w[i + 1] << (64 - sh));  // w[i + 1] holds bytes of the range
    }
  } else if (j.kind == C_OFFSETS) {
    const int32_t* src = (const int32_t*)j.src;
    const int32_t base = src[0], add = (int32_t)j.dbit0;
    for (uint64_t i = tid; i < j.n; i += stride) ((int32_t*)j.dst)[i] = src[i] - base + add;
  } else {
    // output bytes [b0, b1] hold bits of this job; edge bytes may be shared with the neighbouring inputs
    const uint64_t b0 = j.dbit0 >> 3, b1 = (j.dbit0 + j.n - 1) >> 3, src_bytes = (j.n + 7) >> 3;
    for (uint64_t b = b0 + tid; b <= b1; b += stride) {
      const uint64_t lo = std::max<uint64_t>(b << 3, j.dbit0), hi = std::min<uint64_t>((b << 3) + 8, j.dbit0 + j.n);  // output bits [lo, hi)
      uint32_t v;
      if (j.kind == C_ONES) {
        v = 0xffu;
      } else {
        const uint64_t sbit = lo - j.dbit0, sb = sbit >> 3;
        v = j.src[sb];
        if (sb + 1 < src_bytes) v |= (uint32_t)j.src[sb + 1] << 8;
        v >>= (sbit & 7);
      }
      v &= (1u << (hi - lo)) - 1;
      v <<= (lo & 7);
      if (b == b0 || b == b1) or_byte(j.dst + b, v);
      else j.dst[b] = (uint8_t)v;
    }
  }
}

// bytes of utf8 data each input contributes: out[u * n + k] = offsets[rows] - offsets[0] of utf8 column u of input k
struct SpanArg { const int32_t* offsets; int64_t rows; };
__global__ void utf8_span_kernel(const SpanArg* __restrict__ a, int32_t* __restrict__ out, uint32_t n) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = a[i].rows ? a[i].offsets[a[i].rows] - a[i].offsets[0] : 0;
}

}  // namespace
}  // namespace bgpu

using namespace bgpu;

extern "C" int bgpu_batch_concat(bgpu_ctx* ctx, const bgpu_batch* const* batches, int32_t n, bgpu_batch** out) try {
  if (!ctx || !out || n <= 0 || !batches) return BGPU_EINVAL;
  *out = nullptr;
  if (!ctx->has_device) return ctx->fail(BGPU_ENODEVICE, "context has no device (libbgpu has no CPU fallback)");
  for (int32_t k = 0; k < n; k++)
    if (!batches[k]) return ctx->fail(BGPU_EINVAL, "null batch");
  cudaSetDevice(ctx->device);
  GpuTimeScope timed(ctx, T_KERNEL);
  {  // dictionary columns: every input has its own dictionary, so the VALUES are concatenated and re-encoded (dictionary.cu)
    bool any_dict = false;
    for (int32_t k = 0; k < n; k++) any_dict |= batch_has_dictionary(batches[k]);
    if (any_dict) {
      std::vector<std::unique_ptr<bgpu_batch>> decoded((size_t)n);
      std::vector<const bgpu_batch*> ptrs;
      std::vector<size_t> dict_cols;
      for (int32_t k = 0; k < n; k++) {
        std::vector<size_t> dc;
        if (batches[k]->cols.size() != batches[0]->cols.size()) return ctx->fail(BGPU_EINVAL, "concat: batches differ in column count");
        BGPU_CUDA(ctx, cudaSetDevice(ctx->device));
        BGPU_TRY(dict_decode_batch(ctx, batches[k], &decoded[(size_t)k], &dc));
        if (k == 0) dict_cols = dc;
        else if (dc != dict_cols) return ctx->fail(BGPU_EINVAL, "concat: batches differ in their dictionary columns");
        ptrs.push_back(decoded[(size_t)k].get());
      }
      bgpu_batch* joined = nullptr;
      BGPU_TRY(bgpu_batch_concat(ctx, ptrs.data(), n, &joined));
      std::unique_ptr<bgpu_batch> res(joined);
      for (size_t i : dict_cols) BGPU_TRY(dict_identity_encode(ctx, res.get(), i, batches[0]->cols[i], nullptr, 0));
      *out = res.release();
      return BGPU_OK;
    }
  }
  const std::vector<Column>& proto = batches[0]->cols;
  const size_t n_cols = proto.size();
  int64_t total_rows = 0;
  for (int32_t k = 0; k < n; k++) {
    const bgpu_batch* b = batches[k];
    if (b->cols.size() != n_cols) return ctx->fail(BGPU_EINVAL, "concat: batches differ in column count");
    for (size_t i = 0; i < n_cols; i++) {
      if (b->cols[i].kind == K_DICT32) return ctx->fail(BGPU_EUNSUPPORTED, "concat: dictionary column " + proto[i].name);
      if (b->cols[i].kind != proto[i].kind || b->cols[i].width != proto[i].width || b->cols[i].format != proto[i].format)
        return ctx->fail(BGPU_EINVAL, "concat: batches differ in the type of column " + proto[i].name);
    }
    total_rows += b->n_rows;
  }
  BGPU_CUDA(ctx, cudaSetDevice(ctx->device));

  // utf8 byte totals: one small kernel + one copy (the only host round trip)
  std::vector<size_t> utf8_cols;
  for (size_t i = 0; i < n_cols; i++)
    if (proto[i].kind == K_UTF8 || proto[i].kind == K_BINARY) utf8_cols.push_back(i);
  std::vector<int32_t> spans(utf8_cols.size() * (size_t)n, 0);
  if (!utf8_cols.empty() && total_rows > 0) {
    std::vector<SpanArg> args;
    for (size_t u : utf8_cols)
      for (int32_t k = 0; k < n; k++) args.push_back(SpanArg{batches[k]->cols[u].offsets, batches[k]->n_rows});
    DevAllocPtr d_args, d_out;
    BGPU_TRY(dev_alloc(ctx, args.size() * sizeof(SpanArg), &d_args));
    BGPU_TRY(dev_alloc(ctx, spans.size() * 4, &d_out));
    BGPU_CUDA(ctx, cudaMemcpyAsync(d_args->p, args.data(), args.size() * sizeof(SpanArg), cudaMemcpyHostToDevice, ctx->stream));
    utf8_span_kernel<<<(unsigned)((args.size() + 255) / 256), 256, 0, ctx->stream>>>((const SpanArg*)d_args->p, (int32_t*)d_out->p, (uint32_t)args.size());
    ctx->launches++;
    BGPU_CUDA(ctx, cudaMemcpyAsync(spans.data(), d_out->p, spans.size() * 4, cudaMemcpyDeviceToHost, ctx->stream));
    BGPU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  }

  bgpu_batch* res = new bgpu_batch(ctx);
  res->n_rows = total_rows;
  res->cols = proto;
  std::vector<CJob> jobs;
  auto fail = [&](int rc) { delete res; return rc; };
  size_t u = 0;
  for (size_t i = 0; i < n_cols; i++) {
    Column& c = res->cols[i];
    c.owned.clear();
    c.length = total_rows;
    c.validity = nullptr;
    c.offsets = nullptr;
    c.values = nullptr;
    bool any_validity = false, counts_known = true;
    int64_t nulls = 0;
    for (int32_t k = 0; k < n; k++) {
      const Column& s = batches[k]->cols[i];
      if (s.validity && batches[k]->n_rows) {
        any_validity = true;
        if (s.null_count < 0) counts_known = false;
        else nulls += s.null_count;
      }
    }
    c.null_count = any_validity ? (counts_known ? nulls : -1) : 0;
    const size_t bitmap_bytes = ((size_t)((total_rows + 7) / 8) + 3) & ~(size_t)3;  // whole 32-bit words for the edge atomics
    if (any_validity) {
      DevAllocPtr v;
      int rc = dev_alloc(ctx, bitmap_bytes, &v);
      if (rc) return fail(rc);
      cudaMemsetAsync(v->p, 0, bitmap_bytes, ctx->stream);
      c.owned.push_back(v);
      c.validity = (uint8_t*)v->p;
      int64_t row = 0;
      for (int32_t k = 0; k < n; k++) {
        const Column& s = batches[k]->cols[i];
        const int64_t rows = batches[k]->n_rows;
        if (rows) jobs.push_back(CJob{s.validity, c.validity, (uint64_t)rows, (uint64_t)row, s.validity ? C_BITS : C_ONES, 0});
        row += rows;
      }
    }
    if (c.kind == K_UTF8 || c.kind == K_BINARY) {
      int64_t bytes = 0;
      for (int32_t k = 0; k < n; k++) bytes += spans[u * (size_t)n + k];
      if (bytes > INT32_MAX) return fail(ctx->fail(BGPU_EOVERFLOW, "concat: utf8 column " + c.name + " exceeds 2 GiB of data"));
      DevAllocPtr o, d;
      int rc = dev_alloc(ctx, (size_t)(total_rows + 1) * 4, &o);
      if (!rc) rc = dev_alloc(ctx, (size_t)bytes + 16, &d);
      if (rc) return fail(rc);
      c.owned.push_back(o);
      c.owned.push_back(d);
      c.offsets = (int32_t*)o->p;
      c.values = d->p;
      c.data_bytes = bytes;
      int64_t row = 0, at = 0;
      for (int32_t k = 0; k < n; k++) {
        const Column& s = batches[k]->cols[i];
        const int64_t rows = batches[k]->n_rows, span = spans[u * (size_t)n + k];
        if (rows) {
          jobs.push_back(CJob{(const uint8_t*)s.offsets, (uint8_t*)(c.offsets + row), (uint64_t)rows, (uint64_t)at, C_OFFSETS, 0});
          // the input's bytes start at values[offsets[0]], a device-resident number: the kernel adds it
          if (span) jobs.push_back(CJob{(const uint8_t*)s.values, (uint8_t*)c.values + at, (uint64_t)span, (uint64_t)(uintptr_t)s.offsets, C_RAW, 1});
        }
        row += rows;
        at += span;
      }
      // the final offset = total bytes
      cudaMemcpyAsync(c.offsets + total_rows, &c.data_bytes, 4, cudaMemcpyHostToDevice, ctx->stream);
      u++;
    } else if (c.kind == K_BOOL) {
      DevAllocPtr v;
      int rc = dev_alloc(ctx, bitmap_bytes, &v);
      if (rc) return fail(rc);
      cudaMemsetAsync(v->p, 0, bitmap_bytes, ctx->stream);
      c.owned.push_back(v);
      c.values = v->p;
      int64_t row = 0;
      for (int32_t k = 0; k < n; k++) {
        const int64_t rows = batches[k]->n_rows;
        if (rows) jobs.push_back(CJob{(const uint8_t*)batches[k]->cols[i].values, (uint8_t*)c.values, (uint64_t)rows, (uint64_t)row, C_BITS, 0});
        row += rows;
      }
    } else {
      DevAllocPtr v;
      int rc = dev_alloc(ctx, (size_t)total_rows * (size_t)c.width, &v);
      if (rc) return fail(rc);
      c.owned.push_back(v);
      c.values = v->p;
      int64_t row = 0;
      for (int32_t k = 0; k < n; k++) {
        const int64_t rows = batches[k]->n_rows;
        if (rows) jobs.push_back(CJob{(const uint8_t*)batches[k]->cols[i].values, (uint8_t*)c.values + (size_t)row * c.width,
                                      (uint64_t)rows * (uint64_t)c.width, 0, C_RAW, 0});
        row += rows;
      }
    }
  }
  if (!jobs.empty()) {
    DevAllocPtr d_jobs;
    int rc = dev_alloc(ctx, jobs.size() * sizeof(CJob), &d_jobs);
    if (rc) return fail(rc);
    cudaMemcpyAsync(d_jobs->p, jobs.data(), jobs.size() * sizeof(CJob), cudaMemcpyHostToDevice, ctx->stream);
    uint64_t biggest = 0;
    for (const CJob& j : jobs) biggest = std::max<uint64_t>(biggest, j.kind == C_RAW ? j.n : j.kind == C_OFFSETS ? j.n * 4 : j.n / 8);
    const uint32_t gx = (uint32_t)std::min<uint64_t>(std::max<uint64_t>(1, biggest / (64 * 1024)),
                                                     std::max<uint64_t>(1, (uint64_t)ctx->sm_count * 8 / jobs.size() + 1));
    for (size_t j0 = 0; j0 < jobs.size(); j0 += 65535) {
      const uint32_t gy = (uint32_t)std::min<size_t>(65535, jobs.size() - j0);
      concat_kernel<<<dim3(gx, gy), kThreads, 0, ctx->stream>>>((const CJob*)d_jobs->p + j0);
      ctx->launches++;
    }
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);  // the job table and &data_bytes are host temporaries
    if (e != cudaSuccess) return fail(ctx->fail(BGPU_ECUDA, std::string("concat: ") + cudaGetErrorString(e)));
  }
  *out = res;
  return BGPU_OK;
} catch (const std::exception& e) {
  return bgpu::abi_guard(ctx, e);  // nothing is thrown across the C ABI
}
