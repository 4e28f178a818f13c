// Context: one CUDA stream + stream-ordered pool allocations per task (SURVEY.md §8b "Threading").
#include "internal.h"

extern "C" const char* bgpu_version(void) { return "libbgpu 0.1 (sm_100a)"; }

static int ctx_create(int device, cudaStream_t external, bool has_external, bgpu_ctx** out) {
  if (!out) return BGPU_EINVAL;
  *out = nullptr;
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  bgpu_ctx* ctx = new bgpu_ctx();
  *out = ctx;  // returned even on failure so the caller can read the error, must still be destroyed
  if (e != cudaSuccess || n == 0)
    return ctx->fail(BGPU_ENODEVICE, std::string("no CUDA device: ") + cudaGetErrorString(e) +
                                         " (libbgpu has no CPU fallback)");
  if (device < 0 || device >= n) return ctx->fail(BGPU_EINVAL, "device index out of range");
  ctx->device = device;
  BGPU_CUDA(ctx, cudaSetDevice(device));
  cudaDeviceProp prop;
  BGPU_CUDA(ctx, cudaGetDeviceProperties(&prop, device));
  if (prop.major < 10)
    return ctx->fail(BGPU_ENODEVICE, std::string("device ") + prop.name + " is sm_" + std::to_string(prop.major) +
                                         std::to_string(prop.minor) + "; libbgpu is built for sm_100a only");
  ctx->sm_count = prop.multiProcessorCount;
  ctx->smem_optin = prop.sharedMemPerBlockOptin;
  if (has_external) {
    ctx->stream = external;
    ctx->owns_stream = false;
  } else {
    BGPU_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
  }
  // keep freed blocks in the pool: operators allocate/free every batch
  cudaMemPool_t pool;
  BGPU_CUDA(ctx, cudaDeviceGetDefaultMemPool(&pool, device));
  uint64_t thresh = UINT64_MAX;
  BGPU_CUDA(ctx, cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &thresh));
  ctx->has_device = true;
  ctx->pinned_bytes = 1 << 20;
  BGPU_CUDA(ctx, cudaMallocHost(&ctx->pinned, ctx->pinned_bytes));
  return BGPU_OK;
}

extern "C" int bgpu_ctx_create(int device, bgpu_ctx** out) { return ctx_create(device, nullptr, false, out); }
extern "C" int bgpu_ctx_create_on_stream(int device, void* stream, bgpu_ctx** out) {
  return ctx_create(device, (cudaStream_t)stream, true, out);
}

void ctx_retain(bgpu_ctx* ctx) { ctx->refs.fetch_add(1, std::memory_order_relaxed); }

static void ctx_teardown(bgpu_ctx* ctx) {
  if (ctx->has_device) {
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    if (ctx->owns_stream) cudaStreamDestroy(ctx->stream);
  }
  if (ctx->pinned) cudaFreeHost(ctx->pinned);
  for (auto& b : ctx->pinned_pool) cudaFreeHost(b.first);
  for (auto& sp : ctx->times_open) { cudaEventDestroy(sp.a); cudaEventDestroy(sp.b); }
  for (cudaEvent_t e : ctx->times_pool) cudaEventDestroy(e);
  delete ctx;
}

// ---- h2d / kernel / d2h accounting -------------------------------------------------------------------------------
static cudaEvent_t times_event(bgpu_ctx* ctx) {
  if (!ctx->times_pool.empty()) {
    cudaEvent_t e = ctx->times_pool.back();
    ctx->times_pool.pop_back();
    return e;
  }
  cudaEvent_t e = nullptr;
  if (cudaEventCreate(&e) != cudaSuccess) return nullptr;
  return e;
}
GpuTimeScope::GpuTimeScope(bgpu_ctx* c, int cls_, cudaStream_t st) : ctx(c), cls(cls_), stream(st ? st : (c ? c->stream : nullptr)) {
  if (!ctx || !ctx->times_on || !ctx->has_device) { ctx = nullptr; return; }
  if (stream == ctx->stream && ctx->times_depth++ > 0) return;  // an enclosing scope accounts for this work
  a = times_event(ctx);
  if (a) cudaEventRecord(a, stream);
}
GpuTimeScope::~GpuTimeScope() {
  if (!ctx) return;
  if (stream == ctx->stream) ctx->times_depth--;
  if (!a) return;
  cudaEvent_t b = times_event(ctx);
  if (!b) { ctx->times_pool.push_back(a); return; }
  cudaEventRecord(b, stream);
  ctx->times_open.push_back(bgpu_ctx::TimedSpan{a, b, cls});
}

extern "C" int bgpu_ctx_enable_times(bgpu_ctx* ctx, int32_t on) {
  if (!ctx) return BGPU_EINVAL;
  ctx->times_on = on != 0;
  return BGPU_OK;
}

extern "C" int bgpu_ctx_times(bgpu_ctx* ctx, double* out3_ms, int32_t reset) {
  if (!ctx || !out3_ms) return BGPU_EINVAL;
  if (ctx->has_device) {
    BGPU_CUDA(ctx, cudaSetDevice(ctx->device));
    for (auto& sp : ctx->times_open) {
      float ms = 0.f;
      BGPU_CUDA(ctx, cudaEventSynchronize(sp.b));
      if (cudaEventElapsedTime(&ms, sp.a, sp.b) == cudaSuccess) ctx->times_ms[sp.cls] += ms;
      ctx->times_pool.push_back(sp.a);
      ctx->times_pool.push_back(sp.b);
    }
    ctx->times_open.clear();
  }
  for (int i = 0; i < 3; i++) out3_ms[i] = ctx->times_ms[i];
  if (reset)
    for (int i = 0; i < 3; i++) ctx->times_ms[i] = 0.0;
  return BGPU_OK;
}

int ctx_pinned_get(bgpu_ctx* ctx, size_t bytes, void** out) {
  {
    std::lock_guard<std::mutex> lk(ctx->pinned_pool_mu);
    for (size_t i = 0; i < ctx->pinned_pool.size(); i++)
      if (ctx->pinned_pool[i].second >= bytes && ctx->pinned_pool[i].second <= bytes * 2) {
        *out = ctx->pinned_pool[i].first;
        ctx->pinned_pool.erase(ctx->pinned_pool.begin() + (long)i);
        return BGPU_OK;
      }
  }
  BGPU_CUDA(ctx, cudaHostAlloc(out, bytes, cudaHostAllocDefault));
  return BGPU_OK;
}
void ctx_pinned_put(bgpu_ctx* ctx, void* p, size_t bytes) {
  std::lock_guard<std::mutex> lk(ctx->pinned_pool_mu);
  ctx->pinned_pool.emplace_back(p, bytes);
}

void ctx_release(bgpu_ctx* ctx) {
  if (ctx->refs.fetch_sub(1, std::memory_order_acq_rel) == 1) ctx_teardown(ctx);
}

extern "C" void bgpu_ctx_destroy(bgpu_ctx* ctx) {
  if (!ctx) return;
  if (ctx->has_device) {
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
  }
  ctx_release(ctx);
}

extern "C" const char* bgpu_last_error(const bgpu_ctx* ctx) { return ctx ? ctx->err.c_str() : "null context"; }

extern "C" int bgpu_ctx_synchronize(bgpu_ctx* ctx) {
  if (!ctx || !ctx->has_device) return BGPU_EINVAL;
  BGPU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return BGPU_OK;
}

extern "C" void* bgpu_ctx_stream(bgpu_ctx* ctx) { return ctx ? (void*)ctx->stream : nullptr; }

extern "C" uint64_t bgpu_ctx_kernel_launches(const bgpu_ctx* ctx) { return ctx ? ctx->launches : 0; }

int dev_alloc(bgpu_ctx* ctx, size_t bytes, DevAllocPtr* out) {
  if (!ctx->has_device) return ctx->fail(BGPU_ENODEVICE, "context has no device");
  auto a = std::make_shared<DevAlloc>();
  a->bytes = bytes < 16 ? 16 : bytes;
  BGPU_CUDA(ctx, cudaMallocAsync(&a->p, a->bytes, ctx->stream));
  a->ctx = ctx;
  ctx_retain(ctx);
  *out = a;
  return BGPU_OK;
}

int bgpu::abi_guard(bgpu_ctx* ctx, const std::exception& e) {
  const bool oom = dynamic_cast<const std::bad_alloc*>(&e) != nullptr;
  if (!ctx) return BGPU_EINVAL;
  return ctx->fail(oom ? BGPU_ECUDA : BGPU_EINVAL, std::string(oom ? "out of host memory: " : "internal error: ") + e.what());
}
