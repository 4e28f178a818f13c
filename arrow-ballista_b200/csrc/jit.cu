// NVRTC + driver-API plumbing for the runtime-specialised aggregate kernels (see jit.h).
#include "jit.h"

#include <dlfcn.h>
#include <errno.h>
#include <nvrtc.h>
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/stat.h>
#include <unistd.h>

#include <atomic>
#include <map>
#include <mutex>

namespace bgpu {
namespace {

struct Api {
  bool ok = false;
  std::string why;
  // nvrtc
  nvrtcResult (*CreateProgram)(nvrtcProgram*, const char*, const char*, int, const char* const*, const char* const*) = nullptr;
  nvrtcResult (*CompileProgram)(nvrtcProgram, int, const char* const*) = nullptr;
  nvrtcResult (*GetCUBINSize)(nvrtcProgram, size_t*) = nullptr;
  nvrtcResult (*GetCUBIN)(nvrtcProgram, char*) = nullptr;
  nvrtcResult (*GetProgramLogSize)(nvrtcProgram, size_t*) = nullptr;
  nvrtcResult (*GetProgramLog)(nvrtcProgram, char*) = nullptr;
  nvrtcResult (*DestroyProgram)(nvrtcProgram*) = nullptr;
  nvrtcResult (*Version)(int*, int*) = nullptr;
  // driver
  CUresult (*ModuleLoadData)(CUmodule*, const void*) = nullptr;
  CUresult (*ModuleGetFunction)(CUfunction*, CUmodule, const char*) = nullptr;
  CUresult (*LaunchKernel)(CUfunction, unsigned, unsigned, unsigned, unsigned, unsigned, unsigned, unsigned, CUstream,
                           void**, void**) = nullptr;
  CUresult (*FuncSetAttribute)(CUfunction, CUfunction_attribute, int) = nullptr;
  CUresult (*FuncGetAttribute)(int*, CUfunction_attribute, CUfunction) = nullptr;
  CUresult (*Occupancy)(int*, CUfunction, int, size_t) = nullptr;
  CUresult (*GetErrorString)(CUresult, const char**) = nullptr;
};

void* open_first(const std::vector<const char*>& names) {
  for (const char* n : names) {
    void* h = dlopen(n, RTLD_NOW | RTLD_LOCAL);
    if (h) return h;
  }
  return nullptr;
}

const Api& api() {
  static Api a;
  static std::once_flag once;
  std::call_once(once, [] {
    void* rtc = open_first({"libnvrtc.so.12", "libnvrtc.so", "/usr/local/cuda/lib64/libnvrtc.so.12",
                            "/usr/local/cuda/lib64/libnvrtc.so"});
    if (!rtc) { a.why = std::string("cannot load libnvrtc: ") + dlerror(); return; }
    void* drv = open_first({"libcuda.so.1", "libcuda.so"});
    if (!drv) { a.why = std::string("cannot load libcuda: ") + dlerror(); return; }
    bool ok = true;
    auto sym = [&](void* h, const char* n) { void* p = dlsym(h, n); if (!p) { ok = false; a.why = std::string("missing symbol ") + n; } return p; };
    *(void**)&a.CreateProgram = sym(rtc, "nvrtcCreateProgram");
    *(void**)&a.CompileProgram = sym(rtc, "nvrtcCompileProgram");
    *(void**)&a.GetCUBINSize = sym(rtc, "nvrtcGetCUBINSize");
    *(void**)&a.GetCUBIN = sym(rtc, "nvrtcGetCUBIN");
    *(void**)&a.GetProgramLogSize = sym(rtc, "nvrtcGetProgramLogSize");
    *(void**)&a.GetProgramLog = sym(rtc, "nvrtcGetProgramLog");
    *(void**)&a.DestroyProgram = sym(rtc, "nvrtcDestroyProgram");
    *(void**)&a.ModuleLoadData = sym(drv, "cuModuleLoadData");
    *(void**)&a.ModuleGetFunction = sym(drv, "cuModuleGetFunction");
    *(void**)&a.LaunchKernel = sym(drv, "cuLaunchKernel");
    *(void**)&a.FuncSetAttribute = sym(drv, "cuFuncSetAttribute");
    *(void**)&a.FuncGetAttribute = sym(drv, "cuFuncGetAttribute");
    *(void**)&a.Occupancy = sym(drv, "cuOccupancyMaxActiveBlocksPerMultiprocessor");
    *(void**)&a.GetErrorString = sym(drv, "cuGetErrorString");
    a.ok = ok;
  });
  return a;
}

std::string cu_err(const Api& a, CUresult r) {
  const char* s = nullptr;
  if (a.GetErrorString) a.GetErrorString(r, &s);
  return s ? s : ("CUresult " + std::to_string((int)r));
}

std::mutex g_mu;
std::map<std::pair<int, std::string>, JitModule*> g_cache;  // (device, source) -> module; never freed
std::atomic<uint64_t> g_stats[4];  // NVRTC compiles, cubins loaded from the disk cache, in-memory module hits, cubins stored

// ---- on-disk cache of compiled cubins -----------------------------------------------------------------------------------
// NVRTC takes ~2 s per plan: an executor that sees a query once must not pay that in every process. Files live under
// $BGPU_CACHE_DIR (default: $XDG_CACHE_HOME/bgpu, ~/.cache/bgpu, /tmp/bgpu-cache-<uid>; BGPU_CACHE_DIR=off disables),
// named by a 128-bit hash of (format tag incl. arch and NVRTC version, source); the file repeats tag and source in full
// and a load compares them byte for byte, so a hash collision or a stale file can only cost a recompile.
uint64_t fnv1a(const std::string& s, uint64_t h) {
  for (unsigned char c : s) { h ^= c; h *= 1099511628211ULL; }
  return h;
}
std::string disk_cache_dir() {
  const char* e = getenv("BGPU_CACHE_DIR");
  if (e && (!strcmp(e, "off") || !*e)) return "";
  std::string d;
  if (e) d = e;
  else if (const char* x = getenv("XDG_CACHE_HOME")) d = std::string(x) + "/bgpu";
  else if (const char* h = getenv("HOME")) d = std::string(h) + "/.cache/bgpu";
  else d = "/tmp/bgpu-cache-" + std::to_string((unsigned)getuid());
  for (size_t i = 1; i <= d.size(); i++)  // mkdir -p
    if (i == d.size() || d[i] == '/') {
      const std::string part = d.substr(0, i);
      if (mkdir(part.c_str(), 0700) != 0 && errno != EEXIST) return "";
    }
  return d;
}
std::string disk_cache_path(const std::string& tag, const std::string& source) {
  static const std::string dir = disk_cache_dir();
  if (dir.empty()) return "";
  char name[64];
  snprintf(name, sizeof name, "/%016llx%016llx.cubin", (unsigned long long)fnv1a(source, fnv1a(tag, 14695981039346656037ULL)),
           (unsigned long long)fnv1a(source, fnv1a(tag, 0x9e3779b97f4a7c15ULL)));
  return dir + name;
}
bool disk_cache_load(const std::string& path, const std::string& tag, const std::string& source, std::string* cubin) {
  if (path.empty()) return false;
  FILE* f = fopen(path.c_str(), "rb");
  if (!f) return false;
  bool ok = false;
  uint64_t lens[3];
  if (fread(lens, 8, 3, f) == 3 && lens[0] == tag.size() && lens[1] == source.size() && lens[2] > 0 && lens[2] < (1ULL << 30)) {
    std::string t(lens[0], '\0'), s(lens[1], '\0');
    cubin->assign(lens[2], '\0');
    ok = fread(&t[0], 1, t.size(), f) == t.size() && fread(&s[0], 1, s.size(), f) == s.size() &&
         fread(&(*cubin)[0], 1, cubin->size(), f) == cubin->size() && t == tag && s == source;
  }
  fclose(f);
  return ok;
}
bool disk_cache_store(const std::string& path, const std::string& tag, const std::string& source, const std::string& cubin) {
  if (path.empty()) return false;
  const std::string tmp = path + ".tmp" + std::to_string((long)getpid()) + "_" + std::to_string((unsigned long)pthread_self());
  FILE* f = fopen(tmp.c_str(), "wb");
  if (!f) return false;
  const uint64_t lens[3] = {tag.size(), source.size(), cubin.size()};
  bool ok = fwrite(lens, 8, 3, f) == 3 && fwrite(tag.data(), 1, tag.size(), f) == tag.size() &&
            fwrite(source.data(), 1, source.size(), f) == source.size() && fwrite(cubin.data(), 1, cubin.size(), f) == cubin.size();
  ok = fclose(f) == 0 && ok;
  if (ok) ok = rename(tmp.c_str(), path.c_str()) == 0;  // atomic: readers see the whole file or none
  if (!ok) remove(tmp.c_str());
  return ok;
}

}  // namespace

void jit_stats(uint64_t* out4) {
  for (int i = 0; i < 4; i++) out4[i] = g_stats[i].load();
}

// Host-side compile only (no driver needed): used by the CPU test-suite to check that every plan shape the tests
// use generates source NVRTC accepts.
int jit_compile_cubin(bgpu_ctx* ctx, const std::string& source, std::string* cubin) {
  static Api rtc_only;
  static std::once_flag once;
  std::call_once(once, [] {
    void* rtc = open_first({"libnvrtc.so.12", "libnvrtc.so", "/usr/local/cuda/lib64/libnvrtc.so.12",
                            "/usr/local/cuda/lib64/libnvrtc.so"});
    if (!rtc) { rtc_only.why = std::string("cannot load libnvrtc: ") + dlerror(); return; }
    bool ok = true;
    auto sym = [&](const char* n) { void* p = dlsym(rtc, n); if (!p) ok = false; return p; };
    *(void**)&rtc_only.CreateProgram = sym("nvrtcCreateProgram");
    *(void**)&rtc_only.CompileProgram = sym("nvrtcCompileProgram");
    *(void**)&rtc_only.GetCUBINSize = sym("nvrtcGetCUBINSize");
    *(void**)&rtc_only.GetCUBIN = sym("nvrtcGetCUBIN");
    *(void**)&rtc_only.GetProgramLogSize = sym("nvrtcGetProgramLogSize");
    *(void**)&rtc_only.GetProgramLog = sym("nvrtcGetProgramLog");
    *(void**)&rtc_only.DestroyProgram = sym("nvrtcDestroyProgram");
    *(void**)&rtc_only.Version = sym("nvrtcVersion");
    rtc_only.ok = ok;
  });
  const Api& a = rtc_only;
  if (!a.ok) return ctx->fail(BGPU_ECUDA, "NVRTC unavailable: " + a.why);
  int vmaj = 0, vmin = 0;
  a.Version(&vmaj, &vmin);
  const std::string tag = "bgpu-cubin-1 sm_100a nvrtc " + std::to_string(vmaj) + "." + std::to_string(vmin);
  const std::string path = disk_cache_path(tag, source);
  if (disk_cache_load(path, tag, source, cubin)) {
    g_stats[1]++;
    return BGPU_OK;
  }
  nvrtcProgram prog;
  if (a.CreateProgram(&prog, source.c_str(), "bgpu_agg.cu", 0, nullptr, nullptr) != NVRTC_SUCCESS)
    return ctx->fail(BGPU_ECUDA, "nvrtcCreateProgram failed");
  const char* opts[] = {"--gpu-architecture=sm_100a", "-std=c++17", "-lineinfo", "-default-device"};
  const nvrtcResult rc = a.CompileProgram(prog, 4, opts);
  if (rc != NVRTC_SUCCESS) {
    size_t n = 0;
    a.GetProgramLogSize(prog, &n);
    std::string log(n, '\0');
    if (n) a.GetProgramLog(prog, &log[0]);
    a.DestroyProgram(&prog);
    if (log.size() > 4000) log.resize(4000);
    return ctx->fail(BGPU_ECUDA, "NVRTC compile failed:\n" + log);
  }
  size_t n = 0;
  a.GetCUBINSize(prog, &n);
  cubin->assign(n, '\0');
  a.GetCUBIN(prog, &(*cubin)[0]);
  a.DestroyProgram(&prog);
  g_stats[0]++;
  if (disk_cache_store(path, tag, source, *cubin)) g_stats[3]++;
  return BGPU_OK;
}

int jit_compile(bgpu_ctx* ctx, const char* prelude, const std::string& source, const std::vector<std::string>& names,
                const JitModule** out) {
  const Api& a = api();
  if (!a.ok) return ctx->fail(BGPU_ECUDA, "JIT unavailable: " + a.why);
  std::lock_guard<std::mutex> lk(g_mu);
  auto key = std::make_pair(ctx->device, source);
  auto it = g_cache.find(key);
  if (it != g_cache.end()) { g_stats[2]++; *out = it->second; return BGPU_OK; }
  std::string cubin;
  BGPU_TRY(jit_compile_cubin(ctx, std::string(prelude) + source, &cubin));
  BGPU_CUDA(ctx, cudaSetDevice(ctx->device));
  BGPU_CUDA(ctx, cudaFree(0));  // make sure the primary context is current for the driver calls
  auto* m = new JitModule();
  CUresult r = a.ModuleLoadData(&m->module, cubin.data());
  if (r != CUDA_SUCCESS) { delete m; return ctx->fail(BGPU_ECUDA, "cuModuleLoadData: " + cu_err(a, r)); }
  for (const std::string& n : names) {
    CUfunction f;
    r = a.ModuleGetFunction(&f, m->module, n.c_str());
    if (r != CUDA_SUCCESS) { delete m; return ctx->fail(BGPU_ECUDA, "cuModuleGetFunction(" + n + "): " + cu_err(a, r)); }
    m->fn.push_back(f);
  }
  g_cache[key] = m;
  *out = m;
  return BGPU_OK;
}

int jit_launch(bgpu_ctx* ctx, CUfunction fn, int grid, int block, size_t smem, const void* arg_struct) {
  const Api& a = api();
  void* params[] = {const_cast<void*>(arg_struct)};
  const CUresult r = a.LaunchKernel(fn, (unsigned)grid, 1, 1, (unsigned)block, 1, 1, (unsigned)smem, (CUstream)ctx->stream,
                                    params, nullptr);
  if (r != CUDA_SUCCESS) return ctx->fail(BGPU_ECUDA, "cuLaunchKernel: " + cu_err(a, r));
  ctx->launches++;
  return BGPU_OK;
}

// The limit is only ever raised: a CUfunction is shared by every task (thread) whose plan has the same shape, and their
// launches want different amounts (capacity by demand) - lowering it for one task between another task's set and launch
// would fail that launch. The attribute is a cap, not a reservation, so a larger value costs nothing.
int jit_set_max_smem(bgpu_ctx* ctx, CUfunction fn, size_t smem) {
  static std::mutex mu;
  static std::map<CUfunction, size_t> current;
  const Api& a = api();
  std::lock_guard<std::mutex> lk(mu);
  auto it = current.find(fn);
  if (it != current.end() && it->second >= smem) return BGPU_OK;
  const CUresult r = a.FuncSetAttribute(fn, CU_FUNC_ATTRIBUTE_MAX_DYNAMIC_SHARED_SIZE_BYTES, (int)smem);
  if (r != CUDA_SUCCESS) return ctx->fail(BGPU_ECUDA, "cuFuncSetAttribute(max dynamic smem): " + cu_err(a, r));
  current[fn] = smem;
  return BGPU_OK;
}

int jit_local_bytes(bgpu_ctx* ctx, CUfunction fn, int* out) {
  const Api& a = api();
  const CUresult r = a.FuncGetAttribute(out, CU_FUNC_ATTRIBUTE_LOCAL_SIZE_BYTES, fn);
  if (r != CUDA_SUCCESS) return ctx->fail(BGPU_ECUDA, "cuFuncGetAttribute(local size): " + cu_err(a, r));
  return BGPU_OK;
}

int jit_occupancy(bgpu_ctx* ctx, CUfunction fn, int block, size_t smem, int* out) {
  const Api& a = api();
  const CUresult r = a.Occupancy(out, fn, block, smem);
  if (r != CUDA_SUCCESS) return ctx->fail(BGPU_ECUDA, "cuOccupancyMaxActiveBlocksPerMultiprocessor: " + cu_err(a, r));
  return BGPU_OK;
}

}  // namespace bgpu
