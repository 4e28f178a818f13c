// libnccl.so.2 loaded at run time (the copy torch already loaded is reused): libbgpu has no link-time NCCL dependency.
#pragma once
#include <dlfcn.h>

#include <mutex>
#include <string>

#include "internal.h"

typedef struct { char internal[128]; } NcclUniqueId;
constexpr int kNcclUint8 = 1;
struct Nccl {
  bool ok = false;
  std::string why;
  int (*GetUniqueId)(NcclUniqueId*) = nullptr;
  int (*CommInitRank)(void**, int, NcclUniqueId, int) = nullptr;
  int (*CommDestroy)(void*) = nullptr;
  int (*Send)(const void*, size_t, int, int, void*, cudaStream_t) = nullptr;
  int (*Recv)(void*, size_t, int, int, void*, cudaStream_t) = nullptr;
  int (*AllGather)(const void*, void*, size_t, int, void*, cudaStream_t) = nullptr;
  int (*GroupStart)() = nullptr;
  int (*GroupEnd)() = nullptr;
  const char* (*GetErrorString)(int) = nullptr;
};

inline const Nccl& nccl() {
  static Nccl n;
  static std::once_flag once;
  std::call_once(once, [] {
    void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);  // the copy the host process already loaded
    if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_LOCAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_LOCAL);
    if (!h) { n.why = std::string("cannot load libnccl: ") + dlerror(); return; }
    bool ok = true;
    auto sym = [&](const char* s) { void* p = dlsym(h, s); if (!p) { ok = false; n.why = std::string("missing symbol ") + s; } return p; };
    *(void**)&n.GetUniqueId = sym("ncclGetUniqueId");
    *(void**)&n.CommInitRank = sym("ncclCommInitRank");
    *(void**)&n.CommDestroy = sym("ncclCommDestroy");
    *(void**)&n.Send = sym("ncclSend");
    *(void**)&n.Recv = sym("ncclRecv");
    *(void**)&n.AllGather = sym("ncclAllGather");
    *(void**)&n.GroupStart = sym("ncclGroupStart");
    *(void**)&n.GroupEnd = sym("ncclGroupEnd");
    *(void**)&n.GetErrorString = sym("ncclGetErrorString");
    n.ok = ok;
  });
  return n;
}

#define BGPU_NCCL(ctx, call)                                                                                  \
  do {                                                                                                        \
    const int _r = (call);                                                                                    \
    if (_r != 0) return (ctx)->fail(BGPU_ENCCL, std::string(#call) + ": " + nccl().GetErrorString(_r));       \
  } while (0)

