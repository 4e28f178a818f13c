// Runtime specialisation of the aggregate kernels: the hand-written kernel templates of agg_jit_prelude.cuh are
// instantiated over a plan-specific `Plan` struct (emitted by agg_codegen.cu) and compiled for sm_100a with NVRTC.
// A query engine cannot instantiate every (key layout x expression x nullability) combination ahead of time; an
// interpreter over the plan costs ~1000 instructions per row and thrashes the instruction cache (profiles/r1).
//
// libnvrtc and libcuda are dlopen'ed on first use, so the library still loads (and fails loudly per call) on a box
// without a driver. Compiled modules are cached per (device, source) for the life of the process.
#pragma once
#include <cuda.h>

#include <string>
#include <vector>

#include "internal.h"

namespace bgpu {

struct JitModule {
  CUmodule module = nullptr;
  std::vector<CUfunction> fn;  // in the order of the `names` passed to jit_compile
};

// Compile `prelude + source` (cached per device and `source`) and resolve the extern "C" kernels `names`. The
// returned pointer stays valid for the life of the process.
int jit_compile(bgpu_ctx* ctx, const char* prelude, const std::string& source, const std::vector<std::string>& names,
                const JitModule** out);
// {NVRTC compiles, cubins loaded from the on-disk cache, in-memory module hits, cubins stored on disk} of this process
void jit_stats(uint64_t* out4);
// NVRTC only (no driver, no device): source -> sm_100a cubin (through the on-disk cache, see jit.cu)
int jit_compile_cubin(bgpu_ctx* ctx, const std::string& source, std::string* cubin);

// cuLaunchKernel on ctx->stream with a single by-value argument struct.
int jit_launch(bgpu_ctx* ctx, CUfunction fn, int grid, int block, size_t smem, const void* arg_struct);
int jit_set_max_smem(bgpu_ctx* ctx, CUfunction fn, size_t smem);
// bytes of local memory per thread (register spills / stack) of a compiled kernel
int jit_local_bytes(bgpu_ctx* ctx, CUfunction fn, int* out);
int jit_occupancy(bgpu_ctx* ctx, CUfunction fn, int block, size_t smem, int* out);

}  // namespace bgpu
