// K6 intra-node all-to-all over NCCL (filled in a later milestone).
#include "internal.h"
extern "C" int bgpu_all_to_all(bgpu_ctx* ctx, void*, int32_t, int32_t, const bgpu_batch*, const uint64_t*, uint32_t,
                               bgpu_batch** out) {
  if (out) *out = nullptr;
  return ctx ? ctx->fail(BGPU_EUNSUPPORTED, "exchange not built yet") : BGPU_EINVAL;
}
