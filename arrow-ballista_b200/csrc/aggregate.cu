// K4/K5 grouped hash aggregate (filled in next milestone).
#include "internal.h"
extern "C" int bgpu_agg_create(bgpu_ctx* ctx, int32_t, const struct ArrowSchema*, const int32_t*, int32_t,
                               const int32_t*, const int32_t*, int32_t, bgpu_agg** out) {
  if (out) *out = nullptr;
  return ctx ? ctx->fail(BGPU_EUNSUPPORTED, "aggregate not built yet") : BGPU_EINVAL;
}
extern "C" int bgpu_agg_update(bgpu_agg*, const bgpu_batch*) { return BGPU_EUNSUPPORTED; }
extern "C" int bgpu_agg_emit(bgpu_agg*, bgpu_batch** out) { if (out) *out = nullptr; return BGPU_EUNSUPPORTED; }
extern "C" int64_t bgpu_agg_num_groups(bgpu_agg*) { return -1; }
extern "C" void bgpu_agg_destroy(bgpu_agg*) {}
