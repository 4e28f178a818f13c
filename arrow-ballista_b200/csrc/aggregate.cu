// K4/K5: grouped hash aggregate, Partial and Final modes — host side (plan building, create/destroy).
// Update and emit live in aggregate_run.cu; kernels in agg_kernels.cuh.
//
// Replaces DataFusion's AggregateExec -> GroupedHashAggregateStream for the aggregates on Ballista's hot path
// (plan shape /root/reference/ballista/scheduler/src/planner.rs:376-427; proto
// /root/reference/ballista/core/proto/datafusion.proto:1405-1411,1437-1450; restated SURVEY.md A.3).
#include <string.h>

#include <algorithm>

#include "agg_host.h"
#include "interner.h"
#include "jit.h"

using namespace bgpu;

int agg_launch_grid(bgpu_ctx* ctx, int64_t n, int per_block) {
  return (int)std::max<int64_t>(1, std::min<int64_t>(ceil_div64(n, per_block), (int64_t)ctx->sm_count * 8));
}

static std::string decimal_format(int p, int s) { return "d:" + std::to_string(p) + "," + std::to_string(s); }
static bool is_unsigned(int vk) { return vk == V_U8 || vk == V_U16 || vk == V_U32 || vk == V_U64; }

// The packed key of a group table slot: <= 16 bytes, fields in group-column order, never straddling the 8-byte boundary.
//   inline layout   fixed-width columns by value, utf8 / binary columns as [length byte][bytes] sharing what is left
//                   (strings beyond that budget raise a flag at run time);
//   interned layout utf8 / binary columns - and, while the key is still too wide, the widest fixed-width columns - are
//                   replaced by 4-byte ids from an Interner (interner.cu): keys of any width (GroupValuesRows has no limit).
// Returns BGPU_EUNSUPPORTED (without touching ctx->err) when the layout does not fit.
int agg_key_layout(bgpu_agg* g, bool interned) {
  const size_t n_group = g->group_cols.size();
  std::vector<uint8_t> by_id(n_group, 0);
  int fixed_bytes = 0, n_var = 0;
  for (size_t i = 0; i < n_group; i++) {
    const Column& f = g->fields[(size_t)g->group_cols[i]];
    const bool var = f.kind == K_UTF8 || f.kind == K_BINARY;
    if (var && interned) { by_id[i] = 1; fixed_bytes += 4; }
    else if (var) n_var++;
    else fixed_bytes += f.kind == K_BOOL ? 1 : f.width;
  }
  if (interned) {
    while (fixed_bytes > 16) {  // widest fixed-width column not interned yet
      int best = -1;
      for (size_t i = 0; i < n_group; i++) {
        const Column& f = g->fields[(size_t)g->group_cols[i]];
        if (by_id[i] || f.kind != K_FIXED || f.width <= 4) continue;
        if (best < 0 || f.width > g->fields[(size_t)g->group_cols[(size_t)best]].width) best = (int)i;
      }
      if (best < 0) return BGPU_EUNSUPPORTED;
      fixed_bytes -= g->fields[(size_t)g->group_cols[(size_t)best]].width - 4;
      by_id[(size_t)best] = 1;
    }
  }
  const int var_budget = n_var ? (16 - fixed_bytes) / n_var : 0;  // bytes per inline utf8 key incl. its length byte
  if (fixed_bytes > 16 || (n_var && var_budget < 2)) return BGPU_EUNSUPPORTED;
  std::vector<int32_t> key_off, key_width;
  int off = 0;
  for (size_t i = 0; i < n_group; i++) {
    const Column& f = g->fields[(size_t)g->group_cols[i]];
    const bool var = (f.kind == K_UTF8 || f.kind == K_BINARY) && !by_id[i];
    const int w = by_id[i] ? 4 : var ? var_budget : (f.kind == K_BOOL ? 1 : f.width);
    if (!var && w <= 8 && off < 8 && off + w > 8) off = 8;  // do not straddle
    if (off + w > 16) return BGPU_EUNSUPPORTED;
    key_off.push_back(off);
    key_width.push_back(var ? w - 1 : w);
    off += w;
  }
  g->key_off = key_off;
  g->key_width = key_width;
  g->key_words = off > 8 ? 2 : 1;
  g->key_interned = by_id;
  g->interned_mode = interned;
  g->interners.assign(n_group, nullptr);
  return BGPU_OK;
}

static int agg_create_impl(bgpu_ctx* ctx, int32_t mode, const struct ArrowSchema* input_schema,
                           const int32_t* group_cols, int32_t n_group, const int32_t* agg_fns,
                           const bgpu_expr_node* nodes, const int32_t* agg_expr_offsets, const char* const* agg_names,
                           int32_t n_agg, const bgpu_expr_node* filter, int32_t n_filter, bool plan_only,
                           bgpu_agg** out) {
  if (!ctx || !out) return BGPU_EINVAL;
  *out = nullptr;
  if (!plan_only && !ctx->has_device)
    return ctx->fail(BGPU_ENODEVICE, "context has no device (libbgpu has no CPU fallback)");
  if (!input_schema || !input_schema->format || strcmp(input_schema->format, "+s") != 0)
    return ctx->fail(BGPU_EINVAL, "input_schema must be a struct schema");
  if (mode < BGPU_AGG_MODE_PARTIAL || mode > BGPU_AGG_MODE_SINGLE) return ctx->fail(BGPU_EINVAL, "bad aggregate mode");
  if (n_group < 1 || n_group > kMaxGroupCols)
    return ctx->fail(BGPU_EUNSUPPORTED, "1..8 group columns supported (no-GROUP-BY aggregates use AggregateStream)");
  if (n_agg < 0 || n_agg > 16) return ctx->fail(BGPU_EUNSUPPORTED, "at most 16 aggregates");
  if (!plan_only) BGPU_CUDA(ctx, cudaSetDevice(ctx->device));
  std::unique_ptr<bgpu_agg> g(new bgpu_agg(ctx));
  g->mode = mode;
  g->final_mode = mode == BGPU_AGG_MODE_FINAL || mode == BGPU_AGG_MODE_FINAL_PARTITIONED;
  g->fields.resize((size_t)input_schema->n_children);
  for (int64_t i = 0; i < input_schema->n_children; i++)
    BGPU_TRY(parse_field(ctx, input_schema->children[i], &g->fields[(size_t)i]));
  if (n_filter && g->final_mode) return ctx->fail(BGPU_EINVAL, "a filter applies to raw input rows (Partial/Single modes)");
  {  // the plan as given: key of the specialised-kernel cache
    std::string& k = g->plan_key;
    auto put = [&](const void* p, size_t n) { k.append((const char*)p, n); };
    put(&mode, 4); put(&n_group, 4); put(&n_agg, 4); put(&n_filter, 4);
    for (const Column& f : g->fields) { k += f.format; k += f.kind == K_DICT32 ? "|D" + f.dict->format : std::string(); k += ';'; }
    put(group_cols, (size_t)n_group * 4);
    if (n_agg) {
      put(agg_fns, (size_t)n_agg * 4);
      put(agg_expr_offsets, ((size_t)n_agg + 1) * 4);
      for (int32_t i = 0; i < agg_expr_offsets[n_agg]; i++) {  // field by field: the struct has padding
        const bgpu_expr_node& e = nodes[i];
        put(&e.op, 4); put(&e.a, 4); put(&e.precision, 4); put(&e.scale, 4); put(&e.lo, 8); put(&e.hi, 8);
      }
    }
    for (int32_t i = 0; i < n_filter; i++) {
      const bgpu_expr_node& e = filter[i];
      put(&e.op, 4); put(&e.a, 4); put(&e.precision, 4); put(&e.scale, 4); put(&e.lo, 8); put(&e.hi, 8);
    }
  }
  BGPU_TRY(compile_filter(ctx, g->fields, filter, n_filter, &g->filter));
  for (int32_t i = 0; i < n_group; i++) {
    const int32_t c = group_cols[i];
    if (c < 0 || c >= (int32_t)g->fields.size()) return ctx->fail(BGPU_EINVAL, "group column out of range");
    g->group_cols.push_back(c);
  }
  // packed key layout: inline when it fits 16 bytes, otherwise with the wide columns interned (agg_key_layout)
  if (agg_key_layout(g.get(), false) != BGPU_OK && agg_key_layout(g.get(), true) != BGPU_OK)
    return ctx->fail(BGPU_EUNSUPPORTED, "group key wider than 16 packed bytes even with every wide column interned (more than 4 such columns)");
  g->dicts.resize((size_t)n_group);

  // distinct values and accumulators (deduplicated: the reference computes sum(x) and avg(x)'s sum separately; the
  // values are equal)
  auto add_value = [&](const HValue& v) {
    for (size_t i = 0; i < g->values.size(); i++)
      if (values_equal(g->values[i], v)) return (int32_t)i;
    g->values.push_back(v);
    return (int32_t)g->values.size() - 1;
  };
  auto add_phys = [&](int32_t op, int32_t value) {
    for (size_t i = 0; i < g->phys.size(); i++)
      if (g->phys[i].op == op && g->phys[i].value == value) return (int32_t)i;
    PhysAcc p;
    p.op = op; p.value = value; p.words = op == OP_ADD_I128 ? 2 : 1;
    g->phys.push_back(p);
    return (int32_t)g->phys.size() - 1;
  };
  auto state_value = [&](int32_t batch_col, int32_t vk, int32_t cls) {  // Final modes: a state column of the batch
    HValue v;
    HFactor f;
    f.x.is_col = true; f.x.col = batch_col; f.x.vk = vk; f.x.cls = cls;
    v.f.push_back(f);
    v.cls = cls; v.vk = vk;
    return add_value(v);
  };
  int32_t state_col = n_group;  // Final modes: next state column of the incoming batches
  for (int32_t i = 0; i < n_agg; i++) {
    LogicalAgg a;
    a.fn = agg_fns[i];
    a.name = (agg_names && agg_names[i]) ? agg_names[i] : ("agg" + std::to_string(i));
    HValue v;
    const int32_t b = agg_expr_offsets[i], e = agg_expr_offsets[i + 1];
    BGPU_TRY(compile_value(ctx, g->fields, nodes + b, e - b, &v));
    if (a.fn == BGPU_AGG_COUNT) {
      a.result_format = "l";
      a.phys[0] = g->final_mode ? add_phys(OP_ADD_I64, state_value(state_col++, V_I64, CLS_I64))
                                : add_phys(OP_COUNT, add_value(v));
    } else if (a.fn == BGPU_AGG_SUM || a.fn == BGPU_AGG_AVG) {
      if (v.cls == CLS_NONE) return ctx->fail(BGPU_EUNSUPPORTED, "SUM/AVG need a numeric or Decimal128 argument");
      const bool avg = a.fn == BGPU_AGG_AVG;
      int32_t op, state_vk, state_cls;
      if (v.cls == CLS_I128) {
        op = OP_ADD_I128; state_vk = V_I128; state_cls = CLS_I128; a.decimal = true;
        const int p = v.precision, s = v.scale;
        a.sum_format = decimal_format(std::min(38, p + 10), s);
        a.result_format = avg ? decimal_format(std::min(38, p + 4), std::min(38, s + 4)) : a.sum_format;
        a.scale_up = std::min(38, s + 4) - s;
      } else if (avg || v.cls == CLS_F64) {
        op = OP_ADD_F64; state_vk = V_F64; state_cls = CLS_F64; a.sum_format = a.result_format = "g";
        v.cls = CLS_F64;  // planner coercion: avg(int) -> avg(CAST(x AS Float64))
      } else {
        const bool uns = v.plain_column() && is_unsigned(v.vk);
        op = OP_ADD_I64; state_vk = uns ? V_U64 : V_I64; state_cls = CLS_I64;
        a.sum_format = a.result_format = uns ? "L" : "l";
      }
      if (avg)
        a.phys[0] = g->final_mode ? add_phys(OP_ADD_I64, state_value(state_col++, V_U64, CLS_I64))
                                  : add_phys(OP_COUNT, add_value(v));
      a.phys[avg ? 1 : 0] = g->final_mode ? add_phys(op, state_value(state_col++, state_vk, state_cls))
                                          : add_phys(op, add_value(v));
    } else {
      return ctx->fail(BGPU_EUNSUPPORTED, "aggregate function not on the hot path (SUM, AVG, COUNT are)");
    }
    g->aggs.push_back(a);
  }
  if ((int)g->phys.size() > kMaxAcc) return ctx->fail(BGPU_EUNSUPPORTED, "too many accumulators");
  if ((int)g->values.size() > kMaxValues) return ctx->fail(BGPU_EUNSUPPORTED, "too many distinct aggregate arguments");
  // order values so that a value extending another one by one factor directly follows it (its leading factors are
  // then taken from the previous result), and sort the accumulators by value
  {
    const size_t nv = g->values.size();
    std::vector<int32_t> order;
    std::vector<bool> placed(nv, false), from_prev;
    for (size_t i = 0; i < nv; i++) {
      if (placed[i]) continue;
      bool is_ext = false;  // placed later, right after the value it extends
      for (size_t j = 0; j < nv; j++)
        if (j != i && value_extends(g->values[j], g->values[i])) is_ext = true;
      if (is_ext) continue;
      order.push_back((int32_t)i); placed[i] = true; from_prev.push_back(false);
      for (bool more = true; more;) {
        more = false;
        for (size_t j = 0; j < nv; j++)
          if (!placed[j] && value_extends(g->values[(size_t)order.back()], g->values[j])) {
            order.push_back((int32_t)j); placed[j] = true; from_prev.push_back(true);
            more = true;
            break;
          }
      }
    }
    for (size_t i = 0; i < nv; i++)  // extensions whose base was consumed by a sibling chain
      if (!placed[i]) { order.push_back((int32_t)i); placed[i] = true; from_prev.push_back(false); }
    std::vector<int32_t> new_index(nv);
    std::vector<HValue> vals;
    for (size_t k = 0; k < nv; k++) { new_index[(size_t)order[k]] = (int32_t)k; vals.push_back(g->values[(size_t)order[k]]); }
    g->values = vals;
    g->value_from_prev = from_prev;
    for (PhysAcc& p : g->phys) p.value = new_index[(size_t)p.value];
    std::vector<int32_t> pidx(g->phys.size());
    for (size_t i = 0; i < pidx.size(); i++) pidx[i] = (int32_t)i;
    std::stable_sort(pidx.begin(), pidx.end(), [&](int32_t x, int32_t y) { return g->phys[(size_t)x].value < g->phys[(size_t)y].value; });
    std::vector<int32_t> pnew(pidx.size());
    std::vector<PhysAcc> ph;
    for (size_t k = 0; k < pidx.size(); k++) { pnew[(size_t)pidx[k]] = (int32_t)k; ph.push_back(g->phys[(size_t)pidx[k]]); }
    g->phys = ph;
    for (LogicalAgg& a : g->aggs)
      for (int k = 0; k < 2; k++)
        if (a.phys[k] >= 0) a.phys[k] = pnew[(size_t)a.phys[k]];
  }
  {  // accumulator record layout (array of structs): i128 accumulators on 16-byte boundaries
    int32_t off = 0;
    for (PhysAcc& p : g->phys)
      if (p.words == 2) { p.out_off = off; off += 2; }
    for (PhysAcc& p : g->phys)
      if (p.words == 1) { p.out_off = off; off += 1; }
    g->acc_stride = (uint32_t)std::max(1, (off + 1) & ~1);
  }
  if (!plan_only) {
    BGPU_TRY(dev_alloc(ctx, 32, &g->counters));
    BGPU_CUDA(ctx, cudaMemsetAsync(g->counters->p, 0, 32, ctx->stream));
    BGPU_TRY(agg_resize(g.get(), 1u << 16, 0));
    BGPU_CUDA(ctx, cudaEventCreate(&g->ev[0]));
    BGPU_CUDA(ctx, cudaEventCreate(&g->ev[1]));
  }
  *out = g.release();
  return BGPU_OK;
}

extern "C" int bgpu_agg_create_expr(bgpu_ctx* ctx, int32_t mode, const struct ArrowSchema* input_schema,
                                    const int32_t* group_cols, int32_t n_group, const int32_t* agg_fns,
                                    const bgpu_expr_node* nodes, const int32_t* agg_expr_offsets,
                                    const char* const* agg_names, int32_t n_agg, const bgpu_expr_node* filter,
                                    int32_t n_filter, bgpu_agg** out) try {
  if (!group_cols || (n_agg > 0 && (!agg_fns || !agg_expr_offsets))) return BGPU_EINVAL;
  return agg_create_impl(ctx, mode, input_schema, group_cols, n_group, agg_fns, nodes, agg_expr_offsets, agg_names, n_agg,
                         filter, n_filter, false, out);
} catch (const std::exception& e) {
  return bgpu::abi_guard(ctx, e);  // nothing is thrown across the C ABI
}

extern "C" int bgpu_agg_create(bgpu_ctx* ctx, int32_t mode, const struct ArrowSchema* input_schema,
                               const int32_t* group_cols, int32_t n_group, const int32_t* agg_fns,
                               const int32_t* agg_cols, const char* const* agg_names, int32_t n_agg, bgpu_agg** out) try {
  if (!group_cols || n_agg < 0 || n_agg > 16 || (n_agg > 0 && (!agg_fns || !agg_cols))) return BGPU_EINVAL;
  std::vector<bgpu_expr_node> nodes;
  std::vector<int32_t> offs{0};
  for (int32_t i = 0; i < n_agg; i++) {
    if (agg_cols[i] >= 0) {
      bgpu_expr_node n{};
      n.op = BGPU_X_COLUMN;
      n.a = agg_cols[i];
      nodes.push_back(n);
    }
    offs.push_back((int32_t)nodes.size());
  }
  return agg_create_impl(ctx, mode, input_schema, group_cols, n_group, agg_fns, nodes.data(), offs.data(), agg_names, n_agg,
                         nullptr, 0, false, out);
} catch (const std::exception& e) {
  return bgpu::abi_guard(ctx, e);  // nothing is thrown across the C ABI
}

// Introspection (no device needed): generate the kernels of this plan for a synthetic batch shape and compile them
// with NVRTC. flags bit 0: every column carries validity; bit 1: dictionary keys need a remap; bit 2: seen tracking on.
// Writes <out_prefix>.cu and <out_prefix>.cubin when out_prefix is not NULL.
extern "C" int bgpu_agg_debug_compile(bgpu_ctx* ctx, int32_t mode, const struct ArrowSchema* input_schema,
                                      const int32_t* group_cols, int32_t n_group, const int32_t* agg_fns,
                                      const bgpu_expr_node* nodes, const int32_t* agg_expr_offsets, int32_t n_agg,
                                      const bgpu_expr_node* filter, int32_t n_filter, int32_t flags, int32_t threads,
                                      int32_t rows_per_thread, int32_t min_ctas, const char* out_prefix) try {
  bgpu_agg* gp = nullptr;
  BGPU_TRY(agg_create_impl(ctx, mode, input_schema, group_cols, n_group, agg_fns, nodes, agg_expr_offsets, nullptr, n_agg,
                           filter, n_filter, true, &gp));
  std::unique_ptr<bgpu_agg> g(gp);
  const bool nullable = flags & 1;
  JitShape sh;
  auto use = [&](int32_t col, const Column& f, int32_t vk) -> JitSlot& {
    JitSlot& sl = sh.slots[(size_t)jit_slot(&sh, col)];
    sl.kind = f.kind; sl.width = f.width; sl.nullable = nullable;
    if (vk != V_NONE) sl.vk = vk;
    return sl;
  };
  for (size_t i = 0; i < g->group_cols.size(); i++) {
    const int32_t bc = g->final_mode ? (int32_t)i : g->group_cols[i];
    JitSlot& sl = use(bc, g->fields[(size_t)g->group_cols[i]], V_NONE);
    sl.key_use = true;
    sl.remap = sl.kind == K_DICT32 && (flags & 2);
    sh.key_slot.push_back(jit_slot(&sh, bc));
  }
  for (const HFilterTerm& t : g->filter) {
    use(t.col, g->fields[(size_t)t.col], t.vk).filter_use = true;
    sh.filter_slot.push_back(jit_slot(&sh, t.col));
  }
  for (size_t vi = 0; vi < g->values.size(); vi++) {
    bool has_col = g->value_from_prev[vi] ? (bool)sh.value_nullable[vi - 1] : false;
    for (const HFactor& f : g->values[vi].f)
      for (const HAtom* a : {&f.x, &f.y}) {
        if ((a == &f.y && !f.op) || !a->is_col) continue;
        Column c;  // Final modes read state columns: fixed width by ValKind
        c.kind = K_FIXED;
        c.width = a->vk == V_I128 ? 16 : 8;
        use(a->col, g->final_mode ? c : g->fields[(size_t)a->col], a->vk).value_use = true;
        has_col = true;
      }
    sh.value_nullable.push_back(nullable && has_col);
  }
  int32_t first_count = -1, w = 0;
  for (size_t i = 0; i < g->phys.size(); i++) {
    int32_t alias = -1;
    if (g->phys[i].op == OP_COUNT && !sh.value_nullable[(size_t)g->phys[i].value]) {
      if (first_count < 0) first_count = (int32_t)i;
      else alias = first_count;
    }
    sh.acc_alias.push_back(alias);
    sh.acc_word0.push_back(alias < 0 ? w : 0);
    if (alias < 0) w += g->phys[i].words;
  }
  sh.W = w;
  sh.has_seen = flags & 4;
  sh.T = threads; sh.R = rows_per_thread; sh.min_ctas = min_ctas; sh.kernels = 7;
  sh.row32 = !(flags & 8);
  std::string src, cubin;
  BGPU_TRY(agg_generate_source(g.get(), sh, &src));
  src = std::string(agg_jit_prelude()) + src;
  if (out_prefix) {
    FILE* f = fopen((std::string(out_prefix) + ".cu").c_str(), "w");
    if (f) { fwrite(src.data(), 1, src.size(), f); fclose(f); }
  }
  BGPU_TRY(jit_compile_cubin(ctx, src, &cubin));
  if (out_prefix) {
    FILE* f = fopen((std::string(out_prefix) + ".cubin").c_str(), "wb");
    if (f) { fwrite(cubin.data(), 1, cubin.size(), f); fclose(f); }
  }
  return BGPU_OK;
} catch (const std::exception& e) {
  return bgpu::abi_guard(ctx, e);  // nothing is thrown across the C ABI
}

// A new input stream over the same plan (the reference's AggregateExec is created once per task and `execute`d per
// partition, /root/reference/ballista/core/src/execution_plans/shuffle_writer.rs:234-257): the table, the accumulators
// and the counters are cleared; the plan, the compiled kernels, the learned rung / capacities and the staging buffers of
// the streamed input stay.
extern "C" int bgpu_agg_reset(bgpu_agg* g) try {
  if (!g) return BGPU_EINVAL;
  bgpu_ctx* ctx = g->ctx;
  if (!ctx->has_device) return ctx->fail(BGPU_ENODEVICE, "context has no device (libbgpu has no CPU fallback)");
  BGPU_CUDA(ctx, cudaSetDevice(ctx->device));
  if (g->ingest_drain && !g->emitted) (void)g->ingest_drain(g);  // nothing may still be writing the table
  if (g->ingest_reset) g->ingest_reset(g);
  const size_t slot_bytes = g->key_words == 1 ? 16 : 32;
  BGPU_CUDA(ctx, cudaMemsetAsync(g->counters->p, 0, 32, ctx->stream));
  BGPU_CUDA(ctx, cudaMemsetAsync(g->slots->p, 0, (size_t)g->max_groups * 2 * slot_bytes, ctx->stream));
  BGPU_CUDA(ctx, cudaMemsetAsync(g->acc->p, 0, (size_t)g->max_groups * g->acc_stride * 8, ctx->stream));
  if (g->seen_tracking) BGPU_CUDA(ctx, cudaMemsetAsync(g->seen->p, 0, (size_t)g->max_groups * 4, ctx->stream));
  g->n_groups = 0;
  g->passthrough = false;
  g->any_committed = false;
  for (auto& in : g->interners)  // a new stream interns afresh, in the room the last one grew
    if (in) BGPU_TRY(bgpu::interner_clear(in.get()));
  g->pass_chunks.clear();
  g->emitted = false;
  g->hot_kernel_ms = 0.0;
  for (auto& d : g->dicts) {
    std::shared_ptr<bgpu::Interner> spare = d.interner ? d.interner : d.spare;
    if (spare) BGPU_TRY(bgpu::interner_clear(spare.get()));
    d = bgpu_agg::DictState();
    d.spare = spare;
  }
  return BGPU_OK;
} catch (const std::exception& e) {
  return bgpu::abi_guard(g ? g->ctx : nullptr, e);  // nothing is thrown across the C ABI
}

extern "C" void bgpu_jit_stats(uint64_t* out4) {
  if (out4) bgpu::jit_stats(out4);
}

extern "C" void bgpu_agg_destroy(bgpu_agg* g) {
  if (!g) return;
  if (g->ctx) cudaSetDevice(g->ctx->device);
  delete g;
}

extern "C" int64_t bgpu_agg_num_groups(bgpu_agg* g) { return g ? (int64_t)g->n_groups : -1; }
