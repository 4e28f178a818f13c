// Kernel-side plans of the grouped aggregate (POD, passed as __grid_constant__ kernel parameters) and the host-side
// compiler from the ABI's postfix expression programs (include/bgpu.h, bgpu_agg_create_expr) to those plans.
//
// Reference semantics restated here (none of it is in /root/reference):
//   * aggregate input expressions are evaluated inside the aggregate stream (plan text planner.rs:379);
//   * arrow-arith 49 numeric.rs decimal_op: +/- rescale both sides to max(s1,s2) and are i128-checked, * adds the
//     scales and is i128-checked; DataFusion coerces an Int64 literal next to a Decimal128 to Decimal128(20,0);
//   * Int64 arithmetic in DataFusion 34's BinaryExpr is add_wrapping/sub_wrapping/mul_wrapping; Float64 is IEEE.
#pragma once
#include <stdint.h>

#include <string>
#include <vector>

#include "internal.h"

namespace bgpu {

enum ValKind : int32_t { V_NONE = 0, V_I8, V_I16, V_I32, V_I64, V_U8, V_U16, V_U32, V_U64, V_F32, V_F64, V_I128 };
enum AccOp : int32_t { OP_ADD_I64 = 0, OP_ADD_I128 = 1, OP_ADD_F64 = 2, OP_COUNT = 3 };
enum ValClass : int32_t { CLS_NONE = 0 /* COUNT(*): no value */, CLS_I64 = 1, CLS_I128 = 2, CLS_F64 = 3 };

constexpr int kMaxGroupCols = 8;
constexpr int kMaxAcc = 24;
constexpr int kMaxValues = 24;
constexpr int kMaxFactors = 3;
constexpr int kMaxFilterTerms = 4;

// ---- group keys ---------------------------------------------------------------------------------------------
struct KeyCol {
  const void* values;
  const int32_t* offsets;
  const uint8_t* validity;
  const int32_t* remap;  // K_DICT32: batch dictionary index -> canonical index (-1 = null value); nullptr = identity
  int32_t kind;          // ColKind
  int32_t width;         // fixed width; utf8: max inline bytes (excluding the length byte)
  int32_t byte_off;      // position inside the 16-byte packed key
  int32_t pad;
};
struct KeyPlan {
  KeyCol cols[kMaxGroupCols];
  int32_t n;
  int32_t words;  // 1: packed key fits 8 bytes, 2: 16 bytes
};

// The global group table: open addressing with the packed key stored INSIDE the slot, so a probe is one 16-byte
// (one key word) or 32-byte (two key words) access — one DRAM sector — instead of entry + key + null-bits accesses.
//   slot = [k0][k1 if key_words == 2][sm][pad if key_words == 2],  sm = (null bits << 32) | state,
//   state: 0 empty, 1 busy (being published), >= 2: dense group id + 2.
// Accumulators are one record of `acc_stride` u64 per group (array of structs): all of a group's words share one or
// two sectors. Dense key columns and Arrow accumulator columns are produced at emit.
struct GroupTable {
  uint64_t* slots;
  uint32_t mask;        // capacity - 1 (capacity = 2 * max_groups, power of two)
  uint32_t max_groups;
  uint32_t* n_groups;   // device counter
  int32_t* flags;       // [0] capacity overflow, [1] string key too long, [2] arithmetic overflow,
                        // [3] fused pass abort (a group id beyond the per-CTA capacity)
  uint64_t* acc;        // max_groups * acc_stride u64
  uint32_t acc_stride;
  uint32_t slot_words;  // 2 or 4
};

// ---- value expressions: product of up to three factors, each `x`, `x + y` or `x - y` ------------------------------
struct Atom {
  const void* col;          // nullptr: literal
  const uint8_t* validity;  // nullptr: no nulls in this batch
  uint64_t lit_lo, lit_hi;  // literal, already rescaled to the factor's scale
  uint64_t mul;             // column rescale multiplier 10^k (1 = none)
  int32_t vk;               // ValKind of the column
  int32_t pad;
};
struct Factor {
  Atom x, y;
  int32_t op;  // 0: x only, 1: x + y, 2: x - y
  int32_t pad;
};
struct ValueExpr {
  Factor f[kMaxFactors];
  int32_t n_factors;
  int32_t cls;          // ValClass
  int32_t from_prev;    // 1: the first factor is the previous value's result (common sub-expression)
  int32_t acc_begin;    // accumulators [acc_begin, acc_end) consume this value
  int32_t acc_end;
  int32_t pad;
};
struct ValuePlan {
  ValueExpr v[kMaxValues];
  int32_t n;
  int32_t pad;
};

struct FilterTerm {
  const void* col;
  const uint8_t* validity;
  uint64_t lit_lo, lit_hi;  // i128 literal (integers, dates, decimals at the column's scale) or f64 bits in lit_lo
  int32_t vk;
  int32_t cmp;  // BGPU_X_LT .. BGPU_X_NE
};
struct FilterPlan {
  FilterTerm t[kMaxFilterTerms];
  int32_t n;
  int32_t pad;
};

// ---- accumulators -------------------------------------------------------------------------------------------
struct AccK {
  int32_t out_off;   // first word of this accumulator inside a group's record (GroupTable::acc)
  int32_t op;        // AccOp
  int32_t words;     // 1 or 2
  int32_t word0;     // first word of this accumulator inside a group's per-CTA (private / replicated) state
  int32_t seen_bit;  // bit in seen[gid]
  int32_t alias;     // >= 0: this batch it equals accumulator `alias` (counts over null-free values); skipped in the
                     // row loop, the alias' partial is added at flush time. -1 otherwise
};
struct AccPlan {
  AccK a[kMaxAcc];
  int32_t n;
  int32_t total_words;
  uint32_t* seen;   // per-group bitmask of accumulators that saw a non-null input; nullptr = not tracked yet
  uint64_t* acc;    // GroupTable::acc
  uint32_t stride;  // GroupTable::acc_stride
  uint32_t pad;
};

// ---- host-side logical description (independent of any batch) -----------------------------------------------------
struct HAtom {
  bool is_col = false;
  int32_t col = -1;
  __int128 lit = 0;       // literal at its own scale
  double flit = 0.0;
  int32_t precision = 0, scale = 0;  // decimal type of this atom (ints: scale 0)
  int32_t cls = CLS_NONE;
  int32_t vk = V_NONE;    // column: ValKind of the column it must be bound to
};
struct HFactor {
  HAtom x, y;
  int32_t op = 0;
  int32_t precision = 0, scale = 0;
  __int128 x_mul = 1, y_mul = 1;  // rescale multipliers to the factor's scale
};
struct HValue {
  std::vector<HFactor> f;
  int32_t cls = CLS_NONE;
  int32_t precision = 0, scale = 0;  // decimal result type
  int32_t vk = V_NONE;               // plain column: its ValKind (drives Sum's result type)
  bool plain_column() const { return f.size() == 1 && f[0].op == 0 && f[0].x.is_col; }
};
struct HFilterTerm {
  int32_t col = -1, cmp = 0, vk = V_NONE;
  __int128 lit = 0;
  double flit = 0.0;
};

// ---- runtime-specialised kernels (agg_jit_prelude.cuh + agg_codegen.cu) ------------------------------------------------
constexpr int kJitSlots = 24;
constexpr int kJitAcc = 24;
// Host mirror of the prelude's JitArgs: the single by-value argument of every generated kernel (same layout).
struct JitArgs {
  const void* col[kJitSlots];
  const int32_t* aux[kJitSlots];
  const uint8_t* valid[kJitSlots];
  GroupTable t;
  uint32_t* seen;
  uint64_t* scratch;
  uint32_t* gids;
  int64_t n_rows;
  uint32_t g_cap, cache_slots, n_groups, replicas;
  uint16_t* bucket_ids;
  uint32_t* bucket_hist;
  uint32_t bucket_shift, bucket_chunk;
};

// Compile one aggregate-argument program. `fields` is the raw input schema. Returns BGPU_OK or fails ctx.
int compile_value(bgpu_ctx* ctx, const std::vector<Column>& fields, const bgpu_expr_node* nodes, int32_t n,
                  HValue* out);
int compile_filter(bgpu_ctx* ctx, const std::vector<Column>& fields, const bgpu_expr_node* nodes, int32_t n,
                   std::vector<HFilterTerm>* out);
ValKind vk_of(const Column& c);
bool values_equal(const HValue& a, const HValue& b);
// b == a * (one more factor)?
bool value_extends(const HValue& a, const HValue& b);

}  // namespace bgpu
