/* ORACLE (test infrastructure only — never linked into or called by the product path).
 *
 * Plain-C ABI of the CPU restatement of the reference's aggregation + shuffle hot path. The arithmetic
 * restated here lives in crates that are NOT vendored in /root/reference:
 * datafusion git coralogix/arrow-datafusion tag v34.0.0-cx.1 (/root/reference/Cargo.toml:31-36),
 * arrow 49.0.0 (/root/reference/Cargo.toml:25), ahash "0.8" (/root/reference/ballista/core/Cargo.toml:47).
 * Anchors in the reference: shuffle_writer.rs:336-346 (BatchPartitioner call sites),
 * scheduler/src/planner.rs:376-427 (Partial / FinalPartitioned plan shape).
 * PARITY STATUS: the reference's own tests hold no hash value, partition id (n > 2) or grouped result, so DataFusion's glue
 * (seed derivation, combine_hashes, null handling, group order) stays "parity unpinned". Pinned: the ahash hasher against
 * known answers produced by executing compiled upstream ahash-0.8.12 (tests/golden/ahash_0_8_12_kat.json), the Partial ->
 * Final arithmetic against the only aggregate values the reference's tests assert (SUM = 28, AVG = 3.5, COUNT = 8 over its
 * alltypes_plain table, ballista/client/src/context.rs:784-815) and against TPC-H Q1 over its own lineitem test data.
 */
#ifndef ORACLE_H
#define ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

enum {
  ORC_FIXED = 0, /* any fixed-width primitive: width 1,2,4,8,16 bytes (ints, floats as bits, dates, decimal128) */
  ORC_BOOL = 1,  /* bit-packed boolean values */
  ORC_UTF8 = 2,  /* i32 offsets + data */
  ORC_BINARY = 3,
  ORC_DICT_I32 = 4 /* Dictionary<Int32, *>: values = i32 keys, dict = values column */
};

typedef struct orc_col {
  int32_t type;
  int32_t width;           /* ORC_FIXED only */
  const void* values;      /* fixed values / utf8 data / bool bits / dict keys */
  const int32_t* offsets;  /* utf8 / binary */
  const uint8_t* validity; /* Arrow validity bitmap (LSB first) or NULL */
  int64_t offset;          /* Arrow logical offset into all buffers */
  const struct orc_col* dict;
  int64_t dict_len;
} orc_col;

/* the ahash fallback hasher with explicit RandomState keys {k0,k1,k2,k3}: hash_one(&[u8]) and a run of write_u64 */
uint64_t orc_ahash_keyed_bytes(const uint64_t* keys, const uint8_t* d, uint64_t len);
uint64_t orc_ahash_keyed_u64s(const uint64_t* keys, const uint64_t* xs, int32_t n);

/* hash_utils.rs create_hashes: hashes[] must be zero-filled by the caller (as BatchPartitioner does). */
int orc_create_hashes(const orc_col* cols, int32_t n_cols, int64_t n_rows, uint64_t* hashes);

/* repartition/mod.rs BatchPartitioner::partition, hash arm: partition = hash % n; per-partition row index
 * lists in input order. out_counts[n], out_indices[n_rows] grouped by partition ascending (stable). */
int orc_partition_indices(const uint64_t* hashes, int64_t n_rows, uint32_t n, uint64_t* out_counts,
                          uint64_t* out_indices);

/* arrow-select take.rs for fixed-width values and validity; utf8 in two steps. */
int orc_take_fixed(const void* values, int32_t width, int64_t in_offset, const uint64_t* idx, int64_t n,
                   void* out);
int orc_take_bits(const uint8_t* bits, int64_t in_offset, const uint64_t* idx, int64_t n, uint8_t* out_bits);
int64_t orc_take_utf8_offsets(const int32_t* offsets, int64_t in_offset, const uint64_t* idx, int64_t n,
                              int32_t* out_offsets);
int orc_take_utf8_data(const int32_t* offsets, const uint8_t* data, int64_t in_offset, const uint64_t* idx,
                       int64_t n, const int32_t* out_offsets, uint8_t* out_data);

/* aggregates/group_values: intern keys -> dense group index in first-seen order. Null keys form their own
 * group per column-null pattern. State persists in the handle across batches. */
typedef struct orc_groups orc_groups;
orc_groups* orc_groups_new(void);
void orc_groups_free(orc_groups*);
/* returns total number of groups after this batch; out_gid[n_rows]; rows_seen_before = global row number of
 * this batch's first row, used for out_first_row bookkeeping */
int64_t orc_groups_intern(orc_groups*, const orc_col* keys, int32_t n_keys, int64_t n_rows, int64_t* out_gid);
/* global row number (across all interned batches) of each group's first row: out[n_groups] */
int orc_groups_first_rows(const orc_groups*, int64_t* out);

/* GroupsAccumulator::update_batch / merge_batch (prim_op.rs, count.rs, average.rs). All arrays sized
 * n_groups by the caller; seen[] is a byte per group (NullState). validity may be NULL. */
int orc_acc_sum_i64(const int64_t* v, const uint8_t* validity, int64_t offset, const int64_t* gid, int64_t n,
                    int64_t* sum, uint8_t* seen);
int orc_acc_sum_i128(const void* v, const uint8_t* validity, int64_t offset, const int64_t* gid, int64_t n,
                     void* sum /* i128[n_groups] */, uint8_t* seen);
int orc_acc_sum_f64(const double* v, const uint8_t* validity, int64_t offset, const int64_t* gid, int64_t n,
                    double* sum, uint8_t* seen);
int orc_acc_count(const uint8_t* validity, int64_t offset, const int64_t* gid, int64_t n, int64_t* count);
/* min / max for i64 and f64 (not on the north-star list; used for widening) */
int orc_acc_minmax_i64(const int64_t* v, const uint8_t* validity, int64_t offset, const int64_t* gid, int64_t n,
                       int is_max, int64_t* acc, uint8_t* seen);

/* AvgGroupsAccumulator::evaluate for Decimal128: (sum * 10^(scale_up)) / count, truncating. Returns the
 * number of groups whose scaled multiply overflowed i128 (reference: error). */
int64_t orc_avg_decimal_evaluate(const void* sum, const uint64_t* count, int64_t n_groups, int32_t scale_up,
                                 void* out);

/* Timed CPU baseline legs (bench.py cpu_baseline / --impl reference). Multi-threaded, one thread per input
 * partition like Ballista's concurrent_tasks. See oracle/pipeline.cpp. */
#ifdef __cplusplus
}
#endif
#endif
