/* ORACLE (test infrastructure only — never linked into or called by the product path).
 * CPU restatement of DataFusion v34 create_hashes / BatchPartitioner / take / GroupsAccumulators.
 * See oracle.h for provenance and parity status (no reference build on this box). */
#include "oracle.h"

#include <string>
#include <unordered_map>
#include <vector>

#include "ahash_fallback.h"

static inline bool bit_get(const uint8_t* bits, int64_t i) { return (bits[i >> 3] >> (i & 7)) & 1; }
static inline void bit_set(uint8_t* bits, int64_t i) { bits[i >> 3] |= (uint8_t)(1u << (i & 7)); }

/* hash_utils.rs HashValue::hash_one for one non-null value of column c at logical row r */
static inline uint64_t value_hash(const orc_col* c, int64_t r) {
  int64_t i = r + c->offset;
  switch (c->type) {
    case ORC_FIXED: {
      const uint8_t* p = (const uint8_t*)c->values + i * c->width;
      if (c->width == 16) return ah_hash_u128(ah_rd(p, 8), ah_rd(p + 8, 8));
      return ah_hash_u64(ah_rd(p, c->width)); /* write_u{8,16,32,64}: zero-extended bit pattern */
    }
    case ORC_BOOL:
      return ah_hash_u64(bit_get((const uint8_t*)c->values, i) ? 1 : 0); /* bool -> write_u8 */
    case ORC_UTF8: {
      int32_t b = c->offsets[i], e = c->offsets[i + 1];
      return ah_hash_str((const uint8_t*)c->values + b, (uint64_t)(e - b));
    }
    case ORC_BINARY: {
      int32_t b = c->offsets[i], e = c->offsets[i + 1];
      return ah_hash_bytes((const uint8_t*)c->values + b, (uint64_t)(e - b));
    }
    default:
      return 0;
  }
}

/* The hasher with explicit RandomState keys {k0,k1,k2,k3} (as stored, i.e. after with_seeds' xor): what the executed
 * known-answer vectors of tests/golden/ahash_0_8_12_kat.json are checked against. */
static ah_hasher ah_from_keys(const uint64_t* k) {
  ah_hasher h;
  h.buffer = k[1];
  h.pad = k[0];
  h.extra0 = k[2];
  h.extra1 = k[3];
  return h;
}
/* RandomState::hash_one::<&[u8]>: write_usize(len); write(bytes); finish() */
extern "C" uint64_t orc_ahash_keyed_bytes(const uint64_t* keys, const uint8_t* d, uint64_t len) {
  ah_hasher h = ah_from_keys(keys);
  ah_update(&h, len);
  ah_write(&h, d, len);
  return ah_finish(&h);
}
/* write_u64(xs[0]); ...; write_u64(xs[n-1]); finish() */
extern "C" uint64_t orc_ahash_keyed_u64s(const uint64_t* keys, const uint64_t* xs, int32_t n) {
  ah_hasher h = ah_from_keys(keys);
  for (int32_t i = 0; i < n; i++) ah_update(&h, xs[i]);
  return ah_finish(&h);
}

extern "C" int orc_create_hashes(const orc_col* cols, int32_t n_cols, int64_t n_rows, uint64_t* hashes) {
  for (int32_t ci = 0; ci < n_cols; ci++) {
    const orc_col* c = &cols[ci];
    const bool rehash = ci >= 1; /* create_hashes: "combine hashes ... for all columns besides the first" */
    if (c->type == ORC_DICT_I32) {
      /* hash_dictionary: hash each dictionary value once (as a first column: nulls stay 0), then look up
       * per key; null keys leave the running hash untouched. */
      std::vector<uint64_t> dh((size_t)c->dict_len, 0);
      if (orc_create_hashes(c->dict, 1, c->dict_len, dh.data())) return 1;
      const int32_t* keys = (const int32_t*)c->values;
      for (int64_t r = 0; r < n_rows; r++) {
        int64_t i = r + c->offset;
        if (c->validity && !bit_get(c->validity, i)) continue;
        uint64_t vh = dh[(size_t)keys[i]];
        hashes[r] = rehash ? df_combine_hashes(vh, hashes[r]) : vh;
      }
      continue;
    }
    if (c->type < ORC_FIXED || c->type > ORC_BINARY) return 1;
    if (c->type == ORC_FIXED && c->width != 1 && c->width != 2 && c->width != 4 && c->width != 8 &&
        c->width != 16)
      return 1;
    for (int64_t r = 0; r < n_rows; r++) {
      if (c->validity && !bit_get(c->validity, r + c->offset)) continue; /* null rows are skipped */
      uint64_t vh = value_hash(c, r);
      hashes[r] = rehash ? df_combine_hashes(vh, hashes[r]) : vh;
    }
  }
  return 0;
}

extern "C" int orc_partition_indices(const uint64_t* hashes, int64_t n_rows, uint32_t n, uint64_t* out_counts,
                                     uint64_t* out_indices) {
  if (n == 0) return 1;
  std::vector<uint64_t> pos(n, 0);
  for (uint32_t p = 0; p < n; p++) out_counts[p] = 0;
  for (int64_t r = 0; r < n_rows; r++) out_counts[hashes[r] % n]++;
  uint64_t acc = 0;
  for (uint32_t p = 0; p < n; p++) {
    pos[p] = acc;
    acc += out_counts[p];
  }
  for (int64_t r = 0; r < n_rows; r++) out_indices[pos[hashes[r] % n]++] = (uint64_t)r;
  return 0;
}

extern "C" int orc_take_fixed(const void* values, int32_t width, int64_t in_offset, const uint64_t* idx,
                              int64_t n, void* out) {
  const uint8_t* src = (const uint8_t*)values + in_offset * width;
  uint8_t* dst = (uint8_t*)out;
  for (int64_t i = 0; i < n; i++) memcpy(dst + i * width, src + idx[i] * (uint64_t)width, (size_t)width);
  return 0;
}

extern "C" int orc_take_bits(const uint8_t* bits, int64_t in_offset, const uint64_t* idx, int64_t n,
                             uint8_t* out_bits) {
  memset(out_bits, 0, (size_t)((n + 7) / 8));
  for (int64_t i = 0; i < n; i++)
    if (bit_get(bits, in_offset + (int64_t)idx[i])) bit_set(out_bits, i);
  return 0;
}

extern "C" int64_t orc_take_utf8_offsets(const int32_t* offsets, int64_t in_offset, const uint64_t* idx,
                                         int64_t n, int32_t* out_offsets) {
  int64_t acc = 0;
  out_offsets[0] = 0;
  for (int64_t i = 0; i < n; i++) {
    int64_t r = in_offset + (int64_t)idx[i];
    acc += offsets[r + 1] - offsets[r];
    out_offsets[i + 1] = (int32_t)acc;
  }
  return acc;
}

extern "C" int orc_take_utf8_data(const int32_t* offsets, const uint8_t* data, int64_t in_offset,
                                  const uint64_t* idx, int64_t n, const int32_t* out_offsets,
                                  uint8_t* out_data) {
  for (int64_t i = 0; i < n; i++) {
    int64_t r = in_offset + (int64_t)idx[i];
    memcpy(out_data + out_offsets[i], data + offsets[r], (size_t)(offsets[r + 1] - offsets[r]));
  }
  return 0;
}

/* ---- aggregates/group_values/row.rs restated: keys -> dense group index, first-seen order ----------- */
struct orc_groups {
  std::unordered_map<std::string, int64_t> map;
  std::vector<int64_t> first_row;
  int64_t rows_seen = 0;
};

extern "C" orc_groups* orc_groups_new(void) { return new orc_groups(); }
extern "C" void orc_groups_free(orc_groups* g) { delete g; }

/* Row encoding used only for equality: per key column one null byte then the value bytes (utf8/binary with
 * a 4-byte length prefix; dictionary keys are resolved to their values like GroupValuesRows does through
 * arrow-row). */
static void encode_key(const orc_col* c, int64_t r, std::string& out) {
  int64_t i = r + c->offset;
  if (c->validity && !bit_get(c->validity, i)) {
    out.push_back('\0');
    return;
  }
  if (c->type == ORC_DICT_I32) {
    int32_t k = ((const int32_t*)c->values)[i];
    encode_key(c->dict, k, out);
    return;
  }
  out.push_back('\1');
  switch (c->type) {
    case ORC_FIXED:
      out.append((const char*)c->values + i * c->width, (size_t)c->width);
      break;
    case ORC_BOOL:
      out.push_back(bit_get((const uint8_t*)c->values, i) ? '\1' : '\0');
      break;
    case ORC_UTF8:
    case ORC_BINARY: {
      int32_t b = c->offsets[i], e = c->offsets[i + 1];
      int32_t len = e - b;
      out.append((const char*)&len, 4);
      out.append((const char*)c->values + b, (size_t)len);
      break;
    }
  }
}

extern "C" int64_t orc_groups_intern(orc_groups* g, const orc_col* keys, int32_t n_keys, int64_t n_rows,
                                     int64_t* out_gid) {
  std::string k;
  for (int64_t r = 0; r < n_rows; r++) {
    k.clear();
    for (int32_t c = 0; c < n_keys; c++) encode_key(&keys[c], r, k);
    auto it = g->map.find(k);
    if (it == g->map.end()) {
      int64_t id = (int64_t)g->first_row.size();
      g->map.emplace(k, id);
      g->first_row.push_back(g->rows_seen + r);
      out_gid[r] = id;
    } else {
      out_gid[r] = it->second;
    }
  }
  g->rows_seen += n_rows;
  return (int64_t)g->first_row.size();
}

extern "C" int orc_groups_first_rows(const orc_groups* g, int64_t* out) {
  for (size_t i = 0; i < g->first_row.size(); i++) out[i] = g->first_row[i];
  return 0;
}

/* ---- GroupsAccumulator::update_batch restated (prim_op.rs / count.rs / average.rs) ------------------- */
#define VALID(validity, off, r) (!(validity) || bit_get((validity), (off) + (r)))

extern "C" int orc_acc_sum_i64(const int64_t* v, const uint8_t* validity, int64_t offset, const int64_t* gid,
                               int64_t n, int64_t* sum, uint8_t* seen) {
  for (int64_t r = 0; r < n; r++)
    if (VALID(validity, offset, r)) {
      sum[gid[r]] = (int64_t)((uint64_t)sum[gid[r]] + (uint64_t)v[offset + r]); /* add_wrapping */
      seen[gid[r]] = 1;
    }
  return 0;
}

extern "C" int orc_acc_sum_i128(const void* v, const uint8_t* validity, int64_t offset, const int64_t* gid,
                                int64_t n, void* sum, uint8_t* seen) {
  const unsigned __int128* x = (const unsigned __int128*)v;
  unsigned __int128* s = (unsigned __int128*)sum;
  for (int64_t r = 0; r < n; r++)
    if (VALID(validity, offset, r)) {
      unsigned __int128 a;
      memcpy(&a, &x[offset + r], 16); /* values may be only 8-byte aligned */
      unsigned __int128 b;
      memcpy(&b, &s[gid[r]], 16);
      b += a; /* i128 add_wrapping */
      memcpy(&s[gid[r]], &b, 16);
      seen[gid[r]] = 1;
    }
  return 0;
}

extern "C" int orc_acc_sum_f64(const double* v, const uint8_t* validity, int64_t offset, const int64_t* gid,
                               int64_t n, double* sum, uint8_t* seen) {
  for (int64_t r = 0; r < n; r++)
    if (VALID(validity, offset, r)) {
      sum[gid[r]] += v[offset + r]; /* arrival order */
      seen[gid[r]] = 1;
    }
  return 0;
}

extern "C" int orc_acc_count(const uint8_t* validity, int64_t offset, const int64_t* gid, int64_t n,
                             int64_t* count) {
  for (int64_t r = 0; r < n; r++)
    if (VALID(validity, offset, r)) count[gid[r]]++;
  return 0;
}

extern "C" int orc_acc_minmax_i64(const int64_t* v, const uint8_t* validity, int64_t offset, const int64_t* gid,
                                  int64_t n, int is_max, int64_t* acc, uint8_t* seen) {
  for (int64_t r = 0; r < n; r++)
    if (VALID(validity, offset, r)) {
      int64_t x = v[offset + r], g = gid[r];
      if (!seen[g] || (is_max ? x > acc[g] : x < acc[g])) acc[g] = x;
      seen[g] = 1;
    }
  return 0;
}

extern "C" int64_t orc_avg_decimal_evaluate(const void* sum, const uint64_t* count, int64_t n_groups,
                                            int32_t scale_up, void* out) {
  const __int128* s = (const __int128*)sum;
  __int128* o = (__int128*)out;
  __int128 mul = 1;
  for (int32_t i = 0; i < scale_up; i++) mul *= 10;
  int64_t overflowed = 0;
  for (int64_t g = 0; g < n_groups; g++) {
    __int128 a, scaled;
    memcpy(&a, &s[g], 16);
    if (__builtin_mul_overflow(a, mul, &scaled)) overflowed++;
    __int128 q = count[g] ? scaled / (__int128)count[g] : 0; /* truncating i128 division */
    memcpy(&o[g], &q, 16);
  }
  return overflowed;
}
