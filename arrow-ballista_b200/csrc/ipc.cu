// Arrow IPC stream framing of partitioned device batches.
//
// Replaces, for the shuffle write, what ShuffleWriterExec does per (input batch, output partition) with arrow-ipc's
// StreamWriter (/root/reference/ballista/core/src/execution_plans/shuffle_writer.rs:351-399, `IPCWriter` in
// /root/reference/ballista/core/src/utils.rs:63-141): one encapsulated RecordBatch message per non-empty partition.
//
// Division of labour:
//   device  one kernel lays the BODY of every partition's message out exactly as the stream needs it: buffers padded to
//           8 bytes, validity / boolean bitmaps re-aligned to bit 0 of the partition's first row (null counts taken on
//           the way), utf8 offsets rebased to 0, value bytes copied; the staging buffer mirrors the final byte layout of
//           all messages (holes where the metadata goes), so ONE device-to-host copy moves everything
//   host    the flatbuffer metadata (Message / RecordBatch / Schema tables: a few hundred bytes, sizes known up front)
//           is written into the holes by a minimal flatbuffer builder (no flatbuffers dependency)
// The messages are uncompressed (the reference writes LZ4_FRAME bodies; its StreamReader - and pyarrow - read both).
// Dictionary-encoded columns would need DictionaryBatch messages and are BGPU_EUNSUPPORTED here.
#include <string.h>

#include <algorithm>
#include <mutex>
#include <string>
#include <vector>

#include "internal.h"

namespace bgpu {
namespace {

// ---- a minimal flatbuffer builder (back to front, as the reference implementation builds) -----------------------
class Fbb {
 public:
  explicit Fbb(size_t cap = 1024) : buf_(cap), head_(cap) {}
  uint32_t size() const { return (uint32_t)(buf_.size() - head_); }
  const uint8_t* data() const { return buf_.data() + head_; }

  void prep(size_t align, size_t additional) {
    if (align > minalign_) minalign_ = align;
    const size_t pad = (~(size() + additional) + 1) & (align - 1);
    while (head_ < pad + align + additional) grow();
    for (size_t i = 0; i < pad; i++) buf_[--head_] = 0;
  }
  template <typename T>
  void push(T v) {
    prep(sizeof(T), 0);
    head_ -= sizeof(T);
    memcpy(buf_.data() + head_, &v, sizeof(T));
  }
  // relative offset to something already written (`off` = size() right after it was finished)
  void push_ref(uint32_t off) {
    prep(4, 0);
    push<uint32_t>(size() - off + 4);
  }
  uint32_t string(const std::string& s) {
    prep(4, s.size() + 1);
    buf_[--head_] = 0;
    head_ -= s.size();
    memcpy(buf_.data() + head_, s.data(), s.size());
    push<uint32_t>((uint32_t)s.size());
    return size();
  }
  // vector of 16-byte structs of two int64 (FieldNode, Buffer)
  uint32_t struct_vector(const std::vector<std::pair<int64_t, int64_t>>& v) {
    prep(4, v.size() * 16);
    prep(8, v.size() * 16);
    for (size_t i = v.size(); i-- > 0;) {
      push<int64_t>(v[i].second);
      push<int64_t>(v[i].first);
    }
    push<uint32_t>((uint32_t)v.size());
    return size();
  }
  uint32_t offset_vector(const std::vector<uint32_t>& offs) {
    prep(4, offs.size() * 4);
    for (size_t i = offs.size(); i-- > 0;) push_ref(offs[i]);
    push<uint32_t>((uint32_t)offs.size());
    return size();
  }
  // tables: start, add fields in any order (each at most once), end
  void start_table(int n_fields) {
    slots_.assign((size_t)n_fields, 0);
    object_start_ = size();
  }
  template <typename T>
  void add(int id, T v) {  // scalars are always written (no default elision: message sizes stay value independent)
    push<T>(v);
    slots_[(size_t)id] = size();
  }
  void add_ref(int id, uint32_t off) {
    push_ref(off);
    slots_[(size_t)id] = size();
  }
  uint32_t end_table() {
    push<int32_t>(0);  // soffset to the vtable, patched below
    const uint32_t table = size();
    for (size_t i = slots_.size(); i-- > 0;) push<uint16_t>(slots_[i] ? (uint16_t)(table - slots_[i]) : 0);
    push<uint16_t>((uint16_t)(table - object_start_));
    push<uint16_t>((uint16_t)((slots_.size() + 2) * 2));
    const int32_t to_vtable = (int32_t)(size() - table);
    memcpy(buf_.data() + buf_.size() - table, &to_vtable, 4);
    return table;
  }
  void finish(uint32_t root) {
    prep(minalign_, 4);
    push_ref(root);
  }

 private:
  void grow() {
    const size_t old = buf_.size(), used = old - head_;
    std::vector<uint8_t> nb(old * 2);
    memcpy(nb.data() + nb.size() - used, buf_.data() + head_, used);
    head_ = nb.size() - used;
    buf_.swap(nb);
  }
  std::vector<uint8_t> buf_;
  size_t head_;
  size_t minalign_ = 1;
  std::vector<uint32_t> slots_;
  uint32_t object_start_ = 0;
};

// Arrow's Schema.fbs / Message.fbs, the part this path needs
enum FbType : uint8_t { T_Int = 2, T_FloatingPoint = 3, T_Binary = 4, T_Utf8 = 5, T_Bool = 6, T_Decimal = 7, T_Date = 8, T_Time = 9, T_Timestamp = 10 };
enum FbHeader : uint8_t { H_Schema = 1, H_DictionaryBatch = 2, H_RecordBatch = 3 };
constexpr int16_t kMetadataV5 = 4;

// the `type` union member of a Field from the Arrow C format string; false = not supported by this encoder
bool fb_type(Fbb& b, const std::string& f, uint8_t* type_type, uint32_t* type_off) {
  auto int_type = [&](int bits, bool is_signed) {
    b.start_table(2);
    b.add<int32_t>(0, bits);
    b.add<uint8_t>(1, is_signed ? 1 : 0);
    *type_type = T_Int;
    *type_off = b.end_table();
    return true;
  };
  auto empty = [&](uint8_t t) {
    b.start_table(0);
    *type_type = t;
    *type_off = b.end_table();
    return true;
  };
  if (f == "c") return int_type(8, true);
  if (f == "C") return int_type(8, false);
  if (f == "s") return int_type(16, true);
  if (f == "S") return int_type(16, false);
  if (f == "i") return int_type(32, true);
  if (f == "I") return int_type(32, false);
  if (f == "l") return int_type(64, true);
  if (f == "L") return int_type(64, false);
  if (f == "e" || f == "f" || f == "g") {
    b.start_table(1);
    b.add<int16_t>(0, f == "e" ? 0 : f == "f" ? 1 : 2);
    *type_type = T_FloatingPoint;
    *type_off = b.end_table();
    return true;
  }
  if (f == "u") return empty(T_Utf8);
  if (f == "z") return empty(T_Binary);
  if (f == "b") return empty(T_Bool);
  if (f.rfind("d:", 0) == 0) {
    int p = 0, s = 0, bits = 128;
    if (sscanf(f.c_str(), "d:%d,%d,%d", &p, &s, &bits) < 2 || bits != 128) return false;
    b.start_table(3);
    b.add<int32_t>(0, p);
    b.add<int32_t>(1, s);
    b.add<int32_t>(2, 128);
    *type_type = T_Decimal;
    *type_off = b.end_table();
    return true;
  }
  if (f == "tdD" || f == "tdm") {
    b.start_table(1);
    b.add<int16_t>(0, f == "tdD" ? 0 : 1);
    *type_type = T_Date;
    *type_off = b.end_table();
    return true;
  }
  if (f.size() == 3 && f.rfind("tt", 0) == 0) {
    const char u = f[2];
    const int unit = u == 's' ? 0 : u == 'm' ? 1 : u == 'u' ? 2 : u == 'n' ? 3 : -1;
    if (unit < 0) return false;
    b.start_table(2);
    b.add<int16_t>(0, (int16_t)unit);
    b.add<int32_t>(1, unit <= 1 ? 32 : 64);
    *type_type = T_Time;
    *type_off = b.end_table();
    return true;
  }
  if (f.size() >= 4 && f.rfind("ts", 0) == 0 && f[3] == ':') {
    const char u = f[2];
    const int unit = u == 's' ? 0 : u == 'm' ? 1 : u == 'u' ? 2 : u == 'n' ? 3 : -1;
    if (unit < 0) return false;
    const std::string tz = f.substr(4);
    uint32_t tz_off = 0;
    if (!tz.empty()) tz_off = b.string(tz);
    b.start_table(2);
    b.add<int16_t>(0, (int16_t)unit);
    if (tz_off) b.add_ref(1, tz_off);
    *type_type = T_Timestamp;
    *type_off = b.end_table();
    return true;
  }
  return false;
}

// continuation marker + metadata length + flatbuffer, padded so that the body starts 8-byte aligned
void wrap_message(const Fbb& b, std::vector<uint8_t>* out) {
  const uint32_t meta = (b.size() + 7u) & ~7u;
  out->assign(8 + (size_t)meta, 0);
  const uint32_t cont = 0xffffffffu;
  memcpy(out->data(), &cont, 4);
  memcpy(out->data() + 4, &meta, 4);
  memcpy(out->data() + 8, b.data(), b.size());
}

int schema_message(bgpu_ctx* ctx, const std::vector<Column>& cols, std::vector<uint8_t>* out) {
  Fbb b;
  std::vector<uint32_t> fields;
  int64_t next_dict_id = 0;
  for (const Column& c : cols) {
    // a dictionary-encoded field carries the type of its VALUES and a DictionaryEncoding{id, indexType Int32} (Schema.fbs);
    // ids are the position among the dictionary fields, as bgpu_ipc_encode_partitions numbers its dictionary batches
    const bool is_dict = c.kind == K_DICT32;
    if (is_dict && (!c.dict || c.dict->kind == K_DICT32)) return ctx->fail(BGPU_EUNSUPPORTED, "IPC encoder: dictionary column " + c.name);
    const std::string& fmt = is_dict ? c.dict->format : c.format;
    uint8_t tt = 0;
    uint32_t toff = 0;
    if (!fb_type(b, fmt, &tt, &toff)) return ctx->fail(BGPU_EUNSUPPORTED, "IPC encoder: type '" + fmt + "' of column " + c.name);
    uint32_t denc = 0;
    if (is_dict) {
      b.start_table(2);
      b.add<int32_t>(0, 32);
      b.add<uint8_t>(1, 1);
      const uint32_t index_type = b.end_table();
      b.start_table(4);
      b.add<int64_t>(0, next_dict_id++);
      b.add_ref(1, index_type);
      b.add<uint8_t>(2, 0);
      denc = b.end_table();
    }
    const uint32_t children = b.offset_vector({});
    const uint32_t name = b.string(c.name);
    b.start_table(7);
    b.add_ref(0, name);
    b.add<uint8_t>(1, (c.flags & ARROW_FLAG_NULLABLE) ? 1 : 0);
    b.add<uint8_t>(2, tt);
    b.add_ref(3, toff);
    if (denc) b.add_ref(4, denc);
    b.add_ref(5, children);
    fields.push_back(b.end_table());
  }
  const uint32_t fv = b.offset_vector(fields);
  b.start_table(4);
  b.add<int16_t>(0, 0);  // little endian
  b.add_ref(1, fv);
  const uint32_t schema = b.end_table();
  b.start_table(5);
  b.add<int16_t>(0, kMetadataV5);
  b.add<uint8_t>(1, H_Schema);
  b.add_ref(2, schema);
  b.add<int64_t>(3, 0);
  b.finish(b.end_table());
  wrap_message(b, out);
  return BGPU_OK;
}

void record_batch_message(int64_t rows, const std::vector<std::pair<int64_t, int64_t>>& nodes,
                          const std::vector<std::pair<int64_t, int64_t>>& buffers, int64_t body_len, std::vector<uint8_t>* out,
                          int compression = -1, int64_t dict_id = -1) {
  Fbb b;
  uint32_t comp = 0;
  if (compression >= 0) {  // BodyCompression{codec: LZ4_FRAME = 0 | ZSTD = 1, method: BUFFER = 0}
    b.start_table(2);
    b.add<int8_t>(0, (int8_t)compression);
    b.add<int8_t>(1, 0);
    comp = b.end_table();
  }
  const uint32_t bv = b.struct_vector(buffers);
  const uint32_t nv = b.struct_vector(nodes);
  b.start_table(5);
  b.add<int64_t>(0, rows);
  b.add_ref(1, nv);
  b.add_ref(2, bv);
  if (comp) b.add_ref(3, comp);
  uint32_t rb = b.end_table();
  uint8_t header = H_RecordBatch;
  if (dict_id >= 0) {  // DictionaryBatch{id, data: RecordBatch, isDelta = false}: the batch's one column is the dictionary
    b.start_table(3);
    b.add<int64_t>(0, dict_id);
    b.add_ref(1, rb);
    b.add<uint8_t>(2, 0);
    rb = b.end_table();
    header = H_DictionaryBatch;
  }
  b.start_table(5);
  b.add<int16_t>(0, kMetadataV5);
  b.add<uint8_t>(1, header);
  b.add_ref(2, rb);
  b.add<int64_t>(3, body_len);
  b.finish(b.end_table());
  wrap_message(b, out);
}


// ---- device side: the bodies of all messages in one pass ------------------------------------------------------------
enum JobKind : int32_t { J_RAW = 0, J_OFFSETS = 1, J_BITS = 2 };
struct Job {
  const uint8_t* src;  // J_RAW: first byte; J_OFFSETS: the partition's first offset; J_BITS: the bitmap (bit 0 of the batch)
  uint64_t dst;        // byte offset in the staging buffer (8-byte aligned)
  uint64_t n;          // J_RAW: bytes; J_OFFSETS: entries; J_BITS: bits
  uint64_t bit0;       // J_BITS: index of the first bit
  int32_t kind;
  int32_t counter;     // J_BITS: index of the set-bit counter to add to, -1 = none
};
constexpr int kIpcThreads = 256;
constexpr uint64_t kIpcChunk = 64 * 1024;  // bytes (J_BITS: output bytes) per CTA step

__global__ void __launch_bounds__(kIpcThreads) ipc_body_kernel(const Job* __restrict__ jobs, uint8_t* __restrict__ stage,
                                                               unsigned long long* __restrict__ set_bits) {
  const Job j = jobs[blockIdx.y];
  uint8_t* dst = stage + j.dst;
  if (j.kind == J_RAW) {
    const uint64_t a = (uint64_t)j.src;
    if ((a & 7) == 0) {
      const uint64_t words = j.n >> 3;
      for (uint64_t i = (uint64_t)blockIdx.x * kIpcThreads + threadIdx.x; i < words; i += (uint64_t)gridDim.x * kIpcThreads)
        ((uint64_t*)dst)[i] = ((const uint64_t*)j.src)[i];
      if (blockIdx.x == 0 && threadIdx.x < (j.n & 7)) dst[(words << 3) + threadIdx.x] = j.src[(words << 3) + threadIdx.x];
    } else if ((a & 3) == 0) {
      const uint64_t words = j.n >> 2;
      for (uint64_t i = (uint64_t)blockIdx.x * kIpcThreads + threadIdx.x; i < words; i += (uint64_t)gridDim.x * kIpcThreads)
        ((uint32_t*)dst)[i] = ((const uint32_t*)j.src)[i];
      if (blockIdx.x == 0 && threadIdx.x < (j.n & 3)) dst[(words << 2) + threadIdx.x] = j.src[(words << 2) + threadIdx.x];
    } else {  // arbitrary byte start (utf8 data of a partition): aligned 8-byte source words, funnel-shifted
      const uint64_t sh = (a & 7) * 8;
      const uint64_t* w = (const uint64_t*)(a & ~7ULL);
      const uint64_t words = j.n >> 3;
      for (uint64_t i = (uint64_t)blockIdx.x * kIpcThreads + threadIdx.x; i < words; i += (uint64_t)gridDim.x * kIpcThreads)
        ((uint64_t*)dst)[i] = (w[i] >> sh) | (w[i + 1] << (64 - sh));  // w[i + 1] holds bytes of this range: in bounds
      if (blockIdx.x == 0 && threadIdx.x < (j.n & 7)) dst[(words << 3) + threadIdx.x] = j.src[(words << 3) + threadIdx.x];
    }
  } else if (j.kind == J_OFFSETS) {
    const int32_t* src = (const int32_t*)j.src;
    const int32_t base = src[0];
    for (uint64_t i = (uint64_t)blockIdx.x * kIpcThreads + threadIdx.x; i < j.n; i += (uint64_t)gridDim.x * kIpcThreads)
      ((int32_t*)dst)[i] = src[i] - base;
  } else {
    const uint64_t out_bytes = (j.n + 7) >> 3;
    const uint32_t sh = (uint32_t)(j.bit0 & 7);
    const uint8_t* src = j.src + (j.bit0 >> 3);
    const uint64_t src_bytes = (sh + j.n + 7) >> 3;  // bytes of the source that hold bits of this range
    unsigned long long set = 0;
    for (uint64_t i = (uint64_t)blockIdx.x * kIpcThreads + threadIdx.x; i < out_bytes; i += (uint64_t)gridDim.x * kIpcThreads) {
      uint32_t v = src[i] >> sh;
      if (sh && i + 1 < src_bytes) v |= (uint32_t)src[i + 1] << (8 - sh);
      v &= 0xffu;
      const uint64_t left = j.n - (i << 3);
      if (left < 8) v &= (1u << left) - 1;
      dst[i] = (uint8_t)v;
      set += __popc(v);
    }
    if (j.counter >= 0) {
#pragma unroll
      for (int d = 16; d > 0; d >>= 1) set += __shfl_down_sync(0xffffffffu, set, d);
      if ((threadIdx.x & 31) == 0 && set) atomicAdd(&set_bits[j.counter], set);
    }
  }
}

// ---- LZ4_FRAME compression of the message bodies (what the reference's writer applies per buffer,
// /root/reference/ballista/core/src/execution_plans/shuffle_writer.rs:365-376) -------------------------------------------
// One warp compresses one block of at most 64 KB into an LZ4 block (block-independent frames, as lz4_flex writes them):
// the 32 lanes test 32 consecutive positions at once against a 4096-entry hash table of earlier positions in shared
// memory, the first lane with a >= 4-byte match wins, the match is extended 32 bytes per step, and the sequence (token,
// literal run, offset, match length) is written out - header bytes by lane 0, literals strided over the lanes. Greedy,
// no lazy matching: the ratio is a little below the CPU codec's, the format is the same.
constexpr int kLzcWarps = 4;
constexpr uint32_t kLzcBlock = 64 * 1024, kLzcHashBits = 12;
constexpr uint32_t kLzcBound = kLzcBlock + kLzcBlock / 255 + 16;
struct LzcJob {
  const uint8_t* src;
  uint8_t* dst;  // kLzcBound bytes
  uint32_t n;
  uint32_t pad;
};
// 4 bytes at any alignment: the two aligned words around them, funnel-shifted (reads up to 3 bytes before and 7 bytes
// after p: the staging buffer is allocated with slack for that)
__device__ __forceinline__ uint32_t lzc_load4(const uint8_t* p) {
  const uintptr_t a = (uintptr_t)p;
  const uint32_t* w = (const uint32_t*)(a & ~(uintptr_t)3);
  const uint32_t sh = (uint32_t)(a & 3) * 8;
  const uint32_t w0 = w[0];
  return sh ? __funnelshift_r(w0, w[1], sh) : w0;
}
// bytes a length of `v` adds after the token (v >= 15: 1 + (v - 15) / 255)
__device__ __forceinline__ uint32_t lzc_len_bytes(uint32_t v) { return v >= 15 ? 1 + (v - 15) / 255 : 0; }
__device__ __forceinline__ uint32_t lzc_put_len(uint8_t* dst, uint32_t op, uint32_t v) {  // one lane
  if (v < 15) return op;
  v -= 15;
  while (v >= 255) { dst[op++] = 255; v -= 255; }
  dst[op++] = (uint8_t)v;
  return op;
}
__global__ void __launch_bounds__(kLzcWarps * 32) lz4_compress_kernel(const LzcJob* __restrict__ jobs, uint32_t n_jobs,
                                                                     uint32_t* __restrict__ out_sizes) {
  __shared__ uint16_t s_table[kLzcWarps][1u << kLzcHashBits];
  const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31, job = blockIdx.x * kLzcWarps + warp;
  if (job >= n_jobs) return;
  const LzcJob j = jobs[job];
  const uint8_t* __restrict__ src = j.src;
  uint8_t* dst = j.dst;
  const uint32_t n = j.n;
  uint16_t* table = s_table[warp];
  for (uint32_t i = lane; i < (1u << kLzcHashBits); i += 32) table[i] = 0xffffu;
  __syncwarp();
  uint32_t op = 0, anchor = 0, pos = 0;
  // a match must start at least 12 bytes before the end of the block and may not cover its last 5 bytes
  const uint32_t match_start_limit = n >= 13 ? n - 12 : 0, match_end_limit = n >= 5 ? n - 5 : 0;
  while (pos < match_start_limit) {
    const uint32_t p = pos + lane;
    const bool in_range = p < match_start_limit;
    uint32_t v = 0, h = 0, cand = 0xffffu;
    if (in_range) {
      v = lzc_load4(src + p);
      h = (v * 2654435761u) >> (32 - kLzcHashBits);
      cand = table[h];
    }
    __syncwarp();
    if (in_range) table[h] = (uint16_t)p;  // lanes with one hash: any winner will do
    __syncwarp();
    const bool hit = in_range && cand != 0xffffu && cand < p && lzc_load4(src + cand) == v;
    const uint32_t hits = __ballot_sync(0xffffffffu, hit);
    if (!hits) {
      pos += 32;
      continue;
    }
    const int first = __ffs(hits) - 1;
    const uint32_t mp = pos + first, mc = __shfl_sync(0xffffffffu, cand, first);
    // extend the match, 32 bytes per step
    const uint32_t max_len = match_end_limit - mp;  // mp < n - 12, so max_len >= 8
    uint32_t len = 4;
    while (len < max_len) {  // 4 bytes per lane, 128 per step
      const uint32_t i = len + 4 * lane;
      const uint32_t room = i < max_len ? (max_len - i < 4 ? max_len - i : 4) : 0;  // bytes of this lane's word that may match
      uint32_t same = 0;
      if (room) {
        const uint32_t x = lzc_load4(src + mp + i) ^ lzc_load4(src + mc + i);
        same = x ? (uint32_t)(__ffs(x) - 1) >> 3 : 4;
        if (same > room) same = room;
      }
      const uint32_t full = __ballot_sync(0xffffffffu, same == 4);
      if (full == 0xffffffffu) {
        len += 128;
        continue;
      }
      const int stop = __ffs(~full) - 1;  // first lane whose word did not match completely
      len += 4 * (uint32_t)stop + __shfl_sync(0xffffffffu, same, stop);
      break;
    }
    if (len > max_len) len = max_len;
    // the sequence: token | literal length | literals | offset | match length
    const uint32_t lit = mp - anchor, ml = len - 4;
    const uint32_t lit_at = op + 1 + lzc_len_bytes(lit);
    if (lane == 0) {
      dst[op] = (uint8_t)(((lit < 15 ? lit : 15) << 4) | (ml < 15 ? ml : 15));
      lzc_put_len(dst, op + 1, lit);
      const uint32_t off_at = lit_at + lit, offset = mp - mc;
      dst[off_at] = (uint8_t)offset;
      dst[off_at + 1] = (uint8_t)(offset >> 8);
      lzc_put_len(dst, off_at + 2, ml);
    }
    for (uint32_t i = lane; i < lit; i += 32) dst[lit_at + i] = src[anchor + i];
    op = lit_at + lit + 2 + lzc_len_bytes(ml);
    pos = mp + len;
    anchor = pos;
    __syncwarp();
  }
  {  // the last sequence: literals only
    const uint32_t lit = n - anchor, lit_at = op + 1 + lzc_len_bytes(lit);
    if (lane == 0) {
      dst[op] = (uint8_t)((lit < 15 ? lit : 15) << 4);
      lzc_put_len(dst, op + 1, lit);
    }
    for (uint32_t i = lane; i < lit; i += 32) dst[lit_at + i] = src[anchor + i];
    op = lit_at + lit;
  }
  if (lane == 0) out_sizes[job] = op;
}

// byte ranges to byte positions (compressed blocks into their place in the frames)
struct PackJob { const uint8_t* src; uint64_t dst; uint32_t n; uint32_t pad; };
__global__ void __launch_bounds__(kIpcThreads) ipc_pack_kernel(const PackJob* __restrict__ jobs, uint8_t* __restrict__ out) {
  const PackJob j = jobs[blockIdx.x];
  uint8_t* dst = out + j.dst;
  // destination-aligned 4-byte stores fed by bytes; ragged edges by bytes
  const uint32_t head = min(j.n, (uint32_t)((4 - ((uint64_t)dst & 3)) & 3));
  for (uint32_t i = threadIdx.x; i < head; i += kIpcThreads) dst[i] = j.src[i];
  const uint32_t words = (j.n - head) >> 2;
  for (uint32_t w = threadIdx.x; w < words; w += kIpcThreads) {
    const uint8_t* s = j.src + head + 4 * w;
    ((uint32_t*)(dst + head))[w] = (uint32_t)s[0] | ((uint32_t)s[1] << 8) | ((uint32_t)s[2] << 16) | ((uint32_t)s[3] << 24);
  }
  for (uint32_t i = head + 4 * words + threadIdx.x; i < j.n; i += kIpcThreads) dst[i] = j.src[i];
}

__global__ void gather_i32_kernel(const int32_t* __restrict__ src, const uint64_t* __restrict__ idx, int32_t* __restrict__ out, uint32_t n) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = src[idx[i]];
}

inline uint64_t pad8(uint64_t x) { return (x + 7) & ~7ULL; }

// Pinned host buffers for the encoded messages are recycled process-wide: cudaHostAlloc of a few hundred MB costs tens
// of milliseconds, more than laying the bodies out and copying them. bgpu_ipc_free(p, 1) returns a buffer to the pool.
struct PinnedPool {
  std::mutex mu;
  std::vector<std::pair<uint8_t*, size_t>> free_list;   // idle buffers
  std::vector<std::pair<uint8_t*, size_t>> lent;        // handed out, by pointer
  static constexpr size_t kKeepBytes = 1ULL << 30;
  cudaError_t take(size_t bytes, uint8_t** out) {
    std::lock_guard<std::mutex> lock(mu);
    size_t best = free_list.size();
    for (size_t i = 0; i < free_list.size(); i++)
      if (free_list[i].second >= bytes && (best == free_list.size() || free_list[i].second < free_list[best].second)) best = i;
    if (best != free_list.size() && free_list[best].second <= 4 * bytes + (1 << 20)) {
      *out = free_list[best].first;
      lent.push_back(free_list[best]);
      free_list.erase(free_list.begin() + (long)best);
      return cudaSuccess;
    }
    const size_t cap = (bytes + (1 << 20) - 1) & ~(size_t)((1 << 20) - 1);
    const cudaError_t e = cudaHostAlloc((void**)out, cap, cudaHostAllocDefault);
    if (e == cudaSuccess) lent.emplace_back(*out, cap);
    return e;
  }
  void give(uint8_t* p) {
    std::lock_guard<std::mutex> lock(mu);
    for (size_t i = 0; i < lent.size(); i++)
      if (lent[i].first == p) {
        free_list.push_back(lent[i]);
        lent.erase(lent.begin() + (long)i);
        size_t total = 0;
        for (auto& f : free_list) total += f.second;
        while (total > kKeepBytes && !free_list.empty()) {  // keep at most 1 GiB idle: drop the oldest
          total -= free_list.front().second;
          cudaFreeHost(free_list.front().first);
          free_list.erase(free_list.begin());
        }
        return;
      }
    cudaFreeHost(p);  // not ours (should not happen)
  }
};
PinnedPool& pinned_pool() {
  static PinnedPool* p = new PinnedPool();  // leaked on purpose: buffers may be returned during interpreter shutdown
  return *p;
}

}  // namespace
}  // namespace bgpu

using namespace bgpu;

static int to_malloc(bgpu_ctx* ctx, const std::vector<uint8_t>& msg, uint8_t** out_bytes, uint64_t* out_len) {
  uint8_t* p = (uint8_t*)malloc(msg.size());
  if (!p) return ctx->fail(BGPU_EINVAL, "out of host memory");
  memcpy(p, msg.data(), msg.size());
  *out_bytes = p;
  *out_len = msg.size();
  return BGPU_OK;
}

extern "C" int bgpu_ipc_schema_message(bgpu_ctx* ctx, const struct ArrowSchema* schema, uint8_t** out_bytes, uint64_t* out_len) try {
  if (!ctx || !schema || !out_bytes || !out_len) return BGPU_EINVAL;
  std::vector<Column> cols((size_t)schema->n_children);
  for (int64_t i = 0; i < schema->n_children; i++) BGPU_TRY(parse_field(ctx, schema->children[i], &cols[(size_t)i]));
  std::vector<uint8_t> msg;
  BGPU_TRY(schema_message(ctx, cols, &msg));
  return to_malloc(ctx, msg, out_bytes, out_len);
} catch (const std::exception& e) {
  return bgpu::abi_guard(ctx, e);  // nothing is thrown across the C ABI
}

extern "C" int bgpu_ipc_record_batch_metadata(bgpu_ctx* ctx, int64_t rows, const int64_t* nodes, int32_t n_nodes, const int64_t* buffers,
                                              int32_t n_buffers, int64_t body_len, int32_t compression, uint8_t** out_bytes, uint64_t* out_len) try {
  if (!ctx || !out_bytes || !out_len || n_nodes < 0 || n_buffers < 0 || (n_nodes && !nodes) || (n_buffers && !buffers)) return BGPU_EINVAL;
  if (compression < -1 || compression > 0) return ctx->fail(BGPU_EUNSUPPORTED, "compression must be -1 (none) or 0 (LZ4_FRAME)");
  std::vector<std::pair<int64_t, int64_t>> nv, bv;
  for (int32_t i = 0; i < n_nodes; i++) nv.emplace_back(nodes[2 * i], nodes[2 * i + 1]);
  for (int32_t i = 0; i < n_buffers; i++) bv.emplace_back(buffers[2 * i], buffers[2 * i + 1]);
  std::vector<uint8_t> msg;
  record_batch_message(rows, nv, bv, body_len, &msg, compression);
  return to_malloc(ctx, msg, out_bytes, out_len);
} catch (const std::exception& e) {
  return bgpu::abi_guard(ctx, e);  // nothing is thrown across the C ABI
}

// One encapsulated message per non-empty partition of a batch WITHOUT dictionary columns: RecordBatch messages, or
// (dict_id >= 0, one column) DictionaryBatch messages whose body is that column.
static int encode_impl(bgpu_ctx* ctx, const bgpu_batch* batch, const uint64_t* offsets, uint32_t n, int32_t compression,
                       int64_t dict_id, uint8_t** out_bytes, uint64_t* out_msg_off) {
  *out_bytes = nullptr;
  const size_t n_cols = batch->cols.size();
  std::unique_ptr<GpuTimeScope> timed(new GpuTimeScope(ctx, T_KERNEL));  // framing + compression; the copy-out is d2h_time

  // 1. utf8 columns: the byte offsets at the partition boundaries (one small gather, one sync)
  std::vector<size_t> utf8_cols;
  for (size_t i = 0; i < n_cols; i++)
    if (batch->cols[i].kind == K_UTF8 || batch->cols[i].kind == K_BINARY) utf8_cols.push_back(i);
  std::vector<int32_t> bounds(utf8_cols.size() * (size_t)(n + 1), 0);  // [col][p]
  if (!utf8_cols.empty() && batch->n_rows > 0) {
    DevAllocPtr d_idx, d_out;
    BGPU_TRY(dev_alloc(ctx, (size_t)(n + 1) * 8, &d_idx));
    BGPU_TRY(dev_alloc(ctx, bounds.size() * 4, &d_out));
    BGPU_CUDA(ctx, cudaMemcpyAsync(d_idx->p, offsets, (size_t)(n + 1) * 8, cudaMemcpyHostToDevice, ctx->stream));
    for (size_t u = 0; u < utf8_cols.size(); u++) {
      gather_i32_kernel<<<(n + 1 + 255) / 256, 256, 0, ctx->stream>>>(batch->cols[utf8_cols[u]].offsets, (const uint64_t*)d_idx->p,
                                                                    (int32_t*)d_out->p + u * (n + 1), n + 1);
      ctx->launches++;
    }
    BGPU_CUDA(ctx, cudaMemcpyAsync(bounds.data(), d_out->p, bounds.size() * 4, cudaMemcpyDeviceToHost, ctx->stream));
    BGPU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  }

  // 2. layout: message p = [8-byte prefix + metadata (meta_len, the same for every partition)] [body]
  size_t n_buffers = 0;
  for (const Column& c : batch->cols) n_buffers += (c.kind == K_UTF8 || c.kind == K_BINARY) ? 3 : 2;
  std::vector<uint8_t> msg;
  record_batch_message(0, std::vector<std::pair<int64_t, int64_t>>(n_cols), std::vector<std::pair<int64_t, int64_t>>(n_buffers), 0, &msg,
                       compression, dict_id);
  const uint64_t meta_len = msg.size();
  struct Part { std::vector<std::pair<int64_t, int64_t>> buffers; std::vector<int> counter; uint64_t body_len = 0, body0 = 0; };
  std::vector<Part> parts(n);
  std::vector<Job> jobs;
  uint64_t total = 0;
  int n_counters = 0;
  for (uint32_t p = 0; p < n; p++) {
    out_msg_off[p] = total;
    const uint64_t lo = offsets[p], hi = offsets[p + 1];
    if (hi < lo) return ctx->fail(BGPU_EINVAL, "partition offsets must be non-decreasing");
    const uint64_t rows = hi - lo;
    if (rows == 0) continue;  // the writer never sees an empty partition (shuffle_writer.rs:349)
    Part& pt = parts[p];
    const uint64_t body0 = total + meta_len;
    uint64_t body = 0;
    auto buffer = [&](uint64_t len) { pt.buffers.emplace_back((int64_t)body, (int64_t)len); const uint64_t at = body0 + body; body += pad8(len); return at; };
    pt.body0 = body0;
    size_t u = 0;
    for (size_t i = 0; i < n_cols; i++) {
      const Column& c = batch->cols[i];
      if (c.validity) {
        const uint64_t at = buffer((rows + 7) / 8);
        jobs.push_back(Job{c.validity, at, rows, lo, J_BITS, n_counters});
        pt.counter.push_back(n_counters++);
      } else {
        buffer(0);
        pt.counter.push_back(-1);
      }
      if (c.kind == K_UTF8 || c.kind == K_BINARY) {
        const int32_t b0 = bounds[u * (n + 1) + p], b1 = bounds[u * (n + 1) + p + 1];
        u++;
        uint64_t at = buffer((rows + 1) * 4);
        jobs.push_back(Job{(const uint8_t*)(c.offsets + lo), at, rows + 1, 0, J_OFFSETS, -1});
        at = buffer((uint64_t)(b1 - b0));
        if (b1 > b0) jobs.push_back(Job{(const uint8_t*)c.values + b0, at, (uint64_t)(b1 - b0), 0, J_RAW, -1});
      } else if (c.kind == K_BOOL) {
        const uint64_t at = buffer((rows + 7) / 8);
        jobs.push_back(Job{(const uint8_t*)c.values, at, rows, lo, J_BITS, -1});
      } else {
        const uint64_t at = buffer(rows * (uint64_t)c.width);
        jobs.push_back(Job{(const uint8_t*)c.values + lo * (uint64_t)c.width, at, rows * (uint64_t)c.width, 0, J_RAW, -1});
      }
    }
    pt.body_len = body;
    total = body0 + body;
  }
  out_msg_off[n] = total;
  if (total == 0) return BGPU_OK;

  // 3. one kernel builds every body in a staging buffer that mirrors the final layout; one copy brings it to the host
  DevAllocPtr stage, d_jobs, d_cnt;
  BGPU_TRY(dev_alloc(ctx, total + 16, &stage));  // + 16: the LZ4 match search reads whole aligned words around a position
  BGPU_TRY(dev_alloc(ctx, jobs.size() * sizeof(Job), &d_jobs));
  BGPU_TRY(dev_alloc(ctx, (size_t)std::max(1, n_counters) * 8, &d_cnt));
  BGPU_CUDA(ctx, cudaMemsetAsync(stage->p, 0, total, ctx->stream));  // buffer padding is zero
  BGPU_CUDA(ctx, cudaMemsetAsync(d_cnt->p, 0, (size_t)std::max(1, n_counters) * 8, ctx->stream));
  BGPU_CUDA(ctx, cudaMemcpyAsync(d_jobs->p, jobs.data(), jobs.size() * sizeof(Job), cudaMemcpyHostToDevice, ctx->stream));
  uint64_t biggest = 0;
  for (const Job& j : jobs) biggest = std::max<uint64_t>(biggest, j.kind == J_BITS ? (j.n + 7) / 8 : j.kind == J_OFFSETS ? j.n * 4 : j.n);
  const uint32_t gx = (uint32_t)std::min<uint64_t>(std::max<uint64_t>(1, (biggest + kIpcChunk - 1) / kIpcChunk),
                                                   std::max<uint64_t>(1, (uint64_t)ctx->sm_count * 8 / jobs.size() + 1));
  for (size_t j0 = 0; j0 < jobs.size(); j0 += 65535) {
    const uint32_t gy = (uint32_t)std::min<size_t>(65535, jobs.size() - j0);
    ipc_body_kernel<<<dim3(gx, gy), kIpcThreads, 0, ctx->stream>>>((const Job*)d_jobs->p + j0, (uint8_t*)stage->p,
                                                                  (unsigned long long*)d_cnt->p);
    ctx->launches++;
  }
  BGPU_CUDA(ctx, cudaGetLastError());
  if (compression == 0) {
    // ---- LZ4_FRAME: compress every buffer of every message on the device, then lay the compressed messages out ----------
    // a. one job per <= 64 KB block of every non-empty buffer
    struct Buf { uint32_t p; size_t idx; uint64_t at, len; size_t job0, n_jobs; };
    std::vector<Buf> bufs;
    std::vector<LzcJob> cjobs;
    for (uint32_t p = 0; p < n; p++)
      for (size_t i = 0; i < parts[p].buffers.size(); i++) {
        const uint64_t len = (uint64_t)parts[p].buffers[i].second;
        if (!len) continue;
        Buf b{p, i, parts[p].body0 + (uint64_t)parts[p].buffers[i].first, len, cjobs.size(), (size_t)((len + kLzcBlock - 1) / kLzcBlock)};
        for (size_t k = 0; k < b.n_jobs; k++)
          cjobs.push_back(LzcJob{(const uint8_t*)stage->p + b.at + k * kLzcBlock, nullptr, (uint32_t)std::min<uint64_t>(kLzcBlock, len - k * kLzcBlock), 0});
        bufs.push_back(b);
      }
    DevAllocPtr tmp, d_cjobs, d_sizes;
    BGPU_TRY(dev_alloc(ctx, cjobs.size() * (size_t)kLzcBound, &tmp));
    for (size_t k = 0; k < cjobs.size(); k++) cjobs[k].dst = (uint8_t*)tmp->p + k * (size_t)kLzcBound;
    BGPU_TRY(dev_alloc(ctx, cjobs.size() * sizeof(LzcJob), &d_cjobs));
    BGPU_TRY(dev_alloc(ctx, cjobs.size() * 4, &d_sizes));
    BGPU_CUDA(ctx, cudaMemcpyAsync(d_cjobs->p, cjobs.data(), cjobs.size() * sizeof(LzcJob), cudaMemcpyHostToDevice, ctx->stream));
    lz4_compress_kernel<<<(unsigned)((cjobs.size() + kLzcWarps - 1) / kLzcWarps), kLzcWarps * 32, 0, ctx->stream>>>(
        (const LzcJob*)d_cjobs->p, (uint32_t)cjobs.size(), (uint32_t*)d_sizes->p);
    ctx->launches++;
    BGPU_CUDA(ctx, cudaGetLastError());
    std::vector<uint32_t> sizes(cjobs.size());
    std::vector<unsigned long long> set_bits((size_t)std::max(1, n_counters), 0);
    BGPU_CUDA(ctx, cudaMemcpyAsync(sizes.data(), d_sizes->p, sizes.size() * 4, cudaMemcpyDeviceToHost, ctx->stream));
    BGPU_CUDA(ctx, cudaMemcpyAsync(set_bits.data(), d_cnt->p, set_bits.size() * 8, cudaMemcpyDeviceToHost, ctx->stream));
    BGPU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    // b. compressed layout. A buffer becomes [int64 raw length][frame: magic FLG BD HC | (size, block)* | end mark], or
    //    [-1][raw bytes] when the frame would not be smaller (as arrow-ipc does); a block that did not shrink is stored.
    struct Word { uint64_t at; uint32_t v; };   // 4-byte words the host writes after the copy (sizes, marks)
    std::vector<Word> words;
    std::vector<std::pair<uint64_t, int64_t>> prefixes;  // (at, raw length or -1)
    std::vector<PackJob> pjobs;
    std::vector<std::vector<std::pair<int64_t, int64_t>>> cbuffers(n);
    std::vector<uint64_t> cbody_len(n, 0), cmsg_off(n + 1, 0);
    uint64_t ctotal = 0;
    size_t bi = 0;
    constexpr uint32_t kMagic = 0x184D2204u;
    for (uint32_t p = 0; p < n; p++) {
      cmsg_off[p] = ctotal;
      if (offsets[p + 1] == offsets[p]) continue;
      const uint64_t body0 = ctotal + meta_len;
      uint64_t body = 0;
      for (size_t i = 0; i < parts[p].buffers.size(); i++) {
        if (!parts[p].buffers[i].second) {
          cbuffers[p].emplace_back((int64_t)body, 0);
          continue;
        }
        const Buf& b = bufs[bi++];
        uint64_t frame = 7 + 4;  // header + end mark
        for (size_t k = 0; k < b.n_jobs; k++) frame += 4 + std::min<uint32_t>(sizes[b.job0 + k], cjobs[b.job0 + k].n);
        const uint64_t at = body0 + body;
        uint64_t enc;
        if (frame >= b.len) {  // stored
          prefixes.emplace_back(at, -1);
          for (size_t k = 0; k < b.n_jobs; k++)
            pjobs.push_back(PackJob{cjobs[b.job0 + k].src, at + 8 + k * (uint64_t)kLzcBlock, cjobs[b.job0 + k].n, 0});
          enc = 8 + b.len;
        } else {
          prefixes.emplace_back(at, (int64_t)b.len);
          uint64_t q = at + 8;
          words.push_back(Word{q, kMagic});
          // FLG = version 01, block-independent; BD = 64 KB blocks; HC = (xxh32(FLG BD) >> 8) & 0xff = 0x82
          words.push_back(Word{q + 4, 0x60u | (0x40u << 8) | (0x82u << 16)});  // the 4th byte is overwritten by the first block size
          q += 7;
          for (size_t k = 0; k < b.n_jobs; k++) {
            const LzcJob& cj = cjobs[b.job0 + k];
            const bool stored = sizes[b.job0 + k] >= cj.n;
            const uint32_t sz = stored ? cj.n : sizes[b.job0 + k];
            words.push_back(Word{q, sz | (stored ? 0x80000000u : 0u)});
            pjobs.push_back(PackJob{stored ? cj.src : cj.dst, q + 4, sz, 0});
            q += 4 + sz;
          }
          words.push_back(Word{q, 0});  // end mark
          enc = 8 + frame;
        }
        cbuffers[p].emplace_back((int64_t)body, (int64_t)enc);
        body += pad8(enc);
      }
      cbody_len[p] = body;
      ctotal = body0 + body;
    }
    cmsg_off[n] = ctotal;
    // c. pack on the device, one copy to the host, then the small words and the metadata
    DevAllocPtr stage2, d_pjobs;
    BGPU_TRY(dev_alloc(ctx, ctotal, &stage2));
    BGPU_TRY(dev_alloc(ctx, pjobs.size() * sizeof(PackJob), &d_pjobs));
    BGPU_CUDA(ctx, cudaMemsetAsync(stage2->p, 0, ctotal, ctx->stream));
    BGPU_CUDA(ctx, cudaMemcpyAsync(d_pjobs->p, pjobs.data(), pjobs.size() * sizeof(PackJob), cudaMemcpyHostToDevice, ctx->stream));
    ipc_pack_kernel<<<(unsigned)pjobs.size(), kIpcThreads, 0, ctx->stream>>>((const PackJob*)d_pjobs->p, (uint8_t*)stage2->p);
    ctx->launches++;
    BGPU_CUDA(ctx, cudaGetLastError());
    uint8_t* chost = nullptr;
    BGPU_CUDA(ctx, pinned_pool().take(ctotal, &chost));
    timed.reset();
    timed.reset(new GpuTimeScope(ctx, T_D2H));
    cudaError_t ce = cudaMemcpyAsync(chost, stage2->p, ctotal, cudaMemcpyDeviceToHost, ctx->stream);
    timed.reset();
    if (ce == cudaSuccess) ce = cudaStreamSynchronize(ctx->stream);
    if (ce != cudaSuccess) {
      pinned_pool().give(chost);
      return ctx->fail(BGPU_ECUDA, std::string("IPC encode: ") + cudaGetErrorString(ce));
    }
    for (const Word& w : words) {
      memcpy(chost + w.at, &w.v, w.at + 4 <= ctotal ? 4 : ctotal - w.at);
    }
    for (auto& pr : prefixes) memcpy(chost + pr.first, &pr.second, 8);
    for (uint32_t p = 0; p < n; p++) {
      const uint64_t rows = offsets[p + 1] - offsets[p];
      if (rows == 0) continue;
      std::vector<std::pair<int64_t, int64_t>> nodes;
      for (size_t i = 0; i < n_cols; i++)
        nodes.emplace_back((int64_t)rows, parts[p].counter[i] < 0 ? 0 : (int64_t)(rows - set_bits[(size_t)parts[p].counter[i]]));
      record_batch_message((int64_t)rows, nodes, cbuffers[p], (int64_t)cbody_len[p], &msg, 0, dict_id);
      if (msg.size() != meta_len) {
        pinned_pool().give(chost);
        return ctx->fail(BGPU_EINVAL, "internal: IPC metadata size is not constant");
      }
      memcpy(chost + cmsg_off[p], msg.data(), msg.size());
    }
    for (uint32_t p = 0; p <= n; p++) out_msg_off[p] = cmsg_off[p];
    *out_bytes = chost;
    return BGPU_OK;
  }
  uint8_t* host = nullptr;
  BGPU_CUDA(ctx, pinned_pool().take(total, &host));
  std::vector<unsigned long long> set_bits((size_t)std::max(1, n_counters), 0);
  timed.reset();
  timed.reset(new GpuTimeScope(ctx, T_D2H));
  cudaError_t e = cudaMemcpyAsync(host, stage->p, total, cudaMemcpyDeviceToHost, ctx->stream);
  if (e == cudaSuccess) e = cudaMemcpyAsync(set_bits.data(), d_cnt->p, set_bits.size() * 8, cudaMemcpyDeviceToHost, ctx->stream);
  timed.reset();
  if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
  if (e != cudaSuccess) {
    pinned_pool().give(host);
    return ctx->fail(BGPU_ECUDA, std::string("IPC encode: ") + cudaGetErrorString(e));
  }

  // 4. the metadata goes into the holes
  for (uint32_t p = 0; p < n; p++) {
    const uint64_t rows = offsets[p + 1] - offsets[p];
    if (rows == 0) continue;
    const Part& pt = parts[p];
    std::vector<std::pair<int64_t, int64_t>> nodes;
    for (size_t i = 0; i < n_cols; i++)
      nodes.emplace_back((int64_t)rows, pt.counter[i] < 0 ? 0 : (int64_t)(rows - set_bits[(size_t)pt.counter[i]]));
    record_batch_message((int64_t)rows, nodes, pt.buffers, (int64_t)pt.body_len, &msg, -1, dict_id);
    if (msg.size() != meta_len) {
      pinned_pool().give(host);
      return ctx->fail(BGPU_EINVAL, "internal: IPC metadata size is not constant");
    }
    memcpy(host + out_msg_off[p], msg.data(), msg.size());
  }
  *out_bytes = host;
  return BGPU_OK;
}

extern "C" int bgpu_ipc_encode_partitions(bgpu_ctx* ctx, const bgpu_batch* batch, const uint64_t* offsets, uint32_t n,
                                          int32_t compression, uint8_t** out_bytes, uint64_t* out_msg_off) try {
  if (!ctx || !batch || !offsets || !out_bytes || !out_msg_off || n == 0) return BGPU_EINVAL;
  if (compression < -1 || compression > 0) return ctx->fail(BGPU_EUNSUPPORTED, "compression must be -1 (none) or 0 (LZ4_FRAME)");
  *out_bytes = nullptr;
  if (!ctx->has_device) return ctx->fail(BGPU_ENODEVICE, "context has no device (libbgpu has no CPU fallback)");
  BGPU_CUDA(ctx, cudaSetDevice(ctx->device));
  if (offsets[n] > (uint64_t)batch->n_rows) return ctx->fail(BGPU_EINVAL, "partition offsets beyond the batch");
  if (!batch_has_dictionary(batch)) return encode_impl(ctx, batch, offsets, n, compression, -1, out_bytes, out_msg_off);
  // Dictionary columns. A partition's message becomes [DictionaryBatch per dictionary column][RecordBatch]: the
  // partition's VALUES are its dictionary (one entry per row), its indices count 0 .. rows-1 - the stream the reference's
  // reader expects (/root/reference/ballista/core/src/async_reader/mod.rs:168-258: a dictionary batch replaces the
  // dictionary of its id for the record batches that follow).
  std::unique_ptr<bgpu_batch> main;
  std::vector<size_t> dict_cols;
  BGPU_TRY(dict_decode_batch(ctx, batch, &main, &dict_cols));
  struct Piece { uint8_t* bytes = nullptr; std::vector<uint64_t> off; };
  std::vector<Piece> pieces(dict_cols.size() + 1);
  auto cleanup = [&]() { for (Piece& pc : pieces) if (pc.bytes) pinned_pool().give(pc.bytes); };
  DevAllocPtr d_offs;
  BGPU_TRY(dev_alloc(ctx, (size_t)(n + 1) * 8, &d_offs));
  BGPU_CUDA(ctx, cudaMemcpyAsync(d_offs->p, offsets, (size_t)(n + 1) * 8, cudaMemcpyHostToDevice, ctx->stream));
  int rc = BGPU_OK;
  for (size_t k = 0; k < dict_cols.size() && rc == BGPU_OK; k++) {
    bgpu_batch one(ctx);  // the decoded values of dictionary column k, as a one-column batch
    one.n_rows = main->n_rows;
    one.cols.push_back(main->cols[dict_cols[k]]);
    pieces[k].off.assign(n + 1, 0);
    rc = encode_impl(ctx, &one, offsets, n, compression, (int64_t)k, &pieces[k].bytes, pieces[k].off.data());
  }
  if (rc == BGPU_OK) {
    for (size_t i : dict_cols) {
      rc = dict_identity_encode(ctx, main.get(), i, batch->cols[i], (const uint64_t*)d_offs->p, n);
      if (rc != BGPU_OK) break;
      Column& c = main->cols[i];  // encoded as its Int32 indices
      c.kind = K_FIXED; c.format = "i"; c.dict = nullptr;
    }
  }
  if (rc == BGPU_OK) {
    pieces.back().off.assign(n + 1, 0);
    rc = encode_impl(ctx, main.get(), offsets, n, compression, -1, &pieces.back().bytes, pieces.back().off.data());
  }
  if (rc != BGPU_OK) { cleanup(); return rc; }
  uint64_t total = 0;
  for (uint32_t p = 0; p < n; p++) {
    out_msg_off[p] = total;
    for (const Piece& pc : pieces) total += pc.off[p + 1] - pc.off[p];
  }
  out_msg_off[n] = total;
  if (total) {
    uint8_t* host = nullptr;
    const cudaError_t e = pinned_pool().take(total, &host);
    if (e != cudaSuccess) { cleanup(); return ctx->fail(BGPU_ECUDA, std::string("IPC encode: ") + cudaGetErrorString(e)); }
    uint64_t at = 0;
    for (uint32_t p = 0; p < n; p++)
      for (const Piece& pc : pieces) {
        const uint64_t len = pc.off[p + 1] - pc.off[p];
        if (len) memcpy(host + at, pc.bytes + pc.off[p], len);
        at += len;
      }
    *out_bytes = host;
  }
  cleanup();
  return BGPU_OK;
} catch (const std::exception& e) {
  return bgpu::abi_guard(ctx, e);  // nothing is thrown across the C ABI
}

extern "C" void bgpu_ipc_free(uint8_t* bytes, int pinned) {
  if (!bytes) return;
  if (pinned) pinned_pool().give(bytes);
  else free(bytes);
}
