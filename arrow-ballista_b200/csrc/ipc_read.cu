// Arrow IPC stream reader into device batches: the inverse of ipc.cu.
//
// Replaces, for the shuffle read, what ShuffleReaderExec does per fetched partition file with `AsyncStreamReader`
// (/root/reference/ballista/core/src/async_reader/mod.rs:108-258: continuation marker, metadata length, Schema
// message first, then RecordBatch messages until a zero length / EOF) and arrow-ipc's `read_record_batch`, as driven by
// `fetch_partition_local_inner` (/root/reference/ballista/core/src/execution_plans/shuffle_reader.rs:771-785).
//
// Division of labour:
//   host    walks the stream framing and the flatbuffer metadata (a bounds-checked reader, no flatbuffers dependency),
//           and - for LZ4_FRAME bodies as the reference writes them (shuffle_writer.rs:365-376) - the LZ4 frame and
//           block headers, which gives a table of independent decode chains
//   device  uncompressed bodies: one host-to-device copy per buffer into a 256-byte aligned arena (IPC aligns buffers
//           to 8 bytes only; the aggregate kernels load Decimal128 as 16-byte words)
//           LZ4_FRAME bodies: the compressed body is copied once; `lz4_decode_kernel` decodes one chain per warp (a
//           chain = one block of a block-independent frame, or all blocks of a block-linked frame in order)
// Dictionary batches are BGPU_EUNSUPPORTED (as in the encoder); the aggregate states that cross a shuffle carry none.
#include <string.h>

#include <algorithm>
#include <string>
#include <vector>

#include <map>

#include "internal.h"

void export_batch_schema(const bgpu_batch* b, struct ArrowSchema* out);  // arrow_io.cu

namespace bgpu {
namespace {

// ---- bounds-checked flatbuffer reading ----------------------------------------------------------------------------
struct Fb {
  const uint8_t* p = nullptr;
  size_t n = 0;
  bool ok = true;
  template <typename T>
  T rd(size_t at) {
    T v{};
    if (at + sizeof(T) > n || at + sizeof(T) < at) {
      ok = false;
      return v;
    }
    memcpy(&v, p + at, sizeof(T));
    return v;
  }
};
struct Table {
  Fb* fb = nullptr;
  size_t pos = 0;  // 0 = absent
  explicit operator bool() const { return pos != 0; }
  size_t field(int id) const {  // absolute position of field `id`, 0 = absent (default value applies)
    if (!pos) return 0;
    const int32_t so = fb->rd<int32_t>(pos);
    const size_t vt = (size_t)((int64_t)pos - so);
    const uint16_t vsize = fb->rd<uint16_t>(vt);
    const size_t slot = 4 + 2 * (size_t)id;
    if (!fb->ok || slot + 2 > vsize) return 0;
    const uint16_t off = fb->rd<uint16_t>(vt + slot);
    return off ? pos + off : 0;
  }
  template <typename T>
  T scalar(int id, T dflt) const {
    const size_t at = field(id);
    return at ? fb->rd<T>(at) : dflt;
  }
  size_t indirect(int id) const {  // position an offset field points at
    const size_t at = field(id);
    return at ? at + fb->rd<uint32_t>(at) : 0;
  }
  Table table(int id) const { return Table{fb, indirect(id)}; }
  std::string str(int id) const {
    const size_t at = indirect(id);
    if (!at) return "";
    const uint32_t len = fb->rd<uint32_t>(at);
    if (!fb->ok || at + 4 + (size_t)len > fb->n) {
      fb->ok = false;
      return "";
    }
    return std::string((const char*)fb->p + at + 4, len);
  }
  // vector field: element count and position of element 0
  size_t vec(int id, uint32_t* count) const {
    const size_t at = indirect(id);
    *count = at ? fb->rd<uint32_t>(at) : 0;
    // the smallest element (an offset) is 4 bytes: a count the buffer cannot hold is a malformed message, and must not
    // size an allocation
    if (at && (!fb->ok || at + 4 > fb->n || (size_t)*count > (fb->n - at - 4) / 4)) {
      fb->ok = false;
      *count = 0;
    }
    return at ? at + 4 : 0;
  }
};

enum FbType : uint8_t { T_Int = 2, T_FloatingPoint = 3, T_Binary = 4, T_Utf8 = 5, T_Bool = 6, T_Decimal = 7, T_Date = 8, T_Time = 9, T_Timestamp = 10 };
enum FbHeader : uint8_t { H_NONE = 0, H_Schema = 1, H_DictionaryBatch = 2, H_RecordBatch = 3 };

// Field table -> Column descriptor (name, Arrow C format string, kind, width). Schema.fbs defaults apply to absent fields.
// Arrow strings (field names, time zones) are UTF-8; a stream whose strings are not is malformed.
bool valid_utf8(const std::string& s) {
  const size_t n = s.size();
  for (size_t i = 0; i < n;) {
    const uint8_t b = (uint8_t)s[i];
    int extra;
    uint32_t cp;
    if (b < 0x80) { i++; continue; }
    if ((b & 0xe0) == 0xc0) { extra = 1; cp = b & 0x1f; }
    else if ((b & 0xf0) == 0xe0) { extra = 2; cp = b & 0x0f; }
    else if ((b & 0xf8) == 0xf0) { extra = 3; cp = b & 0x07; }
    else return false;
    if (i + (size_t)extra >= n) return false;  // truncated sequence
    for (int k = 1; k <= extra; k++) {
      const uint8_t cb = (uint8_t)s[i + (size_t)k];
      if ((cb & 0xc0) != 0x80) return false;
      cp = (cp << 6) | (cb & 0x3f);
    }
    static const uint32_t kMin[4] = {0, 0x80, 0x800, 0x10000};
    if (cp < kMin[extra] || cp > 0x10ffff || (cp >= 0xd800 && cp <= 0xdfff)) return false;
    i += (size_t)extra + 1;
  }
  return true;
}

int field_type_to_column(bgpu_ctx* ctx, Table f, Column* c);
int field_to_column(bgpu_ctx* ctx, Table f, Column* c) {
  c->name = f.str(0);
  if (!valid_utf8(c->name)) {
    c->name.clear();
    return ctx->fail(BGPU_EINVAL, "IPC reader: field name is not UTF-8");
  }
  c->flags = f.scalar<uint8_t>(1, 0) ? ARROW_FLAG_NULLABLE : 0;
  if (f.field(4)) {
    // DictionaryEncoding{id, indexType (absent = Int32), isOrdered}: the field's type is the type of the VALUES
    const Table enc = f.table(4);
    const Table it = enc.table(1);
    if (it && (it.scalar<int32_t>(0, 0) != 32 || it.scalar<uint8_t>(1, 0) == 0))
      return ctx->fail(BGPU_EUNSUPPORTED, "IPC reader: dictionary indices of column " + c->name + " are not Int32");
    const int64_t id = enc.scalar<int64_t>(0, 0);
      auto values = std::make_shared<Column>();
    BGPU_TRY(field_type_to_column(ctx, f, values.get()));
    values->name = "";
    values->flags = ARROW_FLAG_NULLABLE;
    if (values->kind == K_BOOL) return ctx->fail(BGPU_EUNSUPPORTED, "IPC reader: dictionary of Boolean values (column " + c->name + ")");
    c->format = "i";
    c->kind = K_DICT32;
    c->width = 4;
    c->dict = values;
    c->dict_id = id;
    return BGPU_OK;
  }
  return field_type_to_column(ctx, f, c);
}

// the `type` union of a Field -> format / kind / width
int field_type_to_column(bgpu_ctx* ctx, Table f, Column* c) {
  const uint8_t tt = f.scalar<uint8_t>(2, 0);
  const Table t = f.table(3);
  auto fixed = [&](const std::string& fmt, int w) {
    c->format = fmt;
    c->kind = K_FIXED;
    c->width = w;
    return BGPU_OK;
  };
  switch (tt) {
    case T_Int: {
      const int bits = t.scalar<int32_t>(0, 0);
      const bool sg = t.scalar<uint8_t>(1, 0) != 0;
      switch (bits) {
        case 8: return fixed(sg ? "c" : "C", 1);
        case 16: return fixed(sg ? "s" : "S", 2);
        case 32: return fixed(sg ? "i" : "I", 4);
        case 64: return fixed(sg ? "l" : "L", 8);
        default: break;
      }
      break;
    }
    case T_FloatingPoint: {
      const int prec = t.scalar<int16_t>(0, 0);
      if (prec == 0) return fixed("e", 2);
      if (prec == 1) return fixed("f", 4);
      if (prec == 2) return fixed("g", 8);
      break;
    }
    case T_Binary: c->format = "z"; c->kind = K_BINARY; c->width = 0; return BGPU_OK;
    case T_Utf8: c->format = "u"; c->kind = K_UTF8; c->width = 0; return BGPU_OK;
    case T_Bool: c->format = "b"; c->kind = K_BOOL; c->width = 0; return BGPU_OK;
    case T_Decimal: {
      const int p = t.scalar<int32_t>(0, 0), s = t.scalar<int32_t>(1, 0), bits = t.scalar<int32_t>(2, 128);
      if (bits != 128) break;
      if (p < 1 || p > 38 || s < -38 || s > 38)
        return ctx->fail(BGPU_EINVAL, "IPC reader: Decimal128 precision / scale out of range for column " + c->name);
      return fixed("d:" + std::to_string(p) + "," + std::to_string(s), 16);
    }
    case T_Date: return t.scalar<int16_t>(0, 1) == 0 ? fixed("tdD", 4) : fixed("tdm", 8);
    case T_Time: {
      static const char* u = "smun";
      const int unit = t.scalar<int16_t>(0, 1), bits = t.scalar<int32_t>(1, 32);
      if (unit < 0 || unit > 3) break;
      if (bits != (unit < 2 ? 32 : 64))  // Time32 is seconds / milliseconds, Time64 micro- / nanoseconds
        return ctx->fail(BGPU_EINVAL, "IPC reader: Time unit does not match its bit width for column " + c->name);
      return fixed(std::string("tt") + u[unit], bits == 32 ? 4 : 8);
    }
    case T_Timestamp: {
      static const char* u = "smun";
      const int unit = t.scalar<int16_t>(0, 0);
      if (unit < 0 || unit > 3) break;
      if (!valid_utf8(t.str(1))) return ctx->fail(BGPU_EINVAL, "IPC reader: time zone of column " + c->name + " is not UTF-8");
      return fixed(std::string("ts") + u[unit] + ":" + t.str(1), 8);
    }
    default: break;
  }
  return ctx->fail(BGPU_EUNSUPPORTED, "IPC reader: type id " + std::to_string((int)tt) + " of column " + c->name);
}

// ---- LZ4 frame: host walks the headers, device decodes the blocks ---------------------------------------------------
// One decode step of a chain. kind 0 = LZ4 block, 1 = stored (raw) bytes.
struct Lz4Block {
  uint64_t src;       // byte offset in the (device copy of the) message body
  uint64_t dst;       // byte offset in the decoded arena
  uint32_t src_len;
  uint32_t dst_len;   // bytes this block must produce (checked on the device)
  uint32_t kind;
  uint32_t window;    // bytes of already decoded output before `dst` a match may reach into (linked blocks)
};
struct Lz4Chain {
  uint32_t first, count;  // blocks [first, first + count) decoded in order by one warp
};

// Walks one LZ4 frame (magic, descriptor, blocks, end mark) that must decode to exactly `want` bytes and appends its
// blocks / chains. `at` = offset of the frame in the body, `dst` = where its output goes.
int lz4_frame_plan(bgpu_ctx* ctx, const uint8_t* body, uint64_t at, uint64_t end, uint64_t dst, uint64_t want,
                   std::vector<Lz4Block>* blocks, std::vector<Lz4Chain>* chains) {
  auto bad = [&](const char* what) { return ctx->fail(BGPU_EINVAL, std::string("IPC reader: LZ4 frame: ") + what); };
  auto rd32 = [&](uint64_t o) { uint32_t v; memcpy(&v, body + o, 4); return v; };
  if (at + 7 > end) return bad("truncated header");
  if (rd32(at) != 0x184D2204u) return bad("bad magic");
  const uint8_t flg = body[at + 4], bd = body[at + 5];
  if ((flg >> 6) != 1) return bad("unsupported version");
  const bool independent = flg & 0x20, block_checksum = flg & 0x10, has_size = flg & 0x08, content_checksum = flg & 0x04, has_dict = flg & 0x01;
  const int bsid = (bd >> 4) & 7;
  if (bsid < 4) return bad("bad block size id");
  const uint64_t block_max = 1ULL << (8 + 2 * bsid);  // 4 -> 64 KB ... 7 -> 4 MB
  uint64_t p = at + 6 + (has_size ? 8 : 0) + (has_dict ? 4 : 0) + 1;  // + header checksum byte
  if (p > end) return bad("truncated descriptor");
  const uint32_t first = (uint32_t)blocks->size();
  uint64_t produced = 0;
  for (;;) {
    if (p + 4 > end) return bad("truncated block header");
    const uint32_t h = rd32(p);
    p += 4;
    if (h == 0) break;  // end mark
    const uint32_t len = h & 0x7fffffffu;
    if (p + len > end) return bad("block beyond the buffer");
    if (produced >= want) return bad("more blocks than the uncompressed length allows");
    const uint64_t out = std::min<uint64_t>(block_max, want - produced);  // every block but the last is full
    if ((h >> 31) && len != out) return bad("stored block of unexpected size");
    blocks->push_back(Lz4Block{p, dst + produced, len, (uint32_t)out, (h >> 31) ? 1u : 0u,
                               independent ? 0u : (uint32_t)std::min<uint64_t>(produced, 65535)});
    produced += out;
    p += len + (block_checksum ? 4 : 0);
  }
  if (content_checksum) p += 4;
  if (p > end) return bad("truncated end mark");
  if (produced != want) return bad("blocks do not add up to the uncompressed length");
  const uint32_t n = (uint32_t)blocks->size() - first;
  if (independent)
    for (uint32_t i = 0; i < n; i++) chains->push_back(Lz4Chain{first + i, 1});
  else if (n)
    chains->push_back(Lz4Chain{first, n});
  return BGPU_OK;
}

// Decodes one LZ4 block with `lanes` cooperating threads (lane = this thread's index; 1 lane on the host). Every lane
// parses the sequence headers redundantly (uniform control flow, broadcast loads); the literal and match copies are
// strided over the lanes. Returns the number of bytes produced, or ~0u on malformed input. `sync` orders the lanes'
// writes before later reads of the output (matches read bytes other lanes wrote).
template <typename Sync>
__host__ __device__ inline uint32_t lz4_block_decode(const uint8_t* __restrict__ src, uint32_t src_len, uint8_t* dst, uint32_t dst_cap,
                                                     uint32_t window, uint32_t lane, uint32_t lanes, Sync sync) {
  uint32_t ip = 0, op = 0;
  while (ip < src_len) {
    const uint32_t token = src[ip++];
    uint32_t lit = token >> 4;
    if (lit == 15) {
      uint32_t b;
      do {
        if (ip >= src_len) return ~0u;
        b = src[ip++];
        lit += b;
      } while (b == 255);
    }
    if (lit > src_len - ip || lit > dst_cap - op) return ~0u;
    for (uint32_t i = lane; i < lit; i += lanes) dst[op + i] = src[ip + i];
    ip += lit;
    op += lit;
    if (ip >= src_len) break;  // the last sequence has literals only
    if (ip + 2 > src_len) return ~0u;
    const uint32_t offset = (uint32_t)src[ip] | ((uint32_t)src[ip + 1] << 8);
    ip += 2;
    uint32_t ml = token & 15;
    if (ml == 15) {
      uint32_t b;
      do {
        if (ip >= src_len) return ~0u;
        b = src[ip++];
        ml += b;
      } while (b == 255);
    }
    ml += 4;
    if (offset == 0 || offset > op + window || ml > dst_cap - op) return ~0u;
    sync();  // the literals just written (and everything before) may be match source
    const uint8_t* m = dst + op - offset;  // may reach before dst (linked blocks): that output is already complete
    if (offset >= ml) {
      for (uint32_t i = lane; i < ml; i += lanes) dst[op + i] = m[i];
    } else {  // overlapping match = the last `offset` bytes repeated
      for (uint32_t i = lane; i < ml; i += lanes) dst[op + i] = m[i % offset];
    }
    op += ml;
    sync();
  }
  return op;
}

constexpr int kLz4Threads = 128;  // 4 chains per CTA
__global__ void __launch_bounds__(kLz4Threads) lz4_decode_kernel(const Lz4Block* __restrict__ blocks, const Lz4Chain* __restrict__ chains,
                                                                 uint32_t n_chains, const uint8_t* __restrict__ body, uint8_t* arena,
                                                                 unsigned int* __restrict__ error) {
  const uint32_t chain = blockIdx.x * (kLz4Threads / 32) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (chain >= n_chains) return;
  const Lz4Chain c = chains[chain];
  for (uint32_t b = c.first; b < c.first + c.count; b++) {
    const Lz4Block k = blocks[b];
    uint32_t got;
    if (k.kind == 1) {
      for (uint32_t i = lane; i < k.src_len; i += 32) arena[k.dst + i] = body[k.src + i];
      got = k.src_len;
    } else {
      got = lz4_block_decode(body + k.src, k.src_len, arena + k.dst, k.dst_len, k.window, lane, 32u, [] { __syncwarp(); });
    }
    if (got != k.dst_len) {
      if (lane == 0) atomicExch(error, b + 1);
      return;
    }
    __syncwarp();
  }
}

// Block-independent frames (what lz4_flex - arrow-ipc's codec in the reference build - writes: 64 KB blocks that never
// reach into each other): one warp per block decodes INTO SHARED MEMORY (matches then read shared memory instead of
// output that was written to global memory a few cycles earlier and must come back from L2), and the finished block
// goes out with coalesced 16-byte stores. 64 KB of shared memory per CTA = 3 blocks in flight per SM.
constexpr uint32_t kLz4SmemBlock = 64 * 1024;
__global__ void __launch_bounds__(32) lz4_decode_smem_kernel(const Lz4Block* __restrict__ blocks, const uint32_t* __restrict__ which,
                                                             const uint8_t* __restrict__ body, uint8_t* __restrict__ arena,
                                                             unsigned int* __restrict__ error) {
  extern __shared__ __align__(16) uint8_t win[];
  const uint32_t b = which[blockIdx.x], lane = threadIdx.x;
  const Lz4Block k = blocks[b];
  const uint32_t got = lz4_block_decode(body + k.src, k.src_len, win, k.dst_len, 0u, lane, 32u, [] { __syncwarp(); });
  if (got != k.dst_len) {
    if (lane == 0) atomicExch(error, b + 1);
    return;
  }
  __syncwarp();
  uint8_t* out = arena + k.dst;  // 16-byte aligned: buffers start 256-byte aligned, blocks are 64 KB apart
  const uint32_t vecs = k.dst_len >> 4;
  for (uint32_t i = lane; i < vecs; i += 32) ((uint4*)out)[i] = ((const uint4*)win)[i];
  for (uint32_t i = (vecs << 4) + lane; i < k.dst_len; i += 32) out[i] = win[i];
}

inline uint64_t align_up(uint64_t x, uint64_t a) { return (x + a - 1) & ~(a - 1); }

}  // namespace
}  // namespace bgpu

using namespace bgpu;

struct bgpu_ipc_reader {
  bgpu_ctx* ctx = nullptr;
  const uint8_t* bytes = nullptr;
  uint64_t len = 0, pos = 0;
  bool finished = false;
  bgpu_batch* schema = nullptr;  // zero-row batch carrying the column descriptors
  uint64_t batches = 0, rows = 0;
  std::map<int64_t, std::shared_ptr<Column>> dicts;  // the current dictionary of every id (a later batch of an id replaces it)
};

// Reads the framing of the next message. *meta = flatbuffer bytes, *meta_len = 0 at end of stream.
static int next_message(bgpu_ipc_reader* r, const uint8_t** meta, uint32_t* meta_len) {
  *meta_len = 0;
  if (r->pos + 4 > r->len) return BGPU_OK;  // EOF without the end-of-stream marker is valid (async_reader/mod.rs:180-186)
  uint32_t v;
  memcpy(&v, r->bytes + r->pos, 4);
  r->pos += 4;
  if (v == 0xffffffffu) {
    if (r->pos + 4 > r->len) return r->ctx->fail(BGPU_EINVAL, "IPC reader: truncated message length");
    memcpy(&v, r->bytes + r->pos, 4);
    r->pos += 4;
  }
  if (v == 0) return BGPU_OK;
  if ((int32_t)v < 0 || r->pos + v > r->len) return r->ctx->fail(BGPU_EINVAL, "IPC reader: message metadata beyond the stream");
  *meta = r->bytes + r->pos;
  *meta_len = v;
  r->pos += v;
  return BGPU_OK;
}

static int reader_open_impl(bgpu_ctx* ctx, const uint8_t* bytes, uint64_t len, bgpu_ipc_reader** out) {
  if (!ctx || !out || (!bytes && len)) return BGPU_EINVAL;
  *out = nullptr;
  bgpu_ipc_reader tmp;
  tmp.ctx = ctx;
  tmp.bytes = bytes;
  tmp.len = len;
  const uint8_t* meta = nullptr;
  uint32_t meta_len = 0;
  BGPU_TRY(next_message(&tmp, &meta, &meta_len));
  if (!meta_len) return ctx->fail(BGPU_EINVAL, "IPC reader: empty stream (no schema message)");
  Fb fb{meta, meta_len};
  const Table msg{&fb, (size_t)fb.rd<uint32_t>(0)};
  if (!fb.ok || msg.scalar<uint8_t>(1, 0) != H_Schema) return ctx->fail(BGPU_EINVAL, "IPC reader: Unable to read IPC message as schema");
  const Table schema = msg.table(2);
  if (schema.scalar<int16_t>(0, 0) != 0) return ctx->fail(BGPU_EUNSUPPORTED, "IPC reader: big-endian stream");
  uint32_t n_fields = 0;
  const size_t fields = schema.vec(1, &n_fields);
  std::vector<Column> cols(n_fields);
  for (uint32_t i = 0; i < n_fields; i++) {
    const size_t at = fields + 4 * (size_t)i;
    const Table f{&fb, at + fb.rd<uint32_t>(at)};
    BGPU_TRY(field_to_column(ctx, f, &cols[i]));
  }
  if (!fb.ok) return ctx->fail(BGPU_EINVAL, "IPC reader: malformed schema message");
  const uint64_t schema_body = (uint64_t)msg.scalar<int64_t>(3, 0);  // a schema message has no body; be exact anyway
  if (schema_body > tmp.len - tmp.pos) return ctx->fail(BGPU_EINVAL, "IPC reader: message body beyond the stream");
  tmp.pos += schema_body;
  bgpu_ipc_reader* r = new bgpu_ipc_reader(tmp);
  ctx_retain(ctx);
  r->schema = new bgpu_batch(ctx);
  r->schema->cols = std::move(cols);
  *out = r;
  return BGPU_OK;
}

extern "C" int bgpu_ipc_reader_schema(bgpu_ipc_reader* r, struct ArrowSchema* out) {
  if (!r || !out) return BGPU_EINVAL;
  export_batch_schema(r->schema, out);
  return BGPU_OK;
}

extern "C" void bgpu_ipc_reader_close(bgpu_ipc_reader* r) {
  if (!r) return;
  delete r->schema;
  ctx_release(r->ctx);
  delete r;
}

// One buffer of a RecordBatch message as it will sit in the arena.
struct PlacedBuffer {
  uint64_t arena_off = 0, bytes = 0;
};

// offsets[0] >= 0, offsets non-decreasing, offsets[rows] <= bytes of the data buffer; *bad != 0 otherwise
__global__ void __launch_bounds__(256) check_offsets_kernel(const int32_t* __restrict__ offs, int64_t rows, int64_t data_bytes,
                                                            unsigned int* __restrict__ bad) {
  bool wrong = false;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i <= rows; i += (int64_t)gridDim.x * blockDim.x) {
    const int32_t o = offs[i];
    if (o < 0 || (int64_t)o > data_bytes) wrong = true;
    if (i < rows && offs[i + 1] < o) wrong = true;
  }
  if (wrong) atomicOr(bad, 1u);
}

static int decode_record_batch(bgpu_ipc_reader* r, Fb& fb, const Table rb, const uint8_t* body, uint64_t body_len,
                               const std::vector<Column>& proto, bgpu_batch** out);

// every valid dictionary index lies in [0, dictionary length); *bad != 0 otherwise
__global__ void __launch_bounds__(256) check_indices_kernel(const int32_t* __restrict__ idx, const uint8_t* __restrict__ valid,
                                                            int64_t rows, int64_t dict_len, unsigned int* __restrict__ bad) {
  bool wrong = false;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < rows; i += (int64_t)gridDim.x * blockDim.x) {
    if (valid && !((valid[i >> 3] >> (i & 7)) & 1)) continue;
    const int32_t k = idx[i];
    if (k < 0 || (int64_t)k >= dict_len) wrong = true;
  }
  if (wrong) atomicOr(bad, 1u);
}

static int reader_next_impl(bgpu_ipc_reader* r, bgpu_batch** out) {
  if (!r || !out) return BGPU_EINVAL;
  *out = nullptr;
  bgpu_ctx* ctx = r->ctx;
  if (!ctx->has_device) return ctx->fail(BGPU_ENODEVICE, "context has no device (libbgpu has no CPU fallback)");
  for (;;) {
  if (r->finished) return BGPU_OK;
  const uint8_t* meta = nullptr;
  uint32_t meta_len = 0;
  BGPU_TRY(next_message(r, &meta, &meta_len));
  if (!meta_len) {
    r->finished = true;
    return BGPU_OK;
  }
  Fb fb{meta, meta_len};
  const Table msg{&fb, (size_t)fb.rd<uint32_t>(0)};
  const uint8_t header = msg.scalar<uint8_t>(1, 0);
  const uint64_t body_len = (uint64_t)msg.scalar<int64_t>(3, 0);
  if (!fb.ok || body_len > r->len - r->pos) return ctx->fail(BGPU_EINVAL, "IPC reader: message body beyond the stream");
  const uint8_t* body = r->bytes + r->pos;
  r->pos += body_len;
  if (header == H_Schema) return ctx->fail(BGPU_EINVAL, "IPC reader: Not expecting a schema when messages are read");
  if (header == H_NONE) {
    r->finished = true;
    return BGPU_OK;
  }
  if (header == H_DictionaryBatch) {
    // DictionaryBatch{id, data: RecordBatch of ONE column (the values), isDelta}: becomes the dictionary of every
    // record batch that follows, until a batch of the same id replaces it (async_reader/mod.rs:168-258)
    const Table db = msg.table(2);
    const int64_t id = db.scalar<int64_t>(0, 0);
    if (db.scalar<uint8_t>(2, 0)) return ctx->fail(BGPU_EUNSUPPORTED, "IPC reader: delta dictionary batch");
    const Column* field = nullptr;
    for (const Column& c : r->schema->cols)
      if (c.kind == K_DICT32 && c.dict_id == id) field = &c;
    if (!fb.ok || !field) return ctx->fail(BGPU_EINVAL, "IPC reader: dictionary batch for an id the schema does not have");
    std::vector<Column> one{*field->dict};
    bgpu_batch* values = nullptr;
    BGPU_TRY(decode_record_batch(r, fb, db.table(1), body, body_len, one, &values));
    auto col = std::make_shared<Column>(values->cols[0]);
    col->length = values->n_rows;
    delete values;
    r->dicts[id] = col;
    continue;
  }
  if (header != H_RecordBatch) return ctx->fail(BGPU_EINVAL, "IPC reader: unexpected message header type " + std::to_string((int)header));
  bgpu_batch* b = nullptr;
  BGPU_TRY(decode_record_batch(r, fb, msg.table(2), body, body_len, r->schema->cols, &b));
  for (Column& c : b->cols) {
    if (c.kind != K_DICT32) continue;
    auto it = r->dicts.find(c.dict_id);
    if (it == r->dicts.end()) {
      delete b;
      return ctx->fail(BGPU_EINVAL, "IPC reader: record batch before the dictionary of column " + c.name);
    }
    c.dict = it->second;
  }
  {  // indices come from another process: checked before any kernel looks a value up with them
    DevAllocPtr d_bad;
    for (const Column& c : b->cols) {
      if (c.kind != K_DICT32 || b->n_rows == 0) continue;
      if (!d_bad) {
        const int rc = dev_alloc(ctx, 4, &d_bad);
        if (rc != BGPU_OK) { delete b; return rc; }
        cudaMemsetAsync(d_bad->p, 0, 4, ctx->stream);
      }
      check_indices_kernel<<<(int)std::min<int64_t>(ceil_div64(b->n_rows, 256), (int64_t)ctx->sm_count * 8), 256, 0, ctx->stream>>>(
          (const int32_t*)c.values, c.validity, b->n_rows, c.dict->length, (unsigned int*)d_bad->p);
      ctx->launches++;
    }
    if (d_bad) {
      unsigned int bad = 0;
      cudaError_t e = cudaMemcpyAsync(&bad, d_bad->p, 4, cudaMemcpyDeviceToHost, ctx->stream);
      if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
      if (e != cudaSuccess || bad) {
        delete b;
        if (e != cudaSuccess) return ctx->fail(BGPU_ECUDA, std::string("IPC reader: ") + cudaGetErrorString(e));
        return ctx->fail(BGPU_EINVAL, "IPC reader: dictionary index beyond its dictionary");
      }
    }
  }
  r->batches++;
  r->rows += (uint64_t)b->n_rows;
  *out = b;
  return BGPU_OK;
  }
}

// One RecordBatch table + its body -> a device batch with the columns of `proto` (the stream's fields, or the one value
// column of a dictionary batch). Dictionary-encoded fields are read as their Int32 indices.
static int decode_record_batch(bgpu_ipc_reader* r, Fb& fb, const Table rb, const uint8_t* body, uint64_t body_len,
                               const std::vector<Column>& proto, bgpu_batch** out) {
  bgpu_ctx* ctx = r->ctx;
  const int64_t rows = rb.scalar<int64_t>(0, 0);
  uint32_t n_nodes = 0, n_buffers = 0;
  const size_t nodes = rb.vec(1, &n_nodes), buffers = rb.vec(2, &n_buffers);
  const Table comp = rb.table(3);
  const bool compressed = (bool)comp;
  if (compressed && (comp.scalar<int8_t>(0, 0) != 0 || comp.scalar<int8_t>(1, 0) != 0))
    return ctx->fail(BGPU_EUNSUPPORTED, "IPC reader: only LZ4_FRAME buffer compression is supported");
  size_t want_buffers = 0;
  for (const Column& c : proto) want_buffers += (c.kind == K_UTF8 || c.kind == K_BINARY) ? 3 : 2;
  if (!fb.ok || rows < 0 || n_nodes != proto.size() || n_buffers != want_buffers)
    return ctx->fail(BGPU_EINVAL, "IPC reader: record batch does not match the schema");
  if (rows > 0x7fffffffLL) return ctx->fail(BGPU_EUNSUPPORTED, "IPC reader: more than 2^31 rows in one batch");

  // 1. place every buffer in the arena (256-byte aligned); plan the LZ4 work for compressed ones
  std::vector<PlacedBuffer> placed(n_buffers);
  std::vector<Lz4Block> blocks;
  std::vector<Lz4Chain> chains;
  struct Copy { uint64_t src, dst, bytes; };
  std::vector<Copy> copies;  // uncompressed message: host body -> arena
  uint64_t arena_bytes = 0;
  for (uint32_t i = 0; i < n_buffers; i++) {
    const uint64_t off = (uint64_t)fb.rd<int64_t>(buffers + 16 * (size_t)i), len = (uint64_t)fb.rd<int64_t>(buffers + 16 * (size_t)i + 8);
    if (off + len > body_len || off + len < off) return ctx->fail(BGPU_EINVAL, "IPC reader: buffer beyond the message body");
    PlacedBuffer& pb = placed[i];
    pb.arena_off = arena_bytes;
    if (!compressed) {
      pb.bytes = len;
      if (len) copies.push_back(Copy{off, pb.arena_off, len});
    } else if (len) {
      if (len < 8) return ctx->fail(BGPU_EINVAL, "IPC reader: compressed buffer without its length prefix");
      int64_t raw;
      memcpy(&raw, body + off, 8);
      if (raw > (int64_t)1 << 40) return ctx->fail(BGPU_EINVAL, "IPC reader: implausible uncompressed length");
      if (raw == -1) {  // stored as is
        pb.bytes = len - 8;
        for (uint64_t done = 0; done < pb.bytes; done += 1u << 20) {  // 1 MB steps so that large buffers spread over warps
          const uint32_t step = (uint32_t)std::min<uint64_t>(1u << 20, pb.bytes - done);
          blocks.push_back(Lz4Block{off + 8 + done, pb.arena_off + done, step, step, 1, 0});
          chains.push_back(Lz4Chain{(uint32_t)blocks.size() - 1, 1});
        }
      } else {
        if (raw < 0) return ctx->fail(BGPU_EINVAL, "IPC reader: negative uncompressed length");
        pb.bytes = (uint64_t)raw;
        if (raw) BGPU_TRY(lz4_frame_plan(ctx, body, off + 8, off + len, pb.arena_off, pb.bytes, &blocks, &chains));
      }
    }
    arena_bytes = align_up(arena_bytes + pb.bytes, 256);
  }

  // 2. move / decode (accounted as h2d_time: the copy of the file's bytes dominates; the LZ4 decode rides behind it)
  BGPU_CUDA(ctx, cudaSetDevice(ctx->device));
  GpuTimeScope timed(ctx, T_H2D);
  DevAllocPtr arena;
  BGPU_TRY(dev_alloc(ctx, arena_bytes + 16, &arena));  // + 16: 8-byte-window loads at the end of utf8 data stay in bounds
  uint8_t* base = (uint8_t*)arena->p;
  if (!compressed) {
    for (const Copy& c : copies) BGPU_CUDA(ctx, cudaMemcpyAsync(base + c.dst, body + c.src, c.bytes, cudaMemcpyHostToDevice, ctx->stream));
  } else if (!chains.empty()) {
    DevAllocPtr d_body, d_blocks, d_chains, d_err;
    BGPU_TRY(dev_alloc(ctx, body_len, &d_body));
    BGPU_TRY(dev_alloc(ctx, blocks.size() * sizeof(Lz4Block), &d_blocks));
    BGPU_TRY(dev_alloc(ctx, chains.size() * sizeof(Lz4Chain), &d_chains));
    BGPU_TRY(dev_alloc(ctx, 4, &d_err));
    BGPU_CUDA(ctx, cudaMemcpyAsync(d_body->p, body, body_len, cudaMemcpyHostToDevice, ctx->stream));
    BGPU_CUDA(ctx, cudaMemcpyAsync(d_blocks->p, blocks.data(), blocks.size() * sizeof(Lz4Block), cudaMemcpyHostToDevice, ctx->stream));
    BGPU_CUDA(ctx, cudaMemsetAsync(d_err->p, 0, 4, ctx->stream));
    // single-block chains of at most 64 KB (block-independent frames) decode in shared memory; the rest - linked frames,
    // stored buffers, larger blocks - take the general kernel
    std::vector<uint32_t> small;
    std::vector<Lz4Chain> general;
    for (const Lz4Chain& c : chains) {
      const Lz4Block& k = blocks[c.first];
      if (c.count == 1 && k.kind == 0 && k.dst_len <= kLz4SmemBlock && (k.dst & 15) == 0) small.push_back(c.first);
      else general.push_back(c);
    }
    if (!general.empty()) {
      BGPU_CUDA(ctx, cudaMemcpyAsync(d_chains->p, general.data(), general.size() * sizeof(Lz4Chain), cudaMemcpyHostToDevice, ctx->stream));
      const uint32_t n_chains = (uint32_t)general.size();
      lz4_decode_kernel<<<(n_chains + kLz4Threads / 32 - 1) / (kLz4Threads / 32), kLz4Threads, 0, ctx->stream>>>(
          (const Lz4Block*)d_blocks->p, (const Lz4Chain*)d_chains->p, n_chains, (const uint8_t*)d_body->p, base, (unsigned int*)d_err->p);
      ctx->launches++;
    }
    DevAllocPtr d_small;
    if (!small.empty()) {
      BGPU_TRY(dev_alloc(ctx, small.size() * 4, &d_small));
      BGPU_CUDA(ctx, cudaMemcpyAsync(d_small->p, small.data(), small.size() * 4, cudaMemcpyHostToDevice, ctx->stream));
      BGPU_CUDA(ctx, raise_max_dynamic_smem((const void*)lz4_decode_smem_kernel, ctx->device, kLz4SmemBlock));
      lz4_decode_smem_kernel<<<(unsigned)small.size(), 32, kLz4SmemBlock, ctx->stream>>>(
          (const Lz4Block*)d_blocks->p, (const uint32_t*)d_small->p, (const uint8_t*)d_body->p, base, (unsigned int*)d_err->p);
      ctx->launches++;
    }
    BGPU_CUDA(ctx, cudaGetLastError());
    unsigned int err = 0;
    BGPU_CUDA(ctx, cudaMemcpyAsync(&err, d_err->p, 4, cudaMemcpyDeviceToHost, ctx->stream));
    BGPU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (err) return ctx->fail(BGPU_EINVAL, "IPC reader: malformed LZ4 block " + std::to_string(err - 1));
  }
  // the host bytes are the caller's and only borrowed for the call: the copies must have left them
  if (!compressed && !copies.empty()) BGPU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));

  // 3. columns over the arena
  bgpu_batch* b = new bgpu_batch(ctx);
  b->n_rows = rows;
  b->cols = proto;
  uint32_t bi = 0;
  for (size_t i = 0; i < proto.size(); i++) {
    Column& c = b->cols[i];
    const int64_t length = fb.rd<int64_t>(nodes + 16 * i), nulls = fb.rd<int64_t>(nodes + 16 * i + 8);
    bool ok = length == rows && nulls >= 0 && nulls <= rows;
    c.length = rows;
    c.null_count = nulls;
    c.owned = {arena};
    const PlacedBuffer& v = placed[bi++];
    c.validity = (nulls > 0 && v.bytes) ? base + v.arena_off : nullptr;
    if (nulls > 0) ok = ok && v.bytes >= (uint64_t)(rows + 7) / 8;
    if (c.kind == K_UTF8 || c.kind == K_BINARY) {
      const PlacedBuffer& o = placed[bi++];
      const PlacedBuffer& d = placed[bi++];
      c.offsets = (int32_t*)(base + o.arena_off);
      c.values = base + d.arena_off;
      c.data_bytes = (int64_t)d.bytes;
      ok = ok && (rows == 0 || o.bytes >= (uint64_t)(rows + 1) * 4);
    } else {
      const PlacedBuffer& d = placed[bi++];
      c.values = base + d.arena_off;
      ok = ok && d.bytes >= (c.kind == K_BOOL ? (uint64_t)(rows + 7) / 8 : (uint64_t)rows * (uint64_t)c.width);
    }
    if (!ok) {
      delete b;
      return ctx->fail(BGPU_EINVAL, "IPC reader: buffers of column " + proto[i].name + " are shorter than its length");
    }
  }
  // The CONTENTS of utf8 / binary offset buffers come from another process: every kernel downstream indexes the data
  // buffer with them, so they are checked here as arrow-rs's reader does (first offset >= 0, non-decreasing, last offset
  // within the data buffer) before the batch is handed out.
  {
    DevAllocPtr d_bad;
    for (const Column& c : b->cols) {
      if ((c.kind != K_UTF8 && c.kind != K_BINARY) || rows == 0) continue;
      if (!d_bad) {
        int rc = dev_alloc(ctx, 4, &d_bad);
        if (rc != BGPU_OK) { delete b; return rc; }
        cudaMemsetAsync(d_bad->p, 0, 4, ctx->stream);
      }
      check_offsets_kernel<<<(int)std::min<int64_t>(ceil_div64(rows + 1, 256), (int64_t)ctx->sm_count * 8), 256, 0, ctx->stream>>>(
          c.offsets, rows, c.data_bytes, (unsigned int*)d_bad->p);
      ctx->launches++;
    }
    if (d_bad) {
      unsigned int bad = 0;
      cudaError_t e = cudaMemcpyAsync(&bad, d_bad->p, 4, cudaMemcpyDeviceToHost, ctx->stream);
      if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
      if (e != cudaSuccess || bad) {
        delete b;
        if (e != cudaSuccess) return ctx->fail(BGPU_ECUDA, std::string("IPC reader: ") + cudaGetErrorString(e));
        return ctx->fail(BGPU_EINVAL, "IPC reader: offsets of a utf8/binary column are negative, decreasing or beyond its data buffer");
      }
    }
  }
  *out = b;
  return BGPU_OK;
}

// Test hook: the LZ4 frame walker + block decoder run on the host with one lane. NOT a product path (nothing in the
// library calls it); it lets the CPU test-suite pin the decoder logic against frames written by Arrow's own codec.
static int lz4_frame_decode_host_impl(bgpu_ctx* ctx, const uint8_t* frame, uint64_t len, uint8_t* out, uint64_t out_len) {
  if (!ctx || !frame || (!out && out_len)) return BGPU_EINVAL;
  std::vector<Lz4Block> blocks;
  std::vector<Lz4Chain> chains;
  BGPU_TRY(lz4_frame_plan(ctx, frame, 0, len, 0, out_len, &blocks, &chains));
  for (const Lz4Chain& c : chains)
    for (uint32_t b = c.first; b < c.first + c.count; b++) {
      const Lz4Block& k = blocks[b];
      uint32_t got;
      if (k.kind == 1) {
        memcpy(out + k.dst, frame + k.src, k.src_len);
        got = k.src_len;
      } else {
        got = lz4_block_decode(frame + k.src, k.src_len, out + k.dst, k.dst_len, k.window, 0u, 1u, [] {});
      }
      if (got != k.dst_len) return ctx->fail(BGPU_EINVAL, "LZ4 block " + std::to_string(b) + " is malformed");
    }
  return BGPU_OK;
}

// Nothing is thrown across the C ABI: a malformed stream that drives an allocation out of memory is an error code.
extern "C" int bgpu_ipc_reader_open(bgpu_ctx* ctx, const uint8_t* bytes, uint64_t len, bgpu_ipc_reader** out) {
  try {
    return reader_open_impl(ctx, bytes, len, out);
  } catch (const std::exception& e) {
    if (out) *out = nullptr;
    return ctx ? ctx->fail(BGPU_EINVAL, std::string("IPC reader: ") + e.what()) : BGPU_EINVAL;
  }
}

extern "C" int bgpu_ipc_reader_next(bgpu_ipc_reader* r, bgpu_batch** out) {
  try {
    return reader_next_impl(r, out);
  } catch (const std::exception& e) {
    if (out) *out = nullptr;
    return r && r->ctx ? r->ctx->fail(BGPU_EINVAL, std::string("IPC reader: ") + e.what()) : BGPU_EINVAL;
  }
}

extern "C" int bgpu_debug_lz4_frame_decode_host(bgpu_ctx* ctx, const uint8_t* frame, uint64_t len, uint8_t* out, uint64_t out_len) {
  try {
    return lz4_frame_decode_host_impl(ctx, frame, len, out, out_len);
  } catch (const std::exception& e) {
    return ctx ? ctx->fail(BGPU_EINVAL, std::string("LZ4 frame: ") + e.what()) : BGPU_EINVAL;
  }
}
