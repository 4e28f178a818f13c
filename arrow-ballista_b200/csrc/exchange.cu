// K6: intra-node exchange of hash-partitioned batches — one NCCL all-to-all(v) over NVLink 5 / NVSwitch.
//
// Replaces, inside one 8-GPU node, the reference's shuffle file write (shuffle_writer.rs:351-385), the Flight do_get
// that serves the file (flight_service.rs:124-166) and ShuffleReaderExec's fan-in (shuffle_reader.rs:539-624):
// output partition p of every rank's BatchPartitioner result goes to rank p % world. The received
// batch holds this rank's partitions from all ranks, grouped by (partition ascending, source rank ascending).
//
// Protocol (all on the ctx stream): 1. ncclAllGather of a small per-rank header (row offsets per partition, byte
// offsets per partition of every utf8 column, validity flags); 2. one grouped ncclSend/ncclRecv per (buffer, peer,
// partition) straight out of / into the Arrow buffers — no staging copies; validity bitmaps travel as one byte per
// row (bit ranges are not byte aligned) and utf8 offsets as lengths, re-packed / re-scanned on arrival.
// libnccl.so.2 is dlopen'ed (the copy torch already loaded is reused), so libbgpu has no link-time NCCL dependency.
#include <dlfcn.h>
#include <string.h>

#include <algorithm>
#include <mutex>

#include <cub/device/device_scan.cuh>

#include "internal.h"
#include "nccl_dl.h"

namespace {

constexpr int kThreads = 256;

__global__ void __launch_bounds__(kThreads) bits_to_bytes_kernel(const uint8_t* __restrict__ bits, int64_t n, uint8_t* __restrict__ out) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    out[i] = bits ? ((bits[i >> 3] >> (i & 7)) & 1) : 1;
}
__global__ void __launch_bounds__(kThreads) bytes_to_bits_kernel(const uint8_t* __restrict__ in, int64_t n, uint32_t* __restrict__ bits) {
  const int64_t n_pad = (n + 31) & ~31LL;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_pad; i += (int64_t)gridDim.x * blockDim.x) {
    const uint32_t b = __ballot_sync(0xffffffffu, i < n && in[i] != 0);
    if ((threadIdx.x & 31) == 0) bits[i >> 5] = b;
  }
}
__global__ void __launch_bounds__(kThreads) offsets_to_lengths_kernel(const int32_t* __restrict__ offs, int64_t n, int32_t* __restrict__ len) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    len[i] = offs[i + 1] - offs[i];
}
// rows of partition boundaries -> (absolute) byte offsets into the data buffer of a utf8 column
__global__ void partition_byte_offsets_kernel(const int32_t* __restrict__ offs, const uint64_t* __restrict__ row_off, uint32_t n,
                                              uint64_t* __restrict__ out) {
  for (uint32_t p = threadIdx.x; p <= n; p += blockDim.x) out[p] = (uint64_t)offs[row_off[p]];
}

int grid_for(bgpu_ctx* ctx, int64_t n) {
  return (int)std::max<int64_t>(1, std::min<int64_t>(ceil_div64(n, kThreads), (int64_t)ctx->sm_count * 8));
}

}  // namespace

extern "C" int bgpu_nccl_unique_id(bgpu_ctx* ctx, void* out_id128) {
  if (!ctx || !out_id128) return BGPU_EINVAL;
  const Nccl& n = nccl();
  if (!n.ok) return ctx->fail(BGPU_ENCCL, "NCCL unavailable: " + n.why);
  NcclUniqueId id;
  BGPU_NCCL(ctx, n.GetUniqueId(&id));
  memcpy(out_id128, &id, sizeof id);
  return BGPU_OK;
}

extern "C" int bgpu_nccl_comm_create(bgpu_ctx* ctx, const void* id128, int32_t rank, int32_t world, void** out_comm) try {
  if (!ctx || !id128 || !out_comm) return BGPU_EINVAL;
  *out_comm = nullptr;
  if (!ctx->has_device) return ctx->fail(BGPU_ENODEVICE, "context has no device");
  const Nccl& n = nccl();
  if (!n.ok) return ctx->fail(BGPU_ENCCL, "NCCL unavailable: " + n.why);
  BGPU_CUDA(ctx, cudaSetDevice(ctx->device));
  NcclUniqueId id;
  memcpy(&id, id128, sizeof id);
  BGPU_NCCL(ctx, n.CommInitRank(out_comm, world, id, rank));
  return BGPU_OK;
} catch (const std::exception& e) {
  return bgpu::abi_guard(ctx, e);  // nothing is thrown across the C ABI
}

extern "C" void bgpu_nccl_comm_destroy(void* comm) {
  if (comm && nccl().ok) nccl().CommDestroy(comm);
}

// Host-side plan of the exchange, shared by bgpu_all_to_all and the CPU tests (no device needed).
// all_offsets: [world][n_partitions + 1] row (or byte) offsets of every rank's partitioned buffer.
// For this rank: send_* describe what goes to each (peer, partition); recv_off[(p_local * world + s)] is where the rows
// of owned partition p from source s start in the received buffer; returns the received total.
extern "C" uint64_t bgpu_exchange_plan(const uint64_t* all_offsets, int32_t rank, int32_t world, uint32_t n_partitions,
                                       uint64_t* recv_off /* [owned * world + 1] */, uint32_t* n_owned) {
  uint64_t total = 0;
  uint32_t k = 0;
  for (uint32_t p = (uint32_t)rank; p < n_partitions; p += (uint32_t)world)
    for (int32_t s = 0; s < world; s++) {
      const uint64_t* o = all_offsets + (size_t)s * (n_partitions + 1);
      recv_off[k++] = total;
      total += o[p + 1] - o[p];
    }
  recv_off[k] = total;
  if (n_owned) *n_owned = k / (uint32_t)world;
  return total;
}

extern "C" int bgpu_all_to_all(bgpu_ctx* ctx, void* comm, int32_t rank, int32_t world, const bgpu_batch* in,
                               const uint64_t* offsets, uint32_t n_partitions, bgpu_batch** out) try {
  if (!ctx || !out) return BGPU_EINVAL;
  *out = nullptr;
  if (!comm || !in || !offsets || world < 1 || rank < 0 || rank >= world || n_partitions == 0)
    return ctx->fail(BGPU_EINVAL, "bgpu_all_to_all: bad arguments");
  if (!ctx->has_device) return ctx->fail(BGPU_ENODEVICE, "context has no device");
  const Nccl& nc = nccl();
  if (!nc.ok) return ctx->fail(BGPU_ENCCL, "NCCL unavailable: " + nc.why);
  BGPU_CUDA(ctx, cudaSetDevice(ctx->device));
  GpuTimeScope timed(ctx, T_KERNEL);
  if (bgpu::batch_has_dictionary(in)) {
    // dictionary columns (schema-level: the same decision on every rank): the values travel, the received rows are
    // re-encoded against a dictionary of their own (dictionary.cu)
    std::unique_ptr<bgpu_batch> decoded;
    std::vector<size_t> dict_cols;
    BGPU_TRY(bgpu::dict_decode_batch(ctx, in, &decoded, &dict_cols));
    bgpu_batch* got = nullptr;
    BGPU_TRY(bgpu_all_to_all(ctx, comm, rank, world, decoded.get(), offsets, n_partitions, &got));
    std::unique_ptr<bgpu_batch> res(got);
    for (size_t i : dict_cols) BGPU_TRY(bgpu::dict_identity_encode(ctx, res.get(), i, in->cols[i], nullptr, 0));
    *out = res.release();
    return BGPU_OK;
  }
  cudaStream_t st = ctx->stream;
  const size_t n_cols = in->cols.size();
  const uint32_t P = n_partitions, P1 = n_partitions + 1;
  std::vector<int> var_cols;
  for (size_t c = 0; c < n_cols; c++) {
    const Column& col = in->cols[c];
    if (col.kind == K_DICT32)
      return ctx->fail(BGPU_EUNSUPPORTED, "exchange of dictionary-encoded column " + col.name + " is not supported yet");
    if (col.kind == K_BOOL) return ctx->fail(BGPU_EUNSUPPORTED, "exchange of Boolean column " + col.name + " is not supported yet");
    if (col.kind == K_UTF8 || col.kind == K_BINARY) var_cols.push_back((int)c);
  }
  // ---- 1. header all-gather: [P1 row offsets][n_cols validity flags][P1 byte offsets per utf8 column] -----------------
  const size_t hdr_words = P1 + n_cols + var_cols.size() * P1;
  DevAllocPtr d_hdr, d_all;
  BGPU_TRY(dev_alloc(ctx, hdr_words * 8, &d_hdr));
  BGPU_TRY(dev_alloc(ctx, hdr_words * 8 * (size_t)world, &d_all));
  std::vector<uint64_t> hdr(hdr_words, 0);
  memcpy(hdr.data(), offsets, P1 * 8);
  for (size_t c = 0; c < n_cols; c++) hdr[P1 + c] = in->cols[c].validity ? 1 : 0;
  BGPU_CUDA(ctx, cudaMemcpyAsync(d_hdr->p, hdr.data(), hdr_words * 8, cudaMemcpyHostToDevice, st));
  for (size_t v = 0; v < var_cols.size(); v++) {
    partition_byte_offsets_kernel<<<1, 256, 0, st>>>(in->cols[(size_t)var_cols[v]].offsets, (const uint64_t*)d_hdr->p, P,
                                                     (uint64_t*)d_hdr->p + P1 + n_cols + v * P1);
    ctx->launches++;
  }
  BGPU_NCCL(ctx, nc.AllGather(d_hdr->p, d_all->p, hdr_words * 8, kNcclUint8, comm, st));
  std::vector<uint64_t> all(hdr_words * (size_t)world);
  BGPU_CUDA(ctx, cudaMemcpyAsync(all.data(), d_all->p, all.size() * 8, cudaMemcpyDeviceToHost, st));
  BGPU_CUDA(ctx, cudaStreamSynchronize(st));
  auto row_off = [&](int s, uint32_t p) { return all[(size_t)s * hdr_words + p]; };
  auto byte_off = [&](int s, size_t v, uint32_t p) { return all[(size_t)s * hdr_words + P1 + n_cols + v * P1 + p]; };
  // ---- 2. layout of the received batch ------------------------------------------------------------------------------------
  std::vector<uint32_t> owned;
  for (uint32_t p = (uint32_t)rank; p < P; p += (uint32_t)world) owned.push_back(p);
  std::vector<uint64_t> recv_row(owned.size() * (size_t)world + 1, 0);
  {
    std::vector<uint64_t> rows((size_t)world * P1);
    for (int s = 0; s < world; s++)
      for (uint32_t p = 0; p <= P; p++) rows[(size_t)s * P1 + p] = row_off(s, p);
    bgpu_exchange_plan(rows.data(), rank, world, P, recv_row.data(), nullptr);
  }
  const int64_t n_out = (int64_t)recv_row.back();
  // Limits are checked for EVERY rank from the gathered header, so that all ranks take the same branch: a rank that
  // returned alone would leave its peers blocked in their receives.
  for (int r = 0; r < world; r++) {
    uint64_t rows_r = 0;
    std::vector<uint64_t> bytes_r(var_cols.size(), 0);
    for (uint32_t p = (uint32_t)r; p < P; p += (uint32_t)world)
      for (int s = 0; s < world; s++) {
        rows_r += row_off(s, p + 1) - row_off(s, p);
        for (size_t v = 0; v < var_cols.size(); v++) bytes_r[v] += byte_off(s, v, p + 1) - byte_off(s, v, p);
      }
    if (rows_r > 0x7fffffffULL)
      return ctx->fail(BGPU_EUNSUPPORTED, "received batch of rank " + std::to_string(r) + " too large");
    for (uint64_t b : bytes_r)
      if (b > 0x7fffffffULL)
        return ctx->fail(BGPU_EUNSUPPORTED, "received utf8 column of rank " + std::to_string(r) + " exceeds 2 GiB (i32 offsets)");
  }
  std::unique_ptr<bgpu_batch> ob(new bgpu_batch(ctx));
  ob->n_rows = n_out;
  struct Xfer { const char* src; char* dst; size_t elem; int var; };  // var >= 0: byte ranges of utf8 column var
  std::vector<Xfer> xfers;
  std::vector<DevAllocPtr> keep;
  struct Post { size_t col; DevAllocPtr vbytes, lens; };
  std::vector<Post> posts;
  // Everything below that can fail on ONE rank (allocations, launches) runs inside `prepare`; its outcome is exchanged
  // (one word per rank) before anybody posts a send or a receive, so that a failing rank takes its peers out with it.
  auto prepare = [&]() -> int {
  for (size_t c = 0; c < n_cols; c++) {
    const Column& ic = in->cols[c];
    Column oc;
    oc.name = ic.name; oc.format = ic.format; oc.flags = ic.flags; oc.kind = ic.kind; oc.width = ic.width;
    oc.length = n_out; oc.null_count = -1;
    Post post;
    post.col = c;
    bool any_valid = false;
    for (int s = 0; s < world; s++) any_valid |= all[(size_t)s * hdr_words + P1 + c] != 0;
    if (any_valid) {  // validity as one byte per row
      DevAllocPtr src, dst;
      BGPU_TRY(dev_alloc(ctx, (size_t)std::max<int64_t>(in->n_rows, 1), &src));
      BGPU_TRY(dev_alloc(ctx, (size_t)std::max<int64_t>(n_out, 1), &dst));
      if (in->n_rows) {
        bits_to_bytes_kernel<<<grid_for(ctx, in->n_rows), kThreads, 0, st>>>(ic.validity, in->n_rows, (uint8_t*)src->p);
        ctx->launches++;
      }
      xfers.push_back(Xfer{(const char*)src->p, (char*)dst->p, 1, -1});
      keep.push_back(src);
      post.vbytes = dst;
    } else {
      oc.null_count = 0;
    }
    if (ic.kind == K_FIXED) {
      DevAllocPtr vals;
      BGPU_TRY(dev_alloc(ctx, (size_t)std::max<int64_t>(n_out, 1) * (size_t)ic.width, &vals));
      xfers.push_back(Xfer{(const char*)ic.values, (char*)vals->p, (size_t)ic.width, -1});
      oc.values = vals->p;
      oc.owned.push_back(vals);
    } else {  // utf8 / binary: lengths + bytes
      size_t v = 0;
      while (var_cols[v] != (int)c) v++;
      DevAllocPtr lens_src, lens_dst, data, offs;
      BGPU_TRY(dev_alloc(ctx, (size_t)std::max<int64_t>(in->n_rows, 1) * 4, &lens_src));
      BGPU_TRY(dev_alloc(ctx, (size_t)std::max<int64_t>(n_out, 1) * 4, &lens_dst));
      if (in->n_rows) {
        offsets_to_lengths_kernel<<<grid_for(ctx, in->n_rows), kThreads, 0, st>>>(ic.offsets, in->n_rows, (int32_t*)lens_src->p);
        ctx->launches++;
      }
      uint64_t total_bytes = 0;
      for (uint32_t p : owned)
        for (int s = 0; s < world; s++) total_bytes += byte_off(s, v, p + 1) - byte_off(s, v, p);
      BGPU_TRY(dev_alloc(ctx, (size_t)total_bytes, &data));
      BGPU_TRY(dev_alloc(ctx, (size_t)(n_out + 1) * 4, &offs));
      xfers.push_back(Xfer{(const char*)lens_src->p, (char*)lens_dst->p, 4, -1});
      xfers.push_back(Xfer{(const char*)ic.values, (char*)data->p, 1, (int)v});
      keep.push_back(lens_src);
      post.lens = lens_dst;
      oc.values = data->p; oc.offsets = (int32_t*)offs->p; oc.data_bytes = (int64_t)total_bytes;
      oc.owned.push_back(data); oc.owned.push_back(offs);
    }
    posts.push_back(post);
    ob->cols.push_back(std::move(oc));
  }
  return BGPU_OK;
  };
  const int local_rc = prepare();
  {
    DevAllocPtr d_err;
    uint32_t* pin = (uint32_t*)ctx->pinned + 128;  // [0] mine, [1 .. world] everybody's
    pin[0] = (uint32_t)local_rc;
    // d_all (the gathered header, already read) is reused as the device side of this one-word-per-rank collective
    cudaError_t e = cudaMemcpyAsync(d_hdr->p, pin, 4, cudaMemcpyHostToDevice, st);
    if (e != cudaSuccess) return ctx->fail(BGPU_ECUDA, cudaGetErrorString(e));
    BGPU_NCCL(ctx, nc.AllGather(d_hdr->p, d_all->p, 4, kNcclUint8, comm, st));
    BGPU_CUDA(ctx, cudaMemcpyAsync(pin + 1, d_all->p, 4 * (size_t)world, cudaMemcpyDeviceToHost, st));
    BGPU_CUDA(ctx, cudaStreamSynchronize(st));
    if (local_rc != BGPU_OK) return local_rc;
    for (int r = 0; r < world; r++)
      if (pin[1 + r] != BGPU_OK)
        return ctx->fail(BGPU_ENCCL, "exchange abandoned: rank " + std::to_string(r) + " failed to prepare its buffers (code " +
                                         std::to_string(pin[1 + r]) + ")");
  }
  // ---- 3. the exchange: one grouped send/recv per (buffer, peer, partition) --------------------------------------------
  // (an error inside the group still closes it: an open group would poison every later NCCL call of this thread)
  int nccl_rc = 0;
  BGPU_NCCL(ctx, nc.GroupStart());
  for (const Xfer& x : xfers) {
    // sends: my partition p goes to rank p % world
    for (uint32_t p = 0; p < P; p++) {
      const int peer = (int)(p % (uint32_t)world);
      uint64_t lo, hi;
      if (x.var >= 0) {
        lo = byte_off(rank, (size_t)x.var, p); hi = byte_off(rank, (size_t)x.var, p + 1);
      } else {
        lo = row_off(rank, p) * x.elem; hi = row_off(rank, p + 1) * x.elem;
      }
      if (hi > lo && !nccl_rc) nccl_rc = nc.Send(x.src + lo, (size_t)(hi - lo), kNcclUint8, peer, comm, st);
    }
    // receives, in the received batch's order (partition, source)
    uint64_t at = 0;
    for (uint32_t p : owned)
      for (int s = 0; s < world; s++) {
        const uint64_t n = x.var >= 0 ? byte_off(s, (size_t)x.var, p + 1) - byte_off(s, (size_t)x.var, p)
                                      : (row_off(s, p + 1) - row_off(s, p)) * x.elem;
        if (n && !nccl_rc) nccl_rc = nc.Recv(x.dst + at, (size_t)n, kNcclUint8, s, comm, st);
        at += n;
      }
  }
  {
    const int end_rc = nc.GroupEnd();
    if (nccl_rc) return ctx->fail(BGPU_ENCCL, std::string("ncclSend/ncclRecv: ") + nc.GetErrorString(nccl_rc));
    if (end_rc) return ctx->fail(BGPU_ENCCL, std::string("ncclGroupEnd: ") + nc.GetErrorString(end_rc));
  }
  // ---- 4. re-pack validity, rebuild utf8 offsets --------------------------------------------------------------------------
  for (const Post& post : posts) {
    Column& oc = ob->cols[post.col];
    if (post.vbytes) {
      DevAllocPtr bits;
      BGPU_TRY(dev_alloc(ctx, (size_t)ceil_div64(std::max<int64_t>(n_out, 1), 32) * 4, &bits));
      if (n_out) {
        bytes_to_bits_kernel<<<grid_for(ctx, n_out), kThreads, 0, st>>>((const uint8_t*)post.vbytes->p, n_out, (uint32_t*)bits->p);
        ctx->launches++;
      }
      oc.validity = (uint8_t*)bits->p;
      oc.owned.push_back(bits);
    }
    if (post.lens) {
      BGPU_CUDA(ctx, cudaMemsetAsync(oc.offsets, 0, 4, st));
      if (n_out) {
        size_t tmp_bytes = 0;
        cub::DeviceScan::InclusiveSum(nullptr, tmp_bytes, (const int32_t*)post.lens->p, oc.offsets + 1, (int)n_out, st);
        DevAllocPtr tmp;
        BGPU_TRY(dev_alloc(ctx, tmp_bytes, &tmp));
        BGPU_CUDA(ctx, cub::DeviceScan::InclusiveSum(tmp->p, tmp_bytes, (const int32_t*)post.lens->p, oc.offsets + 1, (int)n_out, st));
        ctx->launches++;
      }
    }
  }
  BGPU_CUDA(ctx, cudaGetLastError());
  // sources and staging buffers must outlive the exchange on the stream
  BGPU_CUDA(ctx, cudaStreamSynchronize(st));
  *out = ob.release();
  return BGPU_OK;
} catch (const std::exception& e) {
  return bgpu::abi_guard(ctx, e);  // nothing is thrown across the C ABI
}
