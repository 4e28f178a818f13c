// Fused hash repartition + exchange over NVLink peer memory (K3 + K6 in one kernel). Part of partition.cu's
// translation unit (it reuses the count / scan kernels and the tile ranking of the scatter).
//
// Replaces, inside one NVSwitch node, BatchPartitioner::partition + shuffle file write + Flight do_get + reader fan-in
// (/root/reference/ballista/core/src/execution_plans/shuffle_writer.rs:336-391,
// /root/reference/ballista/executor/src/flight_service.rs:124-166,
// /root/reference/ballista/core/src/execution_plans/shuffle_reader.rs:539-624) for batches of fixed-width columns
// (aggregate states): the scatter kernel ranks a tile by partition exactly as K3 does and then stores every run
// STRAIGHT INTO THE OWNER RANK'S RECEIVE WINDOW (cudaIpc-mapped peer memory, plain 4/8/16-byte stores over NVLink), so
// rows cross the link once, while the next tile is being ranked - no partitioned copy in local HBM, no send/recv staging.
// The result is identical to bgpu_all_to_all(bgpu_hash_partition(batch)): rows grouped by (owned partition ascending,
// source rank ascending), input order kept inside each group.
//
// Protocol (every rank, on its ctx stream; NCCL is used for the two tiny collectives only):
//   1. count_kernel + scan_kernel: rows per partition of the local batch and base[p][chunk]
//   2. ncclAllGather of the (n + 1) partition offsets + a "can fuse" flag -> every rank knows every count, so every rank
//      computes every receive layout (bgpu_exchange_plan) and the same decisions (fallback, window growth)
//      [this collective also orders the peers' stores after the owner's copy-out of the previous exchange]
//   3. base[p][chunk] += (first row of (p, this rank) in the owner's window) - offsets[p]
//   4. scatter_kernel<.., PEER = true> (partition.cu): out pointer of a run = window of owner(p) + column region + row * width
//   5. ncclAllGather of one word: all ranks' stores have landed
//   6. copy the window's column regions out into the result batch (local HBM copy; the window is reused)
#include "nccl_dl.h"

namespace bgpu {

constexpr int kMaxPeers = 16;

// base[p][chunk] += delta[p] (rows of partition p start at a different place in the owner's window)
__global__ void add_delta_kernel(uint64_t* __restrict__ base, const int64_t* __restrict__ delta, uint32_t n, uint32_t g) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n * g) base[i] += (uint64_t)delta[i / g];
}

struct IpcHandle { char reserved[64]; };

}  // namespace bgpu

namespace bgpu {
// A receive window: cudaMalloc'ed (cudaIpc needs a non-pool allocation), mapped by every peer. A window may be LEASED to the
// result batch of an exchange (its columns point into the window: no copy-out); the lease ends when the batch dies, and a
// window replaced while leased lives on until then.
struct PeerWindow {
  bgpu_ctx* ctx = nullptr;
  void* ptr = nullptr;
  uint64_t bytes = 0;
  std::atomic<int> leased{0};
  ~PeerWindow() {
    if (ptr) {
      cudaSetDevice(ctx->device);
      cudaFree(ptr);  // (synchronises the device: nothing still reads the window)
    }
    if (ctx) ctx_release(ctx);
  }
};
struct WindowLease {
  std::shared_ptr<PeerWindow> w;
  ~WindowLease() { w->leased.store(0, std::memory_order_release); }
};
constexpr int kPeerWindows = 2;
}  // namespace bgpu

struct bgpu_peer_exchange {
  bgpu_ctx* ctx = nullptr;
  void* comm = nullptr;
  int32_t rank = 0, world = 1;
  std::shared_ptr<bgpu::PeerWindow> win[bgpu::kPeerWindows];      // this rank's receive windows (the second one is made on demand)
  void* peer[bgpu::kPeerWindows][bgpu::kMaxPeers] = {};          // peer[w][d] = rank d's window w mapped here (nullptr: none)
  uint64_t exchanges = 0, fused = 0, bytes_sent = 0, leased_results = 0;
};

namespace bgpu {

static int peer_close(bgpu_peer_exchange* x) {
  for (int w = 0; w < kPeerWindows; w++)
    for (int d = 0; d < x->world; d++) {
      if (d != x->rank && x->peer[w][d]) cudaIpcCloseMemHandle(x->peer[w][d]);
      x->peer[w][d] = nullptr;
    }
  return BGPU_OK;
}

// one word from every rank: a barrier on the stream (and the host, after the sync)
static int peer_barrier(bgpu_peer_exchange* x) {
  bgpu_ctx* ctx = x->ctx;
  DevAllocPtr a, b;
  BGPU_TRY(dev_alloc(ctx, 8, &a));
  BGPU_TRY(dev_alloc(ctx, 8 * (size_t)x->world, &b));
  BGPU_NCCL(ctx, nccl().AllGather(a->p, b->p, 8, kNcclUint8, x->comm, ctx->stream));
  return BGPU_OK;
}

// Map every peer's windows (collective; nothing is mapped when it is called).
static int peer_map(bgpu_peer_exchange* x) {
  bgpu_ctx* ctx = x->ctx;
  for (int w = 0; w < kPeerWindows; w++) x->peer[w][x->rank] = x->win[w] ? x->win[w]->ptr : nullptr;
  if (x->world == 1) return BGPU_OK;
  struct Mine { IpcHandle h[kPeerWindows]; uint64_t has[kPeerWindows]; } mine;
  memset(&mine, 0, sizeof mine);
  for (int w = 0; w < kPeerWindows; w++)
    if (x->win[w]) {
      BGPU_CUDA(ctx, cudaIpcGetMemHandle((cudaIpcMemHandle_t*)&mine.h[w], x->win[w]->ptr));
      mine.has[w] = 1;
    }
  DevAllocPtr d_mine, d_all;
  BGPU_TRY(dev_alloc(ctx, sizeof mine, &d_mine));
  BGPU_TRY(dev_alloc(ctx, sizeof mine * (size_t)x->world, &d_all));
  BGPU_CUDA(ctx, cudaMemcpyAsync(d_mine->p, &mine, sizeof mine, cudaMemcpyHostToDevice, ctx->stream));
  BGPU_NCCL(ctx, nccl().AllGather(d_mine->p, d_all->p, sizeof mine, kNcclUint8, x->comm, ctx->stream));
  std::vector<Mine> all((size_t)x->world);
  BGPU_CUDA(ctx, cudaMemcpyAsync(all.data(), d_all->p, sizeof mine * (size_t)x->world, cudaMemcpyDeviceToHost, ctx->stream));
  BGPU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  for (int d = 0; d < x->world; d++) {
    if (d == x->rank) continue;
    for (int w = 0; w < kPeerWindows; w++) {
      if (!all[(size_t)d].has[w]) continue;
      cudaIpcMemHandle_t h;
      memcpy(&h, &all[(size_t)d].h[w], sizeof h);
      const cudaError_t e = cudaIpcOpenMemHandle(&x->peer[w][d], h, cudaIpcMemLazyEnablePeerAccess);
      if (e != cudaSuccess) {
        x->peer[w][d] = nullptr;
        return ctx->fail(BGPU_ECUDA, std::string("cudaIpcOpenMemHandle(rank ") + std::to_string(d) + "): " + cudaGetErrorString(e));
      }
    }
  }
  return BGPU_OK;
}

// Replace this rank's window w by one of `bytes` (0: keep it) and map everybody's windows again. Collective: every rank
// calls it (each with its own w / bytes). A window that is leased to a result batch stays alive until that batch dies.
static int peer_regrow(bgpu_peer_exchange* x, int w, uint64_t bytes) {
  bgpu_ctx* ctx = x->ctx;
  BGPU_CUDA(ctx, cudaSetDevice(ctx->device));
  // nobody may still hold a mapping of a window that is about to be freed
  peer_close(x);
  if (x->world > 1) BGPU_TRY(peer_barrier(x));
  BGPU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  if (bytes) {
    bytes = (bytes + (2u << 20) - 1) & ~(uint64_t)((2u << 20) - 1);
    x->win[w] = nullptr;  // freed now, or when its lease ends
    auto nw = std::make_shared<PeerWindow>();
    nw->ctx = ctx;
    ctx_retain(ctx);
    BGPU_CUDA(ctx, cudaMalloc(&nw->ptr, bytes));
    nw->bytes = bytes;
    x->win[w] = nw;
  }
  return peer_map(x);
}

}  // namespace bgpu

extern "C" int bgpu_peer_exchange_create(bgpu_ctx* ctx, void* nccl_comm, int32_t rank, int32_t world, uint64_t window_bytes,
                                         bgpu_peer_exchange** out) try {
  using namespace bgpu;
  if (!ctx || !out || world < 1 || world > kMaxPeers || rank < 0 || rank >= world || (world > 1 && !nccl_comm)) return BGPU_EINVAL;
  *out = nullptr;
  if (!ctx->has_device) return ctx->fail(BGPU_ENODEVICE, "context has no device (libbgpu has no CPU fallback)");
  if (world > 1 && !nccl().ok) return ctx->fail(BGPU_ENCCL, "NCCL unavailable: " + nccl().why);
  bgpu_peer_exchange* x = new bgpu_peer_exchange();
  x->ctx = ctx;
  x->comm = nccl_comm;
  x->rank = rank;
  x->world = world;
  ctx_retain(ctx);
  const int rc = peer_regrow(x, 0, std::max<uint64_t>(window_bytes, 1u << 20));
  if (rc != BGPU_OK) {
    bgpu_peer_exchange_destroy(x);
    return rc;
  }
  *out = x;
  return BGPU_OK;
} catch (const std::exception& e) {
  return bgpu::abi_guard(ctx, e);  // nothing is thrown across the C ABI
}

extern "C" void bgpu_peer_exchange_destroy(bgpu_peer_exchange* x) {
  using namespace bgpu;
  if (!x) return;
  cudaSetDevice(x->ctx->device);
  // not a collective (it also runs at interpreter shutdown): every exchange ends with all peers' stores landed, so no
  // peer touches these windows any more; the driver keeps the memory alive while a peer still has it mapped. A window
  // leased to a result batch outlives the exchange object (the lease holds it).
  peer_close(x);
  cudaStreamSynchronize(x->ctx->stream);
  for (auto& w : x->win) w = nullptr;
  ctx_release(x->ctx);
  delete x;
}

extern "C" uint64_t bgpu_peer_exchange_window_bytes(const bgpu_peer_exchange* x) {
  uint64_t b = 0;
  if (x)
    for (const auto& w : x->win)
      if (w) b = std::max<uint64_t>(b, w->bytes);
  return b;
}

extern "C" int bgpu_partition_exchange(bgpu_peer_exchange* x, const bgpu_batch* in, const int32_t* key_cols, int32_t n_keys,
                                       uint32_t n_partitions, bgpu_batch** out_received, uint64_t* out_recv_off,
                                       uint64_t* out_sent_bytes) try {
  using namespace bgpu;
  if (!x || !in || !key_cols || !out_received) return BGPU_EINVAL;
  bgpu_ctx* ctx = x->ctx;
  *out_received = nullptr;
  if (out_sent_bytes) *out_sent_bytes = 0;
  const uint32_t n = n_partitions;
  const int world = x->world, me = x->rank;
  if (n == 0 || n > kMaxPartitions) return ctx->fail(BGPU_EINVAL, "n_partitions out of range");
  BGPU_CUDA(ctx, cudaSetDevice(ctx->device));
  GpuTimeScope timed(ctx, T_KERNEL);
  const int64_t n_rows = in->n_rows;
  const size_t n_cols = in->cols.size();
  x->exchanges++;

  // can this rank's batch take the fused path? (every rank must agree: the flag travels with the counts)
  uint64_t can_fuse = n_cols > 0 && n_cols <= (size_t)kMaxCols && n <= 128;
  int row_bytes = 0;
  for (const Column& c : in->cols) {
    if (c.kind != K_FIXED || (c.width != 4 && c.width != 8 && c.width != 16) || c.validity ||
        (n_rows && ((uintptr_t)c.values % (size_t)c.width) != 0))
      can_fuse = 0;
    row_bytes += c.width;
  }
  const int peer_ptrs = (int)n_cols * world;
  ScatterGeom geom = scatter_smem_bytes(n, kGeomSmall.threads, kGeomSmall.items, row_bytes, 0, peer_ptrs) <= kTwoCtasBudget ? kGeomSmall : kGeomNarrow;
  const int64_t kTile = (int64_t)geom.threads * geom.items;
  const size_t smem = scatter_smem_bytes(n, geom.threads, geom.items, row_bytes, 0, peer_ptrs);
  if (smem > kTwoCtasBudget) can_fuse = 0;
  // the window this exchange's rows land in: one that no earlier result batch still holds (the largest such)
  static const bool copy_out = getenv("BGPU_PEER_COPY_OUT") != nullptr;  // round 1's behaviour: one window, copied out
  int my_w = -1;
  for (int w = 0; w < kPeerWindows; w++) {
    const bool free_w = !x->win[w] || x->win[w]->leased.load(std::memory_order_acquire) == 0;
    if (!free_w) continue;
    if (my_w < 0 || (x->win[w] ? x->win[w]->bytes : 0) > (x->win[my_w] ? x->win[my_w]->bytes : 0)) my_w = w;
  }
  if (my_w < 0) can_fuse = 0;  // both windows are still leased to live result batches: the two-step path, on every rank
  const uint64_t my_cap = my_w >= 0 && x->win[my_w] ? x->win[my_w]->bytes : 0;
  ScatterArgs args{};
  ScatterLaunch l{&args, nullptr, n, n_rows, 0, nullptr, nullptr, world};
  int occ = 2;
  if (can_fuse) BGPU_TRY(scatter_dispatch(ctx, geom, false, true, false, 0, smem, l, &occ));
  if (occ < 1) occ = 1;

  // 1. local histogram + scan (grid fixed here: the scatter uses the same chunks)
  KeySet ks{};
  std::vector<DevAllocPtr> temps;
  BGPU_TRY(make_keyset(ctx, in, key_cols, n_keys, &ks, &temps));
  const Modulo mod = make_modulo(n);
  int64_t g = (int64_t)ctx->sm_count * occ;
  const int64_t tiles = std::max<int64_t>(1, ceil_div64(n_rows, kTile));
  if (g > tiles) g = tiles;
  const int64_t chunk_rows = ceil_div64(tiles, g) * kTile;
  g = std::max<int64_t>(1, ceil_div64(n_rows, chunk_rows));
  DevAllocPtr hist, base, hdr, all_hdr, pids;
  BGPU_TRY(dev_alloc(ctx, (size_t)std::max<int64_t>(n_rows, 1) * 2, &pids));
  BGPU_TRY(dev_alloc(ctx, (size_t)n * g * 4, &hist));
  BGPU_TRY(dev_alloc(ctx, (size_t)n * g * 8, &base));
  const size_t hdr_words = (size_t)n + 4;  // offsets[n + 1], can_fuse, the receive window chosen, its capacity
  BGPU_TRY(dev_alloc(ctx, hdr_words * 8, &hdr));
  BGPU_TRY(dev_alloc(ctx, hdr_words * 8 * (size_t)world, &all_hdr));
  if (n_rows > 0) {
    count_kernel<<<(int)g, kCountThreads, n * 4, ctx->stream>>>(ks, mod, n_rows, chunk_rows, (uint32_t*)hist->p, nullptr, (uint16_t*)pids->p);
    scan_kernel<<<1, 1024, 0, ctx->stream>>>((const uint32_t*)hist->p, (uint64_t*)base->p, (uint64_t*)hdr->p, n, (uint32_t)g, n_rows);
    ctx->launches += 2;
    BGPU_CUDA(ctx, cudaGetLastError());
  } else {
    BGPU_CUDA(ctx, cudaMemsetAsync(hdr->p, 0, hdr_words * 8, ctx->stream));
  }
  const uint64_t tail[3] = {can_fuse, (uint64_t)std::max(my_w, 0), my_cap};
  BGPU_CUDA(ctx, cudaMemcpyAsync((uint64_t*)hdr->p + n + 1, tail, sizeof tail, cudaMemcpyHostToDevice, ctx->stream));

  // 2. everybody's counts
  std::vector<uint64_t> all(hdr_words * (size_t)world);
  if (world > 1) {
    BGPU_NCCL(ctx, nccl().AllGather(hdr->p, all_hdr->p, hdr_words * 8, kNcclUint8, x->comm, ctx->stream));
    BGPU_CUDA(ctx, cudaMemcpyAsync(all.data(), all_hdr->p, all.size() * 8, cudaMemcpyDeviceToHost, ctx->stream));
  } else {
    BGPU_CUDA(ctx, cudaMemcpyAsync(all.data(), hdr->p, hdr_words * 8, cudaMemcpyDeviceToHost, ctx->stream));
  }
  BGPU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  for (int s = 0; s < world; s++)
    if (!all[(size_t)s * hdr_words + n + 1])
      return ctx->fail(BGPU_EUNSUPPORTED, "fused exchange needs null-free 4/8/16-byte columns on every rank (rank " + std::to_string(s) +
                                              " cannot): use bgpu_hash_partition + bgpu_all_to_all");

  // receive layout of every rank: rows of (owned partition k, source s) start at recv[d][k * world + s]
  std::vector<uint64_t> offs((size_t)world * (n + 1));
  for (int s = 0; s < world; s++) memcpy(&offs[(size_t)s * (n + 1)], &all[(size_t)s * hdr_words], (size_t)(n + 1) * 8);
  std::vector<std::vector<uint64_t>> recv((size_t)world);
  std::vector<uint64_t> total((size_t)world);
  uint64_t my_need = 0;
  bool grow = false;  // some rank's chosen window is too small: every rank sees that in the gathered headers
  std::vector<int> win_of((size_t)world);
  auto region = [&](uint64_t rows, int width) { return (rows * (uint64_t)width + 255) & ~255ULL; };
  for (int d = 0; d < world; d++) {
    const uint32_t owned = (n + (uint32_t)world - 1 - (uint32_t)d) / (uint32_t)world;
    recv[(size_t)d].assign((size_t)owned * world + 1, 0);
    uint32_t got = 0;
    total[(size_t)d] = bgpu_exchange_plan(offs.data(), d, world, n, recv[(size_t)d].data(), &got);
    uint64_t bytes = 0;
    for (const Column& c : in->cols) bytes += region(total[(size_t)d], c.width);
    win_of[(size_t)d] = (int)all[(size_t)d * hdr_words + n + 2];
    if (bytes > all[(size_t)d * hdr_words + n + 3]) grow = true;
    if (d == me) my_need = bytes;
  }
  if (grow) BGPU_TRY(peer_regrow(x, my_w, my_need > my_cap ? my_need + my_need / 4 : 0));  // collective
  PeerWindow* const mine = x->win[my_w].get();

  // 3. window rows for this rank's partitions + the column regions in every owner's window
  std::vector<int64_t> delta(n);
  uint64_t sent_rows = 0;
  for (uint32_t p = 0; p < n; p++) {
    const int d = (int)(p % (uint32_t)world);
    const uint32_t k = p / (uint32_t)world;
    delta[p] = (int64_t)recv[(size_t)d][(size_t)k * world + me] - (int64_t)offs[(size_t)me * (n + 1) + p];
    if (d != me) sent_rows += offs[(size_t)me * (n + 1) + p + 1] - offs[(size_t)me * (n + 1) + p];
  }
  args.n_cols = (int32_t)n_cols;
  std::vector<size_t> order(n_cols);
  for (size_t i = 0; i < n_cols; i++) order[i] = i;
  std::stable_sort(order.begin(), order.end(), [&](size_t a, size_t b) { return in->cols[a].width > in->cols[b].width; });
  std::vector<void*> ptrs(n_cols * (size_t)world);
  std::vector<uint64_t> my_region(n_cols);  // byte offset of column i in this rank's window
  {
    int off = 0;
    for (size_t a = 0; a < n_cols; a++) {
      const size_t i = order[a];
      args.async_cols[a] = AsyncCol{in->cols[i].values, nullptr, in->cols[i].width, off};
      off += in->cols[i].width;
    }
    args.n_async = (int32_t)n_cols;
    args.async_width = off;
    for (int d = 0; d < world; d++) {
      uint64_t at = 0;
      for (size_t i = 0; i < n_cols; i++) {  // regions in schema order
        size_t a = 0;
        while (order[a] != i) a++;
        ptrs[a * (size_t)world + d] = (char*)x->peer[win_of[(size_t)d]][d] + at;
        if (d == me) my_region[i] = at;
        at += region(total[(size_t)d], in->cols[i].width);
      }
    }
  }
  if (n_rows > 0) {
    DevAllocPtr d_delta, d_ptrs;
    BGPU_TRY(dev_alloc(ctx, (size_t)n * 8, &d_delta));
    BGPU_TRY(dev_alloc(ctx, ptrs.size() * 8, &d_ptrs));
    BGPU_CUDA(ctx, cudaMemcpyAsync(d_delta->p, delta.data(), (size_t)n * 8, cudaMemcpyHostToDevice, ctx->stream));
    BGPU_CUDA(ctx, cudaMemcpyAsync(d_ptrs->p, ptrs.data(), ptrs.size() * 8, cudaMemcpyHostToDevice, ctx->stream));
    add_delta_kernel<<<(unsigned)((n * g + 255) / 256), 256, 0, ctx->stream>>>((uint64_t*)base->p, (const int64_t*)d_delta->p, n, (uint32_t)g);
    // 4. the fused scatter: runs go straight into the owners' windows
    l.pids = (const uint16_t*)pids->p; l.chunk_rows = chunk_rows; l.base = (const uint64_t*)base->p;
    l.peer_out = (void* const*)d_ptrs->p;
    BGPU_TRY(scatter_dispatch(ctx, geom, false, true, true, g, smem, l, nullptr));
    ctx->launches += 2;
    BGPU_CUDA(ctx, cudaGetLastError());
    BGPU_CUDA(ctx, cudaStreamSynchronize(ctx->stream));  // delta / ptrs are host temporaries of this call
  }
  // 5. everybody's stores have landed
  if (world > 1) BGPU_TRY(peer_barrier(x));

  // 6. the result: its columns ARE the window's regions (the window is leased to the batch until the batch dies; the next
  //    exchange takes the other window) - or, with BGPU_PEER_COPY_OUT, copies of them as in round 1
  std::unique_ptr<bgpu_batch> out(new bgpu_batch(ctx));
  out->n_rows = (int64_t)total[(size_t)me];
  out->cols.resize(n_cols);
  DevAllocPtr lease;
  if (!copy_out && out->n_rows && mine) {
    mine->leased.store(1, std::memory_order_release);
    auto keep = std::make_shared<WindowLease>();
    keep->w = x->win[my_w];
    lease = std::make_shared<DevAlloc>();  // no memory of its own: it only carries the lease
    lease->owned = false;
    lease->keep = keep;
    x->leased_results++;
  }
  for (size_t i = 0; i < n_cols; i++) {
    const Column& c = in->cols[i];
    Column& o = out->cols[i];
    o.name = c.name; o.format = c.format; o.flags = c.flags; o.kind = c.kind; o.width = c.width;
    o.length = out->n_rows; o.null_count = 0;
    if (lease) {
      o.owned.push_back(lease);
      o.values = (char*)mine->ptr + my_region[i];
      continue;
    }
    DevAllocPtr a;
    BGPU_TRY(dev_alloc(ctx, (size_t)out->n_rows * c.width, &a));
    o.owned.push_back(a);
    o.values = a->p;
    if (out->n_rows)
      BGPU_CUDA(ctx, cudaMemcpyAsync(a->p, (char*)mine->ptr + my_region[i], (size_t)out->n_rows * c.width, cudaMemcpyDeviceToDevice, ctx->stream));
  }
  if (out_recv_off) memcpy(out_recv_off, recv[(size_t)me].data(), recv[(size_t)me].size() * 8);
  if (out_sent_bytes) *out_sent_bytes = sent_rows * (uint64_t)row_bytes;
  x->fused++;
  x->bytes_sent += sent_rows * (uint64_t)row_bytes;
  *out_received = out.release();
  return BGPU_OK;
} catch (const std::exception& e) {
  return bgpu::abi_guard(x ? x->ctx : nullptr, e);  // nothing is thrown across the C ABI
}
