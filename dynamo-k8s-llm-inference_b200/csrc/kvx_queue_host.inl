// kvx_queue_host.inl — host side of the resident receiver (kvx_queue_*), included at the end of kvx.cu.
#include <atomic>
#include <chrono>
#include <deque>
#include <thread>

#include "kvx_queue.cuh"

namespace {

constexpr uint32_t kQueueRing = 256;             // requests in flight per queue
constexpr uint32_t kQueueTableRing = 1u << 20;   // int32 entries of table ring (host pinned + device)
constexpr uint64_t kQueueSubmitTimeoutNs = 10ull * 1000 * 1000 * 1000;

struct QueueSpan {
  uint64_t seq1;
  uint32_t begin, end;
};

uint64_t now_ns() {
  return static_cast<uint64_t>(std::chrono::duration_cast<std::chrono::nanoseconds>(
                                   std::chrono::steady_clock::now().time_since_epoch()).count());
}

}  // namespace

struct kvx_queue {
  kvx_ctx* ctx = nullptr;
  const kvx_pool* src = nullptr;
  const kvx_pool* dst = nullptr;
  kvx::QueueParams qp{};
  uint32_t grid = 0, block = 0;
  size_t smem = 0;
  unsigned char* d_mem = nullptr;
  unsigned char* h_mem = nullptr;
  kvx::QueueHostSlot* slots = nullptr;   // host views of the pinned block
  int* tables = nullptr;
  volatile uint64_t* done = nullptr;
  volatile uint64_t* parked = nullptr;
  volatile uint32_t* stop = nullptr;
  cudaStream_t stream = nullptr;
  uint64_t seq = 0;        // requests submitted so far
  uint32_t tab_cursor = 0;
  std::deque<QueueSpan> spans;
  uint32_t gen = 0;        // launches so far
  uint64_t launches = 0;
};

namespace {

bool queue_slot_done(const kvx_queue* q, uint64_t seq1) { return q->done[(seq1 - 1) & (kQueueRing - 1)] >= seq1; }

int32_t queue_spin_until_done(kvx_queue* q, uint64_t seq1, uint64_t timeout_ns) {
  if (queue_slot_done(q, seq1)) return KVX_OK;
  const uint64_t t0 = now_ns();
  for (uint32_t spins = 0; !queue_slot_done(q, seq1); ++spins) {
    if ((spins & 0x3ffu) == 0x3ffu) {
      if (now_ns() - t0 > timeout_ns) return KVX_ETIMEOUT;
      std::this_thread::yield();
    }
  }
  return KVX_OK;
}

int32_t queue_launch(kvx_queue* q) {
  kvx_ctx* ctx = q->ctx;
  q->qp.gen = ++q->gen;
  kvx::kvx_queue_kernel<8><<<q->grid, q->block, q->smem, q->stream>>>(q->qp);
  KVX_CUDA(ctx, cudaGetLastError());
  q->launches++;
  ctx->launches++;
  return KVX_OK;
}

void queue_free(kvx_queue* q) {
  if (q->stream) cudaStreamDestroy(q->stream);
  if (q->d_mem) cudaFree(q->d_mem);
  if (q->h_mem) cudaFreeHost(q->h_mem);
  (void)cudaGetLastError();
  delete q;
}

// Stop the kernel if it is resident (it parks at once when it sees the stop word) and wait for it to leave.
int32_t queue_quiesce(kvx_queue* q) {
  *q->stop = 1u;
  std::atomic_thread_fence(std::memory_order_seq_cst);
  cudaError_t e = cudaStreamSynchronize(q->stream);
  *q->stop = 0u;
  *q->parked = 1;  // whatever value: non-zero = "no kernel resident", the next submit launches one
  std::atomic_thread_fence(std::memory_order_seq_cst);
  if (e != cudaSuccess) return cuda_fail(q->ctx, e, "cudaStreamSynchronize(queue)");
  return KVX_OK;
}

}  // namespace

extern "C" {

int32_t kvx_queue_start(kvx_ctx* ctx, const kvx_pool* src, const kvx_pool* dst, const kvx_xform* xf, uint32_t max_ctas,
                        uint32_t idle_us, kvx_queue** out) try {
  if (!ctx || !src || !dst || !out) return KVX_EINVAL;
  *out = nullptr;
  if (!same_shape(src->desc, dst->desc)) return KVX_EINVAL;
  if (src->desc.block_tokens != dst->desc.block_tokens) return KVX_EUNSUPPORTED;  // no block-size regrouping in a queue
  uint32_t so, dof, hc, sl;
  int32_t rc = resolve_xform(&src->desc, &dst->desc, xf, &so, &dof, &hc, &sl, src->desc.layout);
  if (rc) return rc;
  DeviceGuard guard(ctx->device);
  if (!guard.ok) return cuda_fail(ctx, cudaGetLastError(), "cudaSetDevice");
  const kvx_pool_desc& s = src->desc;
  const kvx_pool_desc& d = dst->desc;
  const bool remote = !src->local || !dst->local;
  Plan pl{};
  rc = make_plan(s.block_tokens, s.head_dim * s.elem_bytes, hc, ChunkGeom{s.layout, s.heads, so, s.block_tokens, s.head_stride},
                 ChunkGeom{d.layout, d.heads, dof, d.block_tokens, d.head_stride}, plan_stage_bytes(ctx, remote), &pl);
  if (rc) return rc;
  const Geometry g = resolve_geometry(ctx, remote, !src->local, kEngineTma, false, pl.piece_bytes);
  if (static_cast<uint64_t>(g.cons_warps) * 32u * 8u * 16u < pl.piece_bytes) return KVX_EINVAL;  // raise CONSUMER_WARPS
  if (static_cast<uint64_t>(s.regions) * pl.ppc >= (1ull << 31) / KVX_QUEUE_MAX_BLOCKS) return KVX_EUNSUPPORTED;

  kvx_queue* q = new (std::nothrow) kvx_queue();
  if (!q) return KVX_ENOMEM;
  q->ctx = ctx; q->src = src; q->dst = dst;
  q->block = 32u * (3u + g.cons_warps);
  q->smem = static_cast<size_t>(g.stages) * g.stage_bytes;
  auto fail = [&](int32_t code) {
    queue_free(q);
    return code;
  };
  rc = set_smem(ctx, kvx::kvx_queue_kernel<8>, q->smem);
  if (rc) return fail(rc);
  // every CTA of the grid must be resident at once: a request is complete when ALL of them have counted it
  int per_sm = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kvx::kvx_queue_kernel<8>, static_cast<int>(q->block), q->smem) != cudaSuccess ||
      per_sm < 1)
    return fail(cuda_fail(ctx, cudaGetLastError(), "occupancy of the queue kernel"));
  q->grid = std::min<uint32_t>(g.max_ctas, static_cast<uint32_t>(ctx->sms) * std::min<uint32_t>(g.ctas_per_sm, static_cast<uint32_t>(per_sm)));
  if (max_ctas != 0) q->grid = std::min(q->grid, max_ctas);
  if (q->grid == 0) return fail(KVX_EINVAL);

  // pinned host block: [slots][done words][parked][stop][tables]
  const size_t h_slots = sizeof(kvx::QueueHostSlot) * kQueueRing, h_done = sizeof(uint64_t) * kQueueRing, h_words = 64;
  const size_t h_bytes = h_slots + h_done + h_words + sizeof(int) * static_cast<size_t>(kQueueTableRing);
  if (cudaHostAlloc(reinterpret_cast<void**>(&q->h_mem), h_bytes, cudaHostAllocMapped | cudaHostAllocPortable) != cudaSuccess) {
    (void)cudaGetLastError();
    q->h_mem = nullptr;
    return fail(KVX_ENOMEM);
  }
  memset(q->h_mem, 0, h_bytes);
  q->slots = reinterpret_cast<kvx::QueueHostSlot*>(q->h_mem);
  q->done = reinterpret_cast<volatile uint64_t*>(q->h_mem + h_slots);
  q->parked = reinterpret_cast<volatile uint64_t*>(q->h_mem + h_slots + h_done);
  q->stop = reinterpret_cast<volatile uint32_t*>(q->h_mem + h_slots + h_done + 8);
  q->tables = reinterpret_cast<int*>(q->h_mem + h_slots + h_done + h_words);
  *q->parked = 1;  // no kernel resident yet
  unsigned char* h_dev = nullptr;
  if (cudaHostGetDevicePointer(reinterpret_cast<void**>(&h_dev), q->h_mem, 0) != cudaSuccess)
    return fail(cuda_fail(ctx, cudaGetLastError(), "cudaHostGetDevicePointer"));

  // device block: [descriptors][done counters][pub_seq, park, start[2]][tables]
  const size_t d_desc = sizeof(kvx::QueueDesc) * kQueueRing, d_cnt = sizeof(unsigned) * kQueueRing, d_words = 64;
  const size_t d_bytes = d_desc + d_cnt + d_words + sizeof(int) * static_cast<size_t>(kQueueTableRing);
  if (cudaMalloc(reinterpret_cast<void**>(&q->d_mem), d_bytes) != cudaSuccess) {
    (void)cudaGetLastError();
    q->d_mem = nullptr;
    return fail(KVX_ENOMEM);
  }
  if (cudaMemset(q->d_mem, 0, d_desc + d_cnt + d_words) != cudaSuccess) return fail(cuda_fail(ctx, cudaGetLastError(), "cudaMemset"));
  if (cudaStreamCreateWithFlags(&q->stream, cudaStreamNonBlocking) != cudaSuccess)
    return fail(cuda_fail(ctx, cudaGetLastError(), "cudaStreamCreate"));

  kvx::QueueParams& qp = q->qp;
  XferParams& p = qp.x;
  p.src = paged_side(src, nullptr);
  p.dst = paged_side(dst, nullptr);
  p.src.slice_off = pl.s_slice_off; p.src.piece_stride = pl.s_piece_stride; p.src.ratio = 1; p.src.sub_stride = pl.s_sub_stride;
  p.dst.slice_off = pl.d_slice_off; p.dst.piece_stride = pl.d_piece_stride; p.dst.ratio = 1; p.dst.sub_stride = pl.d_sub_stride;
  p.regions = s.regions; p.ppc = pl.ppc; p.piece_bytes = pl.piece_bytes;
  p.s_runs = pl.s_runs; p.s_run_bytes = pl.s_run_bytes; p.s_stride = pl.s_stride;
  p.d_runs = pl.d_runs; p.d_run_bytes = pl.d_run_bytes; p.d_n1 = pl.d_n1; p.d_stride_a = pl.d_stride_a; p.d_stride_b = pl.d_stride_b;
  p.stages = g.stages; p.stage_bytes = g.stage_bytes;
  p.block_major = ((regions_inside_one_block(*src) || regions_inside_one_block(*dst)) && env_or("KVX_BLOCK_MAJOR", 1) != 0) ? 1u : 0u;
  p.data_local = (dst->local && !ctx->force_sys_fence) ? 1u : 0u;
  p.status = ctx->d_status;
  qp.desc = reinterpret_cast<kvx::QueueDesc*>(q->d_mem);
  qp.done_cnt = reinterpret_cast<unsigned*>(q->d_mem + d_desc);
  qp.pub_seq = reinterpret_cast<unsigned long long*>(q->d_mem + d_desc + d_cnt);
  qp.park = qp.pub_seq + 1;
  qp.start = qp.pub_seq + 2;
  qp.tables = reinterpret_cast<int*>(q->d_mem + d_desc + d_cnt + d_words);
  qp.h_slots = reinterpret_cast<const kvx::QueueHostSlot*>(h_dev);
  qp.h_done = reinterpret_cast<unsigned long long*>(h_dev + h_slots);
  qp.h_parked = reinterpret_cast<unsigned long long*>(h_dev + h_slots + h_done);
  qp.h_stop = reinterpret_cast<const unsigned int*>(h_dev + h_slots + h_done + 8);
  qp.h_tables = reinterpret_cast<const int*>(h_dev + h_slots + h_done + h_words);
  qp.ring_mask = kQueueRing - 1;
  qp.items_per_block = s.regions * pl.ppc;
  // the consumers wait on the ring while the queue is idle: keep the idle time far below mbar_wait's 4 s trap
  qp.idle_ns = static_cast<unsigned long long>(std::min<uint32_t>(std::max<uint32_t>(idle_us ? idle_us : 1000u, 10u), 1000000u)) * 1000ull;
  ctx->queues.push_back(q);
  *out = q;
  return KVX_OK;
} KVX_CATCH_ALL

int32_t kvx_queue_submit(kvx_queue* q, const int32_t* h_src_table, const int32_t* h_dst_table, uint32_t n, uint64_t* flag,
                         uint64_t flag_val, uint64_t* h_done, uint64_t* ticket) try {
  if (!q) return KVX_EINVAL;
  if (n > KVX_QUEUE_MAX_BLOCKS) return KVX_ERANGE;
  if (n != 0 && (!h_src_table || !h_dst_table)) return KVX_EINVAL;
  kvx_ctx* ctx = q->ctx;
  for (uint32_t i = 0; i < n; ++i) {
    if (h_src_table[i] < 0 || static_cast<uint64_t>(h_src_table[i]) >= q->src->desc.num_blocks) return KVX_ERANGE;
    if (h_dst_table[i] < 0 || static_cast<uint64_t>(h_dst_table[i]) >= q->dst->desc.num_blocks) return KVX_ERANGE;
  }
  DeviceGuard guard(ctx->device);
  uint64_t done_dev = 0;
  if (h_done) {  // the kernel writes the word itself: it must be pinned (or registered) host memory
    cudaPointerAttributes at{};
    if (cudaPointerGetAttributes(&at, h_done) != cudaSuccess || at.type != cudaMemoryTypeHost || !at.devicePointer) {
      (void)cudaGetLastError();
      return KVX_EINVAL;
    }
    done_dev = reinterpret_cast<uint64_t>(at.devicePointer);
  }
  const uint64_t seq = q->seq;
  // the ring slot is free once the request that used it last has landed
  if (seq >= kQueueRing) {
    int32_t rc = queue_spin_until_done(q, seq - kQueueRing + 1, kQueueSubmitTimeoutNs);
    if (rc) return rc;
  }
  // a span of the table rings that no request in flight still reads
  const uint32_t need = (2u * n + 3u) & ~3u;
  if (q->tab_cursor + need > kQueueTableRing) q->tab_cursor = 0;
  const uint32_t b = q->tab_cursor, e = b + need;
  while (!q->spans.empty()) {
    const QueueSpan& sp = q->spans.front();
    if (queue_slot_done(q, sp.seq1)) {
      q->spans.pop_front();
    } else if (sp.begin < e && b < sp.end) {
      int32_t rc = queue_spin_until_done(q, sp.seq1, kQueueSubmitTimeoutNs);
      if (rc) return rc;
    } else {
      break;  // oldest span still in flight does not overlap; later ones lie behind it in ring order
    }
  }
  for (const QueueSpan& sp : q->spans)  // after a wrap a younger span can overlap too
    if (sp.begin < e && b < sp.end) {
      int32_t rc = queue_spin_until_done(q, sp.seq1, kQueueSubmitTimeoutNs);
      if (rc) return rc;
    }
  if (n != 0) {
    memcpy(q->tables + b, h_src_table, n * sizeof(int));
    memcpy(q->tables + b + n, h_dst_table, n * sizeof(int));
    q->spans.push_back(QueueSpan{seq + 1, b, e});
  }
  q->tab_cursor = e;
  // the ids first, then the slot: each 16-byte chunk carries its payload and a tag that only fits this sequence number
  std::atomic_thread_fence(std::memory_order_release);
  volatile kvx::QueueHostSlot* hs = q->slots + (seq & (kQueueRing - 1));
  const unsigned long long tag = kvx::queue_tag(seq + 1);
  const unsigned long long w0 = static_cast<unsigned long long>(n) | (static_cast<unsigned long long>(b) << 32);
  const unsigned long long w1 = reinterpret_cast<uint64_t>(flag), w2 = flag_val, w3 = done_dev;
  hs->n_tab = w0; hs->tag0 = tag ^ w0;
  hs->flag = w1; hs->tag1 = tag ^ w1;
  hs->flag_val = w2; hs->tag2 = tag ^ w2;
  hs->done = w3; hs->tag3 = tag ^ w3;
  q->seq = seq + 1;
  if (ticket) *ticket = seq + 1;
  // mirror image of the kernel's parking: slot first, full fence, then look whether a kernel is there to see it
  std::atomic_thread_fence(std::memory_order_seq_cst);
  if (*q->parked != 0) {
    *q->parked = 0;
    std::atomic_thread_fence(std::memory_order_seq_cst);
    const int32_t rc = queue_launch(q);
    if (rc) *q->parked = 1;  // nothing is resident: the next submit must try again (this request stays queued)
    return rc;
  }
  return KVX_OK;
} KVX_CATCH_ALL

int32_t kvx_queue_submit_batch(kvx_queue* q, const kvx_batch_item* items, uint32_t count, uint64_t* last_ticket) try {
  if (!q || (count != 0 && !items)) return KVX_EINVAL;
  for (uint32_t k = 0; k < count; ++k) {
    const kvx_batch_item& it = items[k];
    int32_t rc = kvx_queue_submit(q, it.h_src_table, it.h_dst_table, it.n, it.flag, it.flag_val, it.h_done, last_ticket);
    if (rc) return rc;  // requests before k are queued and will be served
  }
  return KVX_OK;
} KVX_CATCH_ALL

int32_t kvx_queue_wait(kvx_queue* q, uint64_t ticket, uint32_t timeout_ms) try {
  if (!q || ticket == 0 || ticket > q->seq) return KVX_EINVAL;
  if (ticket + kQueueRing <= q->seq) return KVX_OK;  // its slot has long been handed on: it had landed by then
  return queue_spin_until_done(q, ticket, static_cast<uint64_t>(timeout_ms ? timeout_ms : 1000u) * 1000000ull);
} KVX_CATCH_ALL

int32_t kvx_queue_poll(kvx_queue* q, uint64_t ticket) try {
  if (!q || ticket == 0 || ticket > q->seq) return KVX_EINVAL;
  if (ticket + kQueueRing <= q->seq) return KVX_OK;
  return queue_slot_done(q, ticket) ? KVX_OK : KVX_ENOTREADY;
} KVX_CATCH_ALL

int32_t kvx_queue_park(kvx_queue* q) try {
  if (!q) return KVX_EINVAL;
  DeviceGuard guard(q->ctx->device);
  if (q->seq != 0) {
    int32_t rc = queue_spin_until_done(q, q->seq, kQueueSubmitTimeoutNs);
    if (rc) return rc;
  }
  return queue_quiesce(q);
} KVX_CATCH_ALL

int32_t kvx_queue_launch_count(const kvx_queue* q, uint64_t* launches, uint32_t* ctas) try {
  if (!q) return KVX_EINVAL;
  if (launches) *launches = q->launches;
  if (ctas) *ctas = q->grid;
  return KVX_OK;
} KVX_CATCH_ALL

int32_t kvx_queue_stop(kvx_queue* q) try {
  if (!q) return KVX_EINVAL;
  kvx_ctx* ctx = q->ctx;
  DeviceGuard guard(ctx->device);
  int32_t rc = KVX_OK;
  if (q->seq != 0) rc = queue_spin_until_done(q, q->seq, kQueueSubmitTimeoutNs);  // drain what was submitted
  const int32_t rc2 = queue_quiesce(q);
  ctx->queues.erase(std::remove(ctx->queues.begin(), ctx->queues.end(), q), ctx->queues.end());
  queue_free(q);
  return rc ? rc : rc2;
} KVX_CATCH_ALL

}  // extern "C"
