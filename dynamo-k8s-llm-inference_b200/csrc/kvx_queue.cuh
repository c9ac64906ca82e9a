// kvx_queue.cuh — the resident receiver: one persistent kernel per P→D edge that takes requests from a queue in
// pinned host memory instead of being launched once per request.
//
// Why: a handoff of a short prompt is bounded by fixed costs, not bytes — the launch, the pipeline fill, and the
// drain plus completion flag at the end (ISL 1k on one GPU: 45 us against 36 us of bytes).  A resident kernel pays
// none of them between back-to-back requests: its CTAs walk from the last item of request k straight into the
// first item of request k+1 with the shared-memory ring still full, and the completion of request k is counted
// off the critical path.
//
// Structure of a CTA (blockDim = 32 * (3 + consumer warps)):
//   warp 0   producer — exactly the producer of kvx_xfer_tma_kernel (table lookups 32 items at a time, cp.async.bulk
//            into the mbarrier ring), in a loop over REQUESTS
//   warp 1   completer — counts this CTA's part of a request as done once the consumers have released the request's
//            last stage (fence + atomic + the last CTA's notifications: ~1.5 us that neither the producer nor the
//            consumers wait for); fed by the producer through a small shared-memory FIFO
//   warp 2   dispatcher — only in CTA 0: polls the host ring (ld.acquire.sys over PCIe), copies a new request's block
//            ids into the device table ring, publishes it to all CTAs with one release store; decides when the kernel
//            parks (see below)
//   warps 3+ consumers — the consumers of kvx_xfer_tma_kernel (shared memory → permuted 16-byte stores), looping
//            until they meet the exit marker
//
// Parking: a kernel that never ends would block every cudaDeviceSynchronize / cudaFree of the process.  After
// `idle_ns` without a request the dispatcher parks the kernel: it writes seq+1 to the host word `parked`, fences,
// and looks at the host ring ONCE MORE; only if that is still empty does it publish the park to the CTAs and exit.
// The host does the mirror image on submit (write the slot, full fence, read `parked`; if set, clear it and launch
// the kernel again on the queue's stream).  Either the kernel sees the slot or the host sees `parked` — or both, and
// then the extra launch starts where the previous one parked (device-resident `start`), finds nothing to do and parks again.
#pragma once
#include "kvx_kernels.cuh"

namespace kvx {

constexpr unsigned kQueueExit = 0xffffffffu;  // stage marker: consumers leave

struct QueueDesc {  // one request as published to the CTAs (device memory)
  unsigned int n, tab_off, total_items, reserved;
  unsigned long long flag, flag_val, done;  // flag / done: addresses to receive flag_val (0 = none)
  unsigned long long seq1;                  // seq + 1 (what the per-slot host completion word receives)
};

// One request as submitted by the host (pinned host memory, one 64-byte line): four 16-byte chunks, each a
// payload word and a tag = queue_tag(seq + 1) ^ payload.  The dispatcher reads a slot with one 16-byte load per
// chunk — eight slots per poll, one PCIe round trip for "is it there" AND "what is it" — and takes a chunk only
// when its tag fits the sequence number it expects, so a chunk caught half-written, or left from the request that
// used the slot a ring lap ago, is never mistaken for the new one, whatever the granularity of the reads.
struct QueueHostSlot {
  unsigned long long n_tab, tag0;   // n | tab_off << 32: blocks; first entry of [src ids n][dst ids n] in the table rings
  unsigned long long flag, tag1;
  unsigned long long flag_val, tag2;
  unsigned long long done, tag3;
};
static_assert(sizeof(QueueHostSlot) == 64, "one slot = one 64-byte line");
__host__ __device__ __forceinline__ unsigned long long queue_tag(unsigned long long seq1) {
  return seq1 * 0x9E3779B97F4A7C15ull;  // odd multiplier: non-zero for every seq1 >= 1, so zeroed memory never validates
}
constexpr unsigned kQueuePollSlots = 8;  // slots looked at per poll (4 lanes each)

struct QueueParams {
  XferParams x;                  // the edge's geometry (plan, strides, region bases); n, total_items, tables per request
  QueueDesc* desc;               // device [ring]
  int* tables;                   // device table ring [table_ring]
  unsigned int* done_cnt;        // device [ring]: CTAs that finished the request in this slot
  unsigned long long* pub_seq;   // device: requests published to the CTAs so far
  unsigned long long* park;      // device: (generation << 40) | seq, set when the launch of that generation parks
  unsigned long long* start;     // device [2]: the launch of generation g starts at start[g & 1] (never written while
                                 // g runs) and leaves where it parked in start[(g + 1) & 1] for generation g + 1
  const QueueHostSlot* h_slots;  // pinned host ring, through its device alias
  const int* h_tables;           // pinned host table ring
  unsigned long long* h_done;    // pinned host [ring]: seq+1 once the request in that slot has landed
  unsigned long long* h_parked;  // pinned host: seq+1 at which the kernel parked (host clears it on relaunch)
  const unsigned int* h_stop;    // pinned host: 1 = park at once
  unsigned int ring_mask, items_per_block;  // ring size - 1; regions * pieces per chunk
  unsigned int gen, pad;
  unsigned long long idle_ns;
};

__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long ld_acquire_gpu(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_gpu(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ void st_relaxed_sys(unsigned long long* p, unsigned long long v) {
  asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// Block ids of the request being served: the device table ring.  It is rewritten while the kernel lives, so the
// loads go to L2 (relaxed, gpu scope), never through the non-coherent path.
struct TabQueue {
  const int* tab;
  __device__ __forceinline__ int get(const XferSide& s, unsigned blk) const {
    int v;
    asm volatile("ld.relaxed.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(tab + s.tab_off + blk));
    return v;
  }
};

// ---- dispatcher (CTA 0, warp 2) ----------------------------------------------------------------------------
__device__ __forceinline__ void queue_dispatch(const QueueParams& q, unsigned lane) {
  unsigned long long seq = q.start[q.gen & 1u];
  unsigned long long t_idle = global_timer_ns();
  const unsigned grp = lane >> 2, chunk = lane & 3u;
  auto poll = [&](unsigned long long at, unsigned long long& payload) -> unsigned {
    // lanes 4g..4g+3 read the four chunks of slot at+g; returns how many consecutive slots from `at` hold their request
    const unsigned long long* c = reinterpret_cast<const unsigned long long*>(q.h_slots + ((at + grp) & q.ring_mask)) + 2u * chunk;
    unsigned long long tag;
    asm volatile("ld.relaxed.sys.global.v2.u64 {%0,%1}, [%2];" : "=l"(payload), "=l"(tag) : "l"(c) : "memory");
    const unsigned ok = __ballot_sync(0xffffffffu, tag == (queue_tag(at + grp + 1) ^ payload));
    unsigned k = 0;
    while (k < kQueuePollSlots && ((ok >> (4u * k)) & 0xfu) == 0xfu) ++k;
    return k;
  };
  for (;;) {
    unsigned long long payload = 0;
    const unsigned k = poll(seq, payload);
    if (k != 0) {
      // every group's fields, gathered into all of its lanes
      const unsigned g0 = lane & ~3u;
      const unsigned long long n_tab = __shfl_sync(0xffffffffu, payload, g0);
      const unsigned long long flag = __shfl_sync(0xffffffffu, payload, g0 + 1u);
      const unsigned long long val = __shfl_sync(0xffffffffu, payload, g0 + 2u);
      const unsigned long long done = __shfl_sync(0xffffffffu, payload, g0 + 3u);
      // id lists of the k requests, [src n][dst n] padded to 16 bytes by the host: host ring → device ring, same offset;
      // the reads of all requests are in flight together
      for (unsigned r = 0; r < k; ++r) {
        const unsigned long long nt = __shfl_sync(0xffffffffu, n_tab, 4u * r);
        const unsigned n = static_cast<unsigned>(nt), tab_off = static_cast<unsigned>(nt >> 32);
        const unsigned quads = (2u * n + 3u) / 4u;
        const int4* src = reinterpret_cast<const int4*>(q.h_tables + tab_off);
        int4* dst = reinterpret_cast<int4*>(q.tables + tab_off);
        for (unsigned i = lane; i < quads; i += 32u) {
          int4 v;
          asm volatile("ld.relaxed.sys.global.v4.s32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(src + i) : "memory");
          dst[i] = v;
        }
      }
      if (chunk == 0 && grp < k) {
        QueueDesc* d = q.desc + ((seq + grp) & q.ring_mask);
        const unsigned n = static_cast<unsigned>(n_tab);
        d->n = n; d->tab_off = static_cast<unsigned>(n_tab >> 32); d->total_items = n * q.items_per_block;
        d->flag = flag; d->flag_val = val; d->done = done; d->seq1 = seq + grp + 1;
      }
      __threadfence();  // every lane's table and descriptor stores, before the publication
      __syncwarp();
      seq += k;
      if (lane == 0) st_release_gpu(q.pub_seq, seq);
      t_idle = global_timer_ns();
      continue;
    }
    unsigned stop = 0;
    if (lane == 0) stop = *reinterpret_cast<const volatile unsigned int*>(q.h_stop);
    stop = __shfl_sync(0xffffffffu, stop, 0);
    if (stop != 0 || global_timer_ns() - t_idle > q.idle_ns) {
      if (lane == 0) {
        st_release_sys(q.h_parked, seq + 1);
        __threadfence_system();
      }
      __syncwarp();
      const unsigned again = poll(seq, payload);
      if (again != 0 && stop == 0) {  // a request slipped in between the last look and the announcement
        t_idle = global_timer_ns();
        continue;
      }
      if (lane == 0) {
        q.start[(q.gen + 1u) & 1u] = seq;
        __threadfence();
        st_release_gpu(q.park, (static_cast<unsigned long long>(q.gen) << 40) | seq);
      }
      return;
    }
    __nanosleep(100);
  }
}

// A published descriptor, read past L1 (the slot is rewritten while the kernel lives).
__device__ __forceinline__ QueueDesc load_desc_full(const QueueDesc* p) {
  const volatile QueueDesc* v = p;
  QueueDesc d;
  d.n = v->n; d.tab_off = v->tab_off; d.total_items = v->total_items; d.reserved = 0;
  d.flag = v->flag; d.flag_val = v->flag_val; d.done = v->done; d.seq1 = v->seq1;
  return d;
}

// One CTA's share of request `d` is done (all its stores are visible to this thread through the ring's barriers):
// publish them, count, and let the last CTA deliver the request's notifications.
__device__ __forceinline__ void queue_count_done(const QueueParams& q, const QueueDesc& d, unsigned slot) {
  if (q.x.data_local) __threadfence(); else __threadfence_system();
  const unsigned prev = atomicAdd(q.done_cnt + slot, 1u);
  if (prev == gridDim.x - 1) {
    q.done_cnt[slot] = 0u;
    __threadfence_system();
    if (d.flag != 0) st_relaxed_sys(reinterpret_cast<unsigned long long*>(d.flag), d.flag_val);
    if (d.done != 0) st_relaxed_sys(reinterpret_cast<unsigned long long*>(d.done), d.flag_val);
    st_relaxed_sys(q.h_done + slot, d.seq1);
  }
}

template <int KMAX>
__global__ void kvx_queue_kernel(const __grid_constant__ QueueParams q) {
  constexpr int VEC = 16;
  extern __shared__ __align__(1024) unsigned char stage_mem[];
  __shared__ __align__(8) unsigned long long full_bar[kMaxStages];
  __shared__ __align__(8) unsigned long long empty_bar[kMaxStages];
  __shared__ unsigned long long meta_dst[kMaxStages];
  __shared__ unsigned int meta_bytes[kMaxStages];

  const XferParams& p = q.x;
  const unsigned tid = threadIdx.x;
  const unsigned lane = tid & 31u;
  const unsigned warp = tid >> 5;
  const unsigned cons_warps = (blockDim.x >> 5) - 3u;
  const unsigned S = p.stages;
  // producer → completer FIFO: (slot, stage, parity) of every request's last stage in this CTA
  constexpr unsigned kCq = 16, kCqNoStage = 0xffu, kCqExit = 0xffffffffu;
  __shared__ unsigned int cq_slot[kCq], cq_stage[kCq], cq_parity[kCq];
  __shared__ volatile unsigned int cq_tail, cq_head;
  if (tid == 0) {
    for (unsigned s = 0; s < S; ++s) {
      mbar_init(&full_bar[s], 1u);
      mbar_init(&empty_bar[s], cons_warps);
    }
    mbar_fence_init();
    cq_tail = 0u;
    cq_head = 0u;
  }
  __syncthreads();
  const unsigned bid = blockIdx.x;
  const unsigned grid = gridDim.x;

  if (warp == 2) {
    if (bid == 0) queue_dispatch(q, lane);
    return;
  }
  if (warp == 1) {
    // ------------------------------ completer ------------------------------
    for (unsigned head = 0;; ++head) {
      while (cq_tail == head) __nanosleep(64);   // idle until the producer queues a completion (or the exit entry at the park)
      __threadfence_block();
      const unsigned slot = cq_slot[head % kCq], stage = cq_stage[head % kCq], parity = cq_parity[head % kCq];
      if (slot == kCqExit) break;
      if (stage != kCqNoStage) mbar_wait(&empty_bar[stage], parity);  // every consumer warp has stored this CTA's last piece of the request
      if (lane == 0) {
        const QueueDesc d = load_desc_full(q.desc + slot);
        queue_count_done(q, d, slot);
      }
      __syncwarp();
      if (lane == 0) cq_head = head + 1u;
    }
    return;
  }
  if (warp == 0) {
    // ------------------------------ producer: a loop over requests ------------------------------
    // A request's completion goes to the completer as (slot, last stage, that fill's parity).  The completer must see the
    // release of that fill before the stage is filled AGAIN (two phases on, the barrier's parity would read the same):
    // guard[s] remembers the FIFO position of the entry that names stage s, and the producer does not refill s before the
    // completer is past it.
    unsigned cq_pushed = 0;
    unsigned guard[kMaxStages];
#pragma unroll
    for (unsigned k = 0; k < kMaxStages; ++k) guard[k] = 0u;
    auto spin_until_head = [&](unsigned want) {
      if (cq_head >= want) return;
      const unsigned long long t0 = global_timer_ns();
      for (unsigned spins = 1; cq_head < want; ++spins)
        if ((spins & 0xffu) == 0 && global_timer_ns() - t0 > 4000000000ull) __trap();
    };
    auto push = [&](unsigned slot, unsigned stage, unsigned parity) {
      if (cq_pushed >= kCq) spin_until_head(cq_pushed - kCq + 1u);
      if (lane == 0) {
        cq_slot[cq_pushed % kCq] = slot; cq_stage[cq_pushed % kCq] = stage; cq_parity[cq_pushed % kCq] = parity;
        __threadfence_block();
        cq_tail = cq_pushed + 1u;
      }
      ++cq_pushed;
      if (stage != kCqNoStage) guard[stage] = cq_pushed;
    };
    unsigned long long seq = q.start[q.gen & 1u];
    unsigned long long pub_known = 0;  // requests known to be published (saves the poll when several were published at once)
    unsigned s = 0, phase = 0;
    XferParams rp = p;                 // the request being served: the edge's geometry + its n, total_items, table offsets
    const TabQueue tab{q.tables};
    RawItem raw{};
    bool primed = false;               // rp and raw already hold request `seq` (set up while the previous request was issued)
    auto load_desc = [&](unsigned long long at, unsigned& dn, unsigned& dtab, unsigned& dtotal) {
      if (lane == 0) {                 // issued here, consumed by the caller's shuffles later
        const volatile QueueDesc* d = q.desc + (at & q.ring_mask);
        dn = d->n; dtab = d->tab_off; dtotal = d->total_items;
      }
    };
    // Items are dealt to the CTAs round robin ACROSS requests: request k+1's first item goes to the CTA after the one
    // that got request k's last, so no CTA is always the one with the extra item (at 8 blocks per request that extra
    // item is a sixth of a CTA's work).  `rot` = items dealt so far mod grid, `e` = this CTA's place in the current deal.
    unsigned rot = 0, e = bid;
    rp.total_items = 0;
    auto adopt = [&](unsigned dn, unsigned dtab, unsigned dtotal) {
      dn = __shfl_sync(0xffffffffu, dn, 0);
      dtab = __shfl_sync(0xffffffffu, dtab, 0);
      dtotal = __shfl_sync(0xffffffffu, dtotal, 0);
      rot = (rot + rp.total_items % grid) % grid;   // the request that was in rp has been dealt
      e = (bid + grid - rot) % grid;
      rp.n = dn; rp.total_items = dtotal;
      rp.src.tab_off = dtab; rp.dst.tab_off = dtab + dn;
      const unsigned items = e < dtotal ? (dtotal - e + grid - 1u) / grid : 0u;
      if (lane < items) item_load(rp, tab, e + lane * grid, raw);
    };
    for (;;) {
      if (!primed) {
        // wait for request `seq`, or for the park
        bool leave = false;
        while (pub_known <= seq) {
          unsigned long long pub = 0, pk = 0;
          if (lane == 0) {
            pub = ld_acquire_gpu(q.pub_seq);
            pk = ld_acquire_gpu(q.park);
          }
          pub_known = __shfl_sync(0xffffffffu, pub, 0);
          pk = __shfl_sync(0xffffffffu, pk, 0);
          __syncwarp();  // lane 0's acquire orders the other lanes' table and descriptor loads behind the publication
          if (pub_known > seq) break;
          if ((pk >> 40) == q.gen && seq >= (pk & ((1ull << 40) - 1))) { leave = true; break; }
          __nanosleep(100);
        }
        if (leave) break;
        unsigned dn = 0, dtab = 0, dtotal = 0;
        load_desc(seq, dn, dtab, dtotal);
        adopt(dn, dtab, dtotal);
      }
      primed = false;
      const unsigned slot = static_cast<unsigned>(seq & q.ring_mask);
      const unsigned dtotal = rp.total_items;
      const unsigned my_items = e < dtotal ? (dtotal - e + grid - 1u) / grid : 0u;
      const unsigned e_cur = e;   // adopt() moves e on to the next request while this one's last batch is being issued
      unsigned last_stage = 0, last_parity = 0;
      // The chain "is the next request published → its descriptor → its first table lookups" is three dependent L2 round
      // trips; it runs BEHIND this request's own work so that the ring does not run dry between two requests: the
      // descriptor loads are issued here, before the first piece of this request, and consumed in the last batch.
      unsigned ndn = 0, ndtab = 0, ndtotal = 0;
      bool early = false;
      if (my_items != 0) {
        if (pub_known <= seq + 1) {
          unsigned long long pub = 0;
          if (lane == 0) pub = ld_acquire_gpu(q.pub_seq);
          pub_known = __shfl_sync(0xffffffffu, pub, 0);
          __syncwarp();
        }
        early = pub_known > seq + 1;
        if (early) load_desc(seq + 1, ndn, ndtab, ndtotal);
      }
      for (unsigned base = 0; base < my_items; base += 32u) {
        unsigned long long sa = 0, da = 0;
        unsigned bytes = 0;
        if (base + lane < my_items) item_finish(rp, raw, sa, da, bytes);
        const bool last_batch = base + 32u >= my_items;
        bool next_ready = false;
        if (!last_batch) {
          if (base + 32u + lane < my_items) item_load(rp, tab, e_cur + (base + 32u + lane) * grid, raw);
        } else if (early) {
          // every use of rp and raw for THIS request lies behind us (only addresses in registers are issued below): make
          // them the next request's and put its first table lookups in flight under this batch's bulk copies
          adopt(ndn, ndtab, ndtotal);
          primed = true;
        } else {
          // the next request was not there when this one started: look once more, fetch its descriptor under this batch
          if (pub_known <= seq + 1) {
            unsigned long long pub = 0;
            if (lane == 0) pub = ld_acquire_gpu(q.pub_seq);
            pub_known = __shfl_sync(0xffffffffu, pub, 0);
            __syncwarp();
          }
          next_ready = pub_known > seq + 1;
          if (next_ready) load_desc(seq + 1, ndn, ndtab, ndtotal);
        }
        const unsigned cnt = min(32u, my_items - base);
        for (unsigned j = 0; j < cnt; ++j) {
          const unsigned long long src = __shfl_sync(0xffffffffu, sa, j);
          const unsigned long long dst = __shfl_sync(0xffffffffu, da, j);
          const unsigned nbytes = __shfl_sync(0xffffffffu, bytes, j);
          mbar_wait(&empty_bar[s], phase ^ 1u);
          if (guard[s] != 0u) {  // the completer has to be past this stage's last completion entry before the refill
            spin_until_head(guard[s]);
            guard[s] = 0u;
          }
          unsigned char* stage = stage_mem + static_cast<size_t>(s) * p.stage_bytes;
          if (lane == 0) {
            meta_dst[s] = dst;
            meta_bytes[s] = nbytes;
            mbar_arrive_expect_tx(&full_bar[s], nbytes);
          }
          __syncwarp();
          if (nbytes != 0) {
            if (p.s_runs == 1u) {
              if (lane == 0) bulk_g2s(stage, src, nbytes, &full_bar[s]);
            } else {
              for (unsigned x = lane; x < p.s_runs; x += 32u)
                bulk_g2s(stage + x * p.s_run_bytes, src + static_cast<unsigned long long>(x) * p.s_stride, p.s_run_bytes,
                         &full_bar[s]);
            }
          }
          last_stage = s; last_parity = phase;
          if (++s == S) { s = 0; phase ^= 1u; }
        }
        if (next_ready) {
          adopt(ndn, ndtab, ndtotal);
          primed = true;
        }
      }
      // nothing of a request may fall to this CTA (fewer items than CTAs): it still counts
      push(slot, my_items == 0 ? kCqNoStage : last_stage, last_parity);
      ++seq;
    }
    push(kCqExit, kCqNoStage, 0u);
    // exit marker for the consumers
    mbar_wait(&empty_bar[s], phase ^ 1u);
    if (guard[s] != 0u) spin_until_head(guard[s]);
    if (lane == 0) {
      meta_bytes[s] = kQueueExit;
      mbar_arrive_expect_tx(&full_bar[s], 0u);
    }
    return;
  }
  // ------------------------------ consumers ------------------------------
  const unsigned ctid = tid - 96u;
  const unsigned T = blockDim.x - 96u;
  unsigned doff[KMAX];
  const unsigned upr = p.d_run_bytes / VEC;
#pragma unroll
  for (int k = 0; k < KMAX; ++k) {
    const unsigned u = ctid + k * T;
    doff[k] = dst_run_offset(p, u / upr) + (u % upr) * VEC;
  }
  unsigned s = 0, phase = 0;
  for (;;) {
    // an idle queue keeps the consumers here until the park (bounded by idle_ns, far below mbar_wait's trap)
    mbar_wait(&full_bar[s], phase);
    const unsigned nbytes = meta_bytes[s];
    if (nbytes == kQueueExit) break;
    const unsigned long long dst = meta_dst[s];
    const unsigned units = nbytes / VEC;
    const unsigned char* stage = stage_mem + static_cast<size_t>(s) * p.stage_bytes;
    Vec<VEC> v[KMAX];
#pragma unroll
    for (int k = 0; k < KMAX; ++k) {
      const unsigned u = ctid + k * T;
      if (u < units) lds_vec(v[k], stage + static_cast<size_t>(u) * VEC);
    }
#pragma unroll
    for (int k = 0; k < KMAX; ++k) {
      const unsigned u = ctid + k * T;
      if (u < units) stg_stream(dst + doff[k], v[k]);
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(&empty_bar[s]);
    if (++s == S) { s = 0; phase ^= 1u; }
  }
}

}  // namespace kvx
