// kvx_kernels.cuh — sm_100a device code of the KV handoff path.
//
// One generic "transfer" (xfer) moves work items from a source side to a destination side.  A work
// item is a PIECE of one CHUNK; a chunk is one block of one region (K or V of one layer) of one
// rank: N tokens x H heads x D x E bytes.  Each side is either PAGED (address = region base +
// block id x stride, the block id looked up in the request's block table) or CONTIGUOUS (a staging
// image [regions][n][chunk]).  gather = paged→contiguous, scatter = contiguous→paged, fused
// handoff = paged→paged (the destination may be a peer GPU's memory mapped over NVLink), copy =
// contiguous→contiguous.  Layout transform and TP head slicing are expressed as run descriptors
// (see XferParams); the kernels contain no arithmetic on the payload.
//
// Two engines:
//   kvx_xfer_tma_kernel — persistent CTAs, one producer warp feeding an mbarrier ring of
//     shared-memory stages with cp.async.bulk (TMA bulk copy), consumer warps draining each stage
//     either with one cp.async.bulk store per destination run or with 16/32-byte vector stores.
//   kvx_xfer_ldg_kernel — persistent CTAs, 16/32-byte vector loads straight to vector stores,
//     offsets hoisted into registers.
// Both resolve block tables a warp at a time (32 items per lookup, broadcast with shuffles).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace kvx {

constexpr int kMaxStages = 8;
constexpr unsigned kStatusBadBlockId = 1u;
constexpr unsigned kStatusWaitTimeout = 2u;

struct XferSide {
  const unsigned long long* bases;  // paged: device array [regions] of region base addresses; null = contiguous
  const int* table;                 // paged: device array [n] of block ids
  unsigned long long base0;         // contiguous: base address of the image
  unsigned long long block_stride;  // paged: bytes per block id; contiguous: bytes per (region, i) chunk
  unsigned long long num_blocks;    // paged: ids must be < num_blocks
  unsigned int slice_off;           // byte offset of the moved slice inside a chunk
  unsigned int piece_stride;        // bytes between consecutive pieces of one chunk
  unsigned int ratio;               // paged: granules per block on this side (block-size regrouping), >= 1
  unsigned int sub_stride;          // paged: bytes between consecutive granules inside one block
  unsigned int tab_off;             // paged, tables carried in the kernel parameters: first entry of this side's ids
  unsigned int reserved;
};

struct XferParams {
  XferSide src, dst;
  unsigned int total_items;  // regions * n * ppc
  unsigned int n, regions, ppc;
  unsigned int piece_bytes;  // bytes moved per item
  unsigned int tail_bytes;   // bytes of the very last item (copy mode), 0 = piece_bytes
  // Source runs of one piece: run x (< s_runs) starts at x * s_stride, s_run_bytes long; the piece
  // image in shared memory / in item order is the runs back to back.
  unsigned int s_runs, s_run_bytes, s_stride;
  // Destination runs of one piece: run x (< d_runs) covers image bytes [x*d_run_bytes, +d_run_bytes)
  // and lands at (x / d_n1) * d_stride_a + (x % d_n1) * d_stride_b.
  unsigned int d_runs, d_run_bytes, d_n1, d_stride_a, d_stride_b;
  unsigned int stages, stage_bytes;
  unsigned int block_major;  // 1: items run block by block (all regions of block i, then block i+1) instead of region by region
  unsigned int l2_hint;
  unsigned int data_local;   // 1: every destination byte lands in THIS GPU's memory → a device-scope fence per CTA publishes it
  unsigned int flag_local;   // 1: the flag lives on this GPU too and nobody outside it observes it → device scope to the end
  unsigned long long* flag;  // completion flag (local or peer memory) or null
  unsigned long long* flag2; // second completion word (e.g. the caller's pinned host memory) or null
  unsigned long long flag_val;
  // Batched requests (kvx_handoff_batch): notify_count (address, value) pairs of 64-bit words each, stored
  // in the parameter block's table area from int32 entry notify_off (even), all written by the last CTA.
  unsigned int notify_count, notify_off;
  unsigned int* done_counter;  // scratch: CTAs finished (zero on entry, zero again on exit)
  unsigned int* status;        // scratch: sticky error bits
  // Tensor-map engine (kvx_xfer_tmap_kernel): one CUtensorMap per destination region over [block][N][H][row],
  // the permutation HND → NHD done by the TMA unit: one box store per head of a piece.
  const void* dst_tmaps;       // device array [regions] of CUtensorMap (128 bytes each), or null
  const void* src_tmaps;       // the same for an NHD SOURCE pool (kvx_xfer_tmapld_kernel: NHD → HND, permutation on the load side)
  unsigned int tm_head_bytes;  // bytes of one head's slab in the shared-memory image: N * row
  unsigned int tm_head0;       // first destination head of piece 0 (TP slot); piece k starts at tm_head0 + k * heads per piece
};

// ----------------------------------------------------------------------------------------------
// PTX wrappers
// ----------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(unsigned long long* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(unsigned long long* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive(unsigned long long* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(unsigned long long* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
// Bounded wait: a pipeline bug must trap, not hang the GPU.
__device__ __forceinline__ unsigned long long global_timer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, uint32_t parity) {
  if (mbar_try_wait(bar, parity)) return;
  const unsigned long long t0 = global_timer_ns();
  for (uint32_t spins = 1; !mbar_try_wait(bar, parity); ++spins) {
    if ((spins & 0xffu) == 0 && global_timer_ns() - t0 > 4000000000ull) __trap();  // 4 s: never a hang
  }
}
__device__ __forceinline__ unsigned long long l2_policy_evict_first() {
  unsigned long long pol;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}
// TMA bulk copy global → shared, completion counted in bytes on an mbarrier.
__device__ __forceinline__ void bulk_g2s(void* smem_dst, unsigned long long gsrc, uint32_t bytes,
                                         unsigned long long* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
          smem_u32(smem_dst)),
      "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}
__device__ __forceinline__ void bulk_g2s_hint(void* smem_dst, unsigned long long gsrc, uint32_t bytes,
                                              unsigned long long* bar, unsigned long long pol) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], "
      "%4;" ::"r"(smem_u32(smem_dst)),
      "l"(gsrc), "r"(bytes), "r"(smem_u32(bar)), "l"(pol)
      : "memory");
}
// TMA bulk copy shared → global (local HBM or a peer's, the address decides).
__device__ __forceinline__ void bulk_s2g(unsigned long long gdst, const void* smem_src, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst),
               "r"(smem_u32(smem_src)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
// TMA tensor store shared → global through a 4-d tensor map (coordinates innermost first): the box's rows leave
// shared memory densely packed and land wherever the map's strides put them.
__device__ __forceinline__ void tensor_s2g_4d(const void* tmap, const void* smem_src, int c0, int c1, int c2, int c3) {
  asm volatile("cp.async.bulk.tensor.4d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5}], [%1];" ::"l"(
                   reinterpret_cast<unsigned long long>(tmap)),
               "r"(smem_u32(smem_src)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
               : "memory");
}
// TMA tensor load global → shared through a 4-d tensor map, completion counted in bytes on an mbarrier (one CTA's
// own shared memory; never multicast).
__device__ __forceinline__ void tensor_g2s_4d(void* smem_dst, const void* tmap, int c0, int c1, int c2, int c3, unsigned long long* bar) {
  asm volatile("cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(
                   smem_u32(smem_dst)),
               "l"(reinterpret_cast<unsigned long long>(tmap)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
               : "memory");
}
template <int N>
__device__ __forceinline__ void bulk_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
template <int N>
__device__ __forceinline__ void bulk_wait() {
  asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory");
}

// Programmatic dependent launch: a kernel launched with the stream-serialization attribute may become
// resident while the previous kernel on the stream is still draining; nothing it reads or writes in
// global memory may be touched before grid_dep_wait() returns (= the previous grid has completed and
// its writes are visible).  Both are no-ops for a normally launched kernel.
__device__ __forceinline__ void grid_dep_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void grid_dep_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

template <int VEC>
struct Vec;
template <>
struct Vec<16> {
  uint32_t v[4];
};
template <>
struct Vec<32> {
  uint32_t v[8];
};

__device__ __forceinline__ void ldg_stream(Vec<16>& r, unsigned long long a) {
  asm volatile("ld.global.nc.L1::no_allocate.v4.b32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.v[0]), "=r"(r.v[1]), "=r"(r.v[2]), "=r"(r.v[3])
               : "l"(a));
}
__device__ __forceinline__ void ldg_stream(Vec<32>& r, unsigned long long a) {
  asm volatile("ld.global.nc.L1::no_allocate.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(r.v[0]), "=r"(r.v[1]), "=r"(r.v[2]), "=r"(r.v[3]), "=r"(r.v[4]), "=r"(r.v[5]),
                 "=r"(r.v[6]), "=r"(r.v[7])
               : "l"(a));
}
__device__ __forceinline__ void stg_stream(unsigned long long a, const Vec<16>& r) {
  asm volatile("st.global.L1::no_allocate.v4.b32 [%0], {%1,%2,%3,%4};" ::"l"(a), "r"(r.v[0]), "r"(r.v[1]),
               "r"(r.v[2]), "r"(r.v[3])
               : "memory");
}
__device__ __forceinline__ void stg_stream(unsigned long long a, const Vec<32>& r) {
  asm volatile("st.global.L1::no_allocate.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(a), "r"(r.v[0]),
               "r"(r.v[1]), "r"(r.v[2]), "r"(r.v[3]), "r"(r.v[4]), "r"(r.v[5]), "r"(r.v[6]), "r"(r.v[7])
               : "memory");
}
__device__ __forceinline__ void lds_vec(Vec<16>& r, const unsigned char* s) {
  const uint4 t = *reinterpret_cast<const uint4*>(s);
  r.v[0] = t.x; r.v[1] = t.y; r.v[2] = t.z; r.v[3] = t.w;
}
__device__ __forceinline__ void lds_vec(Vec<32>& r, const unsigned char* s) {
  const uint4 t0 = *reinterpret_cast<const uint4*>(s);
  const uint4 t1 = *reinterpret_cast<const uint4*>(s + 16);
  r.v[0] = t0.x; r.v[1] = t0.y; r.v[2] = t0.z; r.v[3] = t0.w;
  r.v[4] = t1.x; r.v[5] = t1.y; r.v[6] = t1.z; r.v[7] = t1.w;
}

// ----------------------------------------------------------------------------------------------
// Item decode: item w → (piece, i, region) → source / destination byte address.  A block id outside
// [0, num_blocks) yields bytes == 0 (the item is skipped) and raises kStatusBadBlockId.
// The decode is split in two so that the producer can have the NEXT batch's table and base loads in
// flight while it issues the current batch: item_load only issues loads (all four back to back, no
// branch between them), item_finish consumes them.
// ----------------------------------------------------------------------------------------------
struct RawSide {
  unsigned long long base;
  int id;
};
struct RawItem {
  RawSide s, d;
  unsigned w, r, i, piece;
};

// Where the block ids come from: device memory (the caller's tables), or the kernel's own parameter
// block (kvx_handoff_host copies short host lists there, so no H2D copy precedes the launch).
struct TabGlobal {
  __device__ __forceinline__ int get(const XferSide& s, unsigned blk) const {
    return s.table ? __ldg(s.table + blk) : static_cast<int>(blk);
  }
  __device__ __forceinline__ void notify(unsigned, unsigned, unsigned) const {}  // batches always carry their tables in the parameters
};
struct TabParam {
  const int* tab;  // generic address of the table array inside the __grid_constant__ parameter
  // A generic load, not an indexed constant-bank read: the 32 lanes of a lookup want 32 different
  // entries, which the constant cache would serve one address at a time.
  __device__ __forceinline__ int get(const XferSide& s, unsigned blk) const {
    int v;
    asm volatile("ld.s32 %0, [%1];" : "=r"(v) : "l"(tab + s.tab_off + blk));
    return v;
  }
  // Store every batched request's completion word: the lanes of one warp share the list, each store a
  // release at system scope (ordered after everything the last CTA observed before it got here).
  __device__ __forceinline__ void notify(unsigned count, unsigned off, unsigned lane) const {
    const unsigned long long* rec = reinterpret_cast<const unsigned long long*>(tab + off);
    for (unsigned k = lane; k < count; k += 32u) {
      unsigned long long addr, val;
      asm volatile("ld.u64 %0, [%1];" : "=l"(addr) : "l"(rec + 2 * k));
      asm volatile("ld.u64 %0, [%1];" : "=l"(val) : "l"(rec + 2 * k + 1));
      asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(addr), "l"(val) : "memory");
    }
  }
};

template <class Tab>
__device__ __forceinline__ void side_load(const XferSide& s, const Tab& tab, unsigned r, unsigned i, RawSide& out) {
  out.id = 0;
  out.base = 0;
  if (s.bases != nullptr) {
    out.id = tab.get(s, i / s.ratio);
    out.base = __ldg(s.bases + r);
  }
}

__device__ __forceinline__ unsigned long long side_finish(const XferSide& s, const RawSide& raw, unsigned r, unsigned i,
                                                          unsigned piece, unsigned n, bool& ok) {
  unsigned long long a;
  if (s.bases != nullptr) {
    if (raw.id < 0 || static_cast<unsigned long long>(raw.id) >= s.num_blocks) {
      ok = false;
      return 0;
    }
    a = raw.base + static_cast<unsigned long long>(raw.id) * s.block_stride +
        static_cast<unsigned long long>(i % s.ratio) * s.sub_stride;
  } else {
    a = s.base0 + (static_cast<unsigned long long>(r) * n + i) * s.block_stride;
  }
  return a + s.slice_off + static_cast<unsigned long long>(piece) * s.piece_stride;
}

template <class Tab>
__device__ __forceinline__ void item_load(const XferParams& p, const Tab& tab, unsigned w, RawItem& raw) {
  raw.w = w;
  raw.piece = w % p.ppc;
  const unsigned c = w / p.ppc;
  if (p.block_major) {  // cross-layer pools: a block's regions are adjacent in memory, visit them together
    raw.r = c % p.regions;
    raw.i = c / p.regions;
  } else {
    raw.i = c % p.n;
    raw.r = c / p.n;
  }
  side_load(p.src, tab, raw.r, raw.i, raw.s);
  side_load(p.dst, tab, raw.r, raw.i, raw.d);
}

__device__ __forceinline__ void item_finish(const XferParams& p, const RawItem& raw, unsigned long long& sa,
                                            unsigned long long& da, unsigned& bytes) {
  bool ok = true;
  sa = side_finish(p.src, raw.s, raw.r, raw.i, raw.piece, p.n, ok);
  da = side_finish(p.dst, raw.d, raw.r, raw.i, raw.piece, p.n, ok);
  bytes = (p.tail_bytes != 0 && raw.w == p.total_items - 1) ? p.tail_bytes : p.piece_bytes;
  if (!ok) {
    bytes = 0;
    atomicOr(p.status, kStatusBadBlockId);
  }
}

__device__ __forceinline__ unsigned dst_run_offset(const XferParams& p, unsigned x) {
  return (x / p.d_n1) * p.d_stride_a + (x % p.d_n1) * p.d_stride_b;
}

// Completion (H6): every CTA publishes its stores, the last one to finish stores the flag(s) and
// re-zeroes the counter for the next launch that gets this counter slot.  Scopes: the per-CTA fence is
// device-wide when the data lands in this GPU's own memory (its L2 is the point of coherence for
// every observer, local or remote) and system-wide when the stores crossed NVLink; the last CTA's
// fence is system-wide unless the flag itself stays on this GPU.
template <class Tab>
__device__ __forceinline__ void finish_and_flag(const XferParams& p, const Tab& tab) {
  if (p.flag == nullptr && p.flag2 == nullptr && p.notify_count == 0u) return;
  // bar.sync orders every thread's stores before thread 0's fence, which is cumulative (the pattern
  // cooperative-groups grid sync uses): one fence per CTA instead of one per thread.
  __syncthreads();
  if (threadIdx.x < 32u) {  // warp 0: lane 0 counts, all lanes share a batch's notifications
    unsigned last = 0u;
    if (threadIdx.x == 0) {
      if (p.data_local) __threadfence(); else __threadfence_system();
      const unsigned prev = atomicAdd(p.done_counter, 1u);
      if (prev == gridDim.x - 1) {
        last = 1u;
        *p.done_counter = 0u;
        if (p.flag_local) {
          // data and flag both live in this GPU's memory: every observer, local or remote, reads them
          // through this GPU's L2, so one device-scope release orders them for all of them
          asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(p.flag), "l"(p.flag_val) : "memory");
        } else {
          // one system-wide fence (acquire for the other CTAs' counts, release for everything they
          // published), then plain system-scope stores: they cannot pass the fence
          __threadfence_system();
          if (p.flag != nullptr)
            asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p.flag), "l"(p.flag_val) : "memory");
          if (p.flag2 != nullptr)
            asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p.flag2), "l"(p.flag_val) : "memory");
        }
      }
    }
    if (p.notify_count != 0u) {
      last = __shfl_sync(0xffffffffu, last, 0);  // also orders lane 0's fence before the other lanes' stores
      if (last) tab.notify(p.notify_count, p.notify_off, threadIdx.x);
    }
  }
}

// ----------------------------------------------------------------------------------------------
// TMA engine.  blockDim.x = 32 * (1 + consumer warps); warp 0 produces.
// STORE: 16 / 32 = vector stores of that many bytes by the consumer threads; 0 = bulk stores.
// ----------------------------------------------------------------------------------------------
template <int STORE, int KMAX, class Tab>
__device__ __forceinline__ void xfer_tma_body(const XferParams& p, const Tab tab) {
  extern __shared__ __align__(1024) unsigned char stage_mem[];
  __shared__ __align__(8) unsigned long long full_bar[kMaxStages];
  __shared__ __align__(8) unsigned long long empty_bar[kMaxStages];
  __shared__ unsigned long long meta_dst[kMaxStages];
  __shared__ unsigned int meta_bytes[kMaxStages];

  const unsigned tid = threadIdx.x;
  const unsigned lane = tid & 31u;
  const unsigned warp = tid >> 5;
  const unsigned cons_warps = (blockDim.x >> 5) - 1u;
  const unsigned S = p.stages;

  grid_dep_launch_dependents();
  if (tid == 0) {
    for (unsigned s = 0; s < S; ++s) {
      mbar_init(&full_bar[s], 1u);
      mbar_init(&empty_bar[s], cons_warps);
    }
    mbar_fence_init();
  }
  __syncthreads();

  const unsigned bid = blockIdx.x;
  const unsigned grid = gridDim.x;
  const unsigned my_items = bid < p.total_items ? (p.total_items - bid + grid - 1u) / grid : 0u;

  if (warp == 0) {
    // ------------------------------ producer ------------------------------
    const unsigned long long pol = p.l2_hint ? l2_policy_evict_first() : 0ull;
    grid_dep_wait();
    RawItem raw{};
    if (lane < my_items) item_load(p, tab, bid + lane * grid, raw);
    unsigned s = 0, phase = 0;  // ring position and its parity, advanced by hand: no division per item
    for (unsigned base = 0; base < my_items; base += 32u) {
      unsigned long long sa = 0, da = 0;
      unsigned bytes = 0;
      if (base + lane < my_items) item_finish(p, raw, sa, da, bytes);
      // next batch's lookups travel while this batch is issued
      if (base + 32u + lane < my_items) item_load(p, tab, bid + (base + 32u + lane) * grid, raw);
      const unsigned cnt = min(32u, my_items - base);
      for (unsigned j = 0; j < cnt; ++j) {
        const unsigned long long src = __shfl_sync(0xffffffffu, sa, j);
        const unsigned long long dst = __shfl_sync(0xffffffffu, da, j);
        const unsigned nbytes = __shfl_sync(0xffffffffu, bytes, j);
        mbar_wait(&empty_bar[s], phase ^ 1u);
        unsigned char* stage = stage_mem + static_cast<size_t>(s) * p.stage_bytes;
        if (lane == 0) {
          meta_dst[s] = dst;
          meta_bytes[s] = nbytes;
          mbar_arrive_expect_tx(&full_bar[s], nbytes);
        }
        __syncwarp();
        if (nbytes != 0) {
          if (p.s_runs == 1u) {
            if (lane == 0) {
              if (p.l2_hint) bulk_g2s_hint(stage, src, nbytes, &full_bar[s], pol);
              else bulk_g2s(stage, src, nbytes, &full_bar[s]);
            }
          } else {
            for (unsigned x = lane; x < p.s_runs; x += 32u) {
              if (p.l2_hint)
                bulk_g2s_hint(stage + x * p.s_run_bytes, src + static_cast<unsigned long long>(x) * p.s_stride,
                              p.s_run_bytes, &full_bar[s], pol);
              else
                bulk_g2s(stage + x * p.s_run_bytes, src + static_cast<unsigned long long>(x) * p.s_stride,
                         p.s_run_bytes, &full_bar[s]);
            }
          }
        }
        if (++s == S) { s = 0; phase ^= 1u; }
      }
    }
  } else {
    // ------------------------------ consumers -----------------------------
    const unsigned ctid = tid - 32u;
    const unsigned T = blockDim.x - 32u;
    if constexpr (STORE == 0) {
      // One cp.async.bulk store per destination run, runs spread over the consumer threads.
      // A stage is handed back one item late so that two stores stay in flight per thread.
      unsigned prev_stage = 0;
      bool have_prev = false;
      grid_dep_wait();
      unsigned s = 0, phase = 0;
      for (unsigned it = 0; it < my_items; ++it) {
        mbar_wait(&full_bar[s], phase);
        const unsigned long long dst = meta_dst[s];
        const unsigned nbytes = meta_bytes[s];
        const unsigned char* stage = stage_mem + static_cast<size_t>(s) * p.stage_bytes;
        if (nbytes != 0) {
          if (p.d_runs == 1u) {
            if (ctid == 0) bulk_s2g(dst, stage, nbytes);
          } else {
            for (unsigned x = ctid; x < p.d_runs; x += T)
              bulk_s2g(dst + dst_run_offset(p, x), stage + x * p.d_run_bytes, p.d_run_bytes);
          }
        }
        bulk_commit();
        if (have_prev) {
          bulk_wait_read<1>();
          __syncwarp();
          if (lane == 0) mbar_arrive(&empty_bar[prev_stage]);
        }
        prev_stage = s;
        have_prev = true;
        if (++s == S) { s = 0; phase ^= 1u; }
      }
      bulk_wait<0>();  // all bytes written before the completion fence
    } else {
      constexpr int VEC = STORE;
      // Destination offsets of this thread's units are the same for every piece: hoist them.
      unsigned doff[KMAX];
      const unsigned upr = p.d_run_bytes / VEC;  // units per destination run
#pragma unroll
      for (int k = 0; k < KMAX; ++k) {
        const unsigned u = ctid + k * T;
        const unsigned x = u / upr;
        doff[k] = dst_run_offset(p, x) + (u % upr) * VEC;
      }
      grid_dep_wait();
      unsigned s = 0, phase = 0;
      for (unsigned it = 0; it < my_items; ++it) {
        mbar_wait(&full_bar[s], phase);
        const unsigned long long dst = meta_dst[s];
        const unsigned units = meta_bytes[s] / VEC;
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
  }
  finish_and_flag(p, tab);
}

template <int STORE, int KMAX>
__global__ void kvx_xfer_tma_kernel(const __grid_constant__ XferParams p) {
  xfer_tma_body<STORE, KMAX>(p, TabGlobal{});
}

// The same kernel with the request's block ids inside the parameter block (NT int32 entries: the
// source side's ids from tab_off of p.src, the destination side's from tab_off of p.dst).
template <int NT>
struct XferParamsTab {
  XferParams p;
  int tab[NT];
};
template <int STORE, int KMAX, int NT>
__global__ void kvx_xfer_tma_ptab_kernel(const __grid_constant__ XferParamsTab<NT> pt) {
  xfer_tma_body<STORE, KMAX>(pt.p, TabParam{pt.tab});
}

// ----------------------------------------------------------------------------------------------
// Tensor-map engine (HND source → NHD paged destination).  Two warps per CTA: the producer of the TMA engine, and
// one warp that hands every landed piece to the TMA unit again — one cp.async.bulk.tensor store per head, the box
// (row, 1 head, N tokens, 1 block) of the destination region's tensor map.  No thread touches the payload.
// ----------------------------------------------------------------------------------------------
template <class Tab>
__device__ __forceinline__ void xfer_tmap_body(const XferParams& p, const Tab tab) {
  extern __shared__ __align__(1024) unsigned char stage_mem[];
  __shared__ __align__(8) unsigned long long full_bar[kMaxStages];
  __shared__ __align__(8) unsigned long long empty_bar[kMaxStages];
  __shared__ int meta_blk[kMaxStages];
  __shared__ unsigned int meta_reg[kMaxStages], meta_head[kMaxStages], meta_bytes[kMaxStages];

  const unsigned tid = threadIdx.x;
  const unsigned lane = tid & 31u;
  const unsigned warp = tid >> 5;
  const unsigned S = p.stages;
  grid_dep_launch_dependents();
  if (tid == 0) {
    for (unsigned s = 0; s < S; ++s) {
      mbar_init(&full_bar[s], 1u);
      mbar_init(&empty_bar[s], 1u);
    }
    mbar_fence_init();
  }
  __syncthreads();
  const unsigned bid = blockIdx.x;
  const unsigned grid = gridDim.x;
  const unsigned my_items = bid < p.total_items ? (p.total_items - bid + grid - 1u) / grid : 0u;
  const unsigned heads_per_piece = p.piece_bytes / p.tm_head_bytes;

  if (warp == 0) {
    grid_dep_wait();
    RawItem raw{};
    if (lane < my_items) item_load(p, tab, bid + lane * grid, raw);
    unsigned s = 0, phase = 0;
    for (unsigned base = 0; base < my_items; base += 32u) {
      unsigned long long sa = 0, da = 0;
      unsigned bytes = 0;
      if (base + lane < my_items) item_finish(p, raw, sa, da, bytes);
      const int my_blk = raw.d.id;
      const unsigned my_reg = raw.r, my_piece = raw.piece;
      if (base + 32u + lane < my_items) item_load(p, tab, bid + (base + 32u + lane) * grid, raw);
      const unsigned cnt = min(32u, my_items - base);
      for (unsigned j = 0; j < cnt; ++j) {
        const unsigned long long src = __shfl_sync(0xffffffffu, sa, j);
        const unsigned nbytes = __shfl_sync(0xffffffffu, bytes, j);
        const int blk = __shfl_sync(0xffffffffu, my_blk, j);
        const unsigned reg = __shfl_sync(0xffffffffu, my_reg, j);
        const unsigned piece = __shfl_sync(0xffffffffu, my_piece, j);
        mbar_wait(&empty_bar[s], phase ^ 1u);
        unsigned char* stage = stage_mem + static_cast<size_t>(s) * p.stage_bytes;
        if (lane == 0) {
          meta_blk[s] = blk;
          meta_reg[s] = reg;
          meta_head[s] = p.tm_head0 + piece * heads_per_piece;
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
        if (++s == S) { s = 0; phase ^= 1u; }
      }
    }
  } else {
    // a stage is handed back one item late, so that two items' stores stay in flight
    unsigned prev_stage = 0;
    bool have_prev = false;
    grid_dep_wait();
    unsigned s = 0, phase = 0;
    for (unsigned it = 0; it < my_items; ++it) {
      mbar_wait(&full_bar[s], phase);
      const unsigned char* stage = stage_mem + static_cast<size_t>(s) * p.stage_bytes;
      if (meta_bytes[s] != 0) {
        const unsigned char* tm = static_cast<const unsigned char*>(p.dst_tmaps) + static_cast<size_t>(meta_reg[s]) * 128u;
        for (unsigned h = lane; h < heads_per_piece; h += 32u)
          tensor_s2g_4d(tm, stage + static_cast<size_t>(h) * p.tm_head_bytes, 0, static_cast<int>(meta_head[s] + h), 0, meta_blk[s]);
      }
      bulk_commit();
      if (have_prev) {
        bulk_wait_read<1>();
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty_bar[prev_stage]);
      }
      prev_stage = s;
      have_prev = true;
      if (++s == S) { s = 0; phase ^= 1u; }
    }
    bulk_wait<0>();  // every byte written before the completion fence
  }
  finish_and_flag(p, tab);
}

__global__ void __launch_bounds__(64) kvx_xfer_tmap_kernel(const __grid_constant__ XferParams p) {
  xfer_tmap_body(p, TabGlobal{});
}

// The other direction, NHD source → HND destination: the producer fetches every head of a piece with one
// cp.async.bulk.tensor LOAD (box = that head's N rows of one block of the source region), which lands them head-major in
// shared memory; the second warp writes the piece out with plain bulk stores (one per destination run).
__global__ void __launch_bounds__(64) kvx_xfer_tmapld_kernel(const __grid_constant__ XferParams p) {
  extern __shared__ __align__(1024) unsigned char stage_mem[];
  __shared__ __align__(8) unsigned long long full_bar[kMaxStages];
  __shared__ __align__(8) unsigned long long empty_bar[kMaxStages];
  __shared__ unsigned long long meta_dst[kMaxStages];
  __shared__ unsigned int meta_bytes[kMaxStages];
  const TabGlobal tab{};
  const unsigned tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
  const unsigned S = p.stages;
  grid_dep_launch_dependents();
  if (tid == 0) {
    for (unsigned s = 0; s < S; ++s) {
      mbar_init(&full_bar[s], 1u);
      mbar_init(&empty_bar[s], 1u);
    }
    mbar_fence_init();
  }
  __syncthreads();
  const unsigned bid = blockIdx.x, grid = gridDim.x;
  const unsigned my_items = bid < p.total_items ? (p.total_items - bid + grid - 1u) / grid : 0u;
  const unsigned heads_per_piece = p.piece_bytes / p.tm_head_bytes;
  if (warp == 0) {
    grid_dep_wait();
    RawItem raw{};
    if (lane < my_items) item_load(p, tab, bid + lane * grid, raw);
    unsigned s = 0, phase = 0;
    for (unsigned base = 0; base < my_items; base += 32u) {
      unsigned long long sa = 0, da = 0;
      unsigned bytes = 0;
      if (base + lane < my_items) item_finish(p, raw, sa, da, bytes);
      const int my_blk = raw.s.id;
      const unsigned my_reg = raw.r, my_piece = raw.piece;
      if (base + 32u + lane < my_items) item_load(p, tab, bid + (base + 32u + lane) * grid, raw);
      const unsigned cnt = min(32u, my_items - base);
      for (unsigned j = 0; j < cnt; ++j) {
        const unsigned long long dst = __shfl_sync(0xffffffffu, da, j);
        const unsigned nbytes = __shfl_sync(0xffffffffu, bytes, j);
        const int blk = __shfl_sync(0xffffffffu, my_blk, j);
        const unsigned reg = __shfl_sync(0xffffffffu, my_reg, j);
        const unsigned piece = __shfl_sync(0xffffffffu, my_piece, j);
        mbar_wait(&empty_bar[s], phase ^ 1u);
        unsigned char* stage = stage_mem + static_cast<size_t>(s) * p.stage_bytes;
        if (lane == 0) {
          meta_dst[s] = dst;
          meta_bytes[s] = nbytes;
          mbar_arrive_expect_tx(&full_bar[s], nbytes);
        }
        __syncwarp();
        if (nbytes != 0) {
          const unsigned char* tm = static_cast<const unsigned char*>(p.src_tmaps) + static_cast<size_t>(reg) * 128u;
          const unsigned head0 = p.tm_head0 + piece * heads_per_piece;
          for (unsigned h = lane; h < heads_per_piece; h += 32u)
            tensor_g2s_4d(stage + static_cast<size_t>(h) * p.tm_head_bytes, tm, 0, static_cast<int>(head0 + h), 0, blk, &full_bar[s]);
        }
        if (++s == S) { s = 0; phase ^= 1u; }
      }
    }
  } else {
    unsigned prev_stage = 0;
    bool have_prev = false;
    grid_dep_wait();
    unsigned s = 0, phase = 0;
    for (unsigned it = 0; it < my_items; ++it) {
      mbar_wait(&full_bar[s], phase);
      const unsigned long long dst = meta_dst[s];
      const unsigned nbytes = meta_bytes[s];
      const unsigned char* stage = stage_mem + static_cast<size_t>(s) * p.stage_bytes;
      if (nbytes != 0) {
        if (p.d_runs == 1u) {
          if (lane == 0) bulk_s2g(dst, stage, nbytes);
        } else {
          for (unsigned x = lane; x < p.d_runs; x += 32u)
            bulk_s2g(dst + dst_run_offset(p, x), stage + x * p.d_run_bytes, p.d_run_bytes);
        }
      }
      bulk_commit();
      if (have_prev) {
        bulk_wait_read<1>();
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty_bar[prev_stage]);
      }
      prev_stage = s;
      have_prev = true;
      if (++s == S) { s = 0; phase ^= 1u; }
    }
    bulk_wait<0>();
  }
  finish_and_flag(p, tab);
}

// ----------------------------------------------------------------------------------------------
// LDG engine: no shared memory; each thread owns KMAX vector units of every piece.
// ----------------------------------------------------------------------------------------------
template <int VEC, int KMAX>
__global__ void __launch_bounds__(256) kvx_xfer_ldg_kernel(const __grid_constant__ XferParams p) {
  grid_dep_launch_dependents();
  const unsigned tid = threadIdx.x;
  const unsigned lane = tid & 31u;
  const unsigned T = blockDim.x;
  unsigned soff[KMAX], doff[KMAX];
  {
    const unsigned supr = p.s_run_bytes / VEC;
    const unsigned dupr = p.d_run_bytes / VEC;
#pragma unroll
    for (int k = 0; k < KMAX; ++k) {
      const unsigned u = tid + k * T;
      soff[k] = (u / supr) * p.s_stride + (u % supr) * VEC;
      doff[k] = dst_run_offset(p, u / dupr) + (u % dupr) * VEC;
    }
  }
  const unsigned bid = blockIdx.x;
  const unsigned grid = gridDim.x;
  const unsigned my_items = bid < p.total_items ? (p.total_items - bid + grid - 1u) / grid : 0u;
  grid_dep_wait();
  const TabGlobal tab{};
  RawItem raw{};
  if (lane < my_items) item_load(p, tab, bid + lane * grid, raw);
  for (unsigned base = 0; base < my_items; base += 32u) {
    unsigned long long sa = 0, da = 0;
    unsigned bytes = 0;
    if (base + lane < my_items) item_finish(p, raw, sa, da, bytes);
    if (base + 32u + lane < my_items) item_load(p, tab, bid + (base + 32u + lane) * grid, raw);
    const unsigned cnt = min(32u, my_items - base);
    for (unsigned j = 0; j < cnt; ++j) {
      const unsigned long long src = __shfl_sync(0xffffffffu, sa, j);
      const unsigned long long dst = __shfl_sync(0xffffffffu, da, j);
      const unsigned units = __shfl_sync(0xffffffffu, bytes, j) / VEC;
      Vec<VEC> v[KMAX];
#pragma unroll
      for (int k = 0; k < KMAX; ++k)
        if (tid + k * T < units) ldg_stream(v[k], src + soff[k]);
#pragma unroll
      for (int k = 0; k < KMAX; ++k)
        if (tid + k * T < units) stg_stream(dst + doff[k], v[k]);
    }
  }
  finish_and_flag(p, tab);
}

// Stream-ordered, bounded wait on a completion flag written by another GPU.
__global__ void kvx_flag_wait_kernel(const unsigned long long* flag, unsigned long long value,
                                     unsigned long long timeout_ns, unsigned int* status) {
  unsigned long long t0, now, cur;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
  for (;;) {
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(cur) : "l"(flag) : "memory");
    if (cur >= value) return;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
    if (now - t0 > timeout_ns) {
      atomicOr(status, kStatusWaitTimeout);
      return;
    }
    __nanosleep(256);
  }
}

}  // namespace kvx
