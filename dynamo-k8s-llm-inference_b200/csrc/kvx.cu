// kvx.cu — host side of libkvx.so: the C ABI declared in include/kvx.h over the sm_100a kernels in
// kvx_kernels.cuh.  No torch types, no CPU fallback: every compute entry needs a CUDA device.
#include "kvx.h"

#include <cuda.h>
#include <cuda_runtime.h>
#include <unistd.h>

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <new>
#include <string>
#include <vector>

#include "kvx_kernels.cuh"

using kvx::XferParams;
using kvx::XferSide;

namespace {

constexpr uint32_t kNumFlags = 4096;     // completion flags per context (uint64 each), all visible to peers
constexpr uint32_t kPrivFlags = 64;      // private words behind them (never exported): kvx_handoff_host's pageable h_done
// Finished-CTA counters.  A counter is only touched while its kernel RUNS (one atomicAdd per CTA at the end, reset
// by the last CTA), and kernels issued to one stream run one after the other — a dependent launch waits for its
// predecessor's completion before its first global access — so two counters per stream, used alternately, can
// never be shared by two running kernels however deep the backlog.  Launches recorded into a CUDA graph may be
// replayed on any stream later, so each gets a counter of its own for the life of the context.
constexpr uint32_t kStreamSlots = 448;   // distinct streams with flagged launches per context (2 counters each)
constexpr uint32_t kGraphCounters = 128; // flagged launches recorded into CUDA graphs per context
constexpr uint32_t kNumCounters = 2 * kStreamSlots + kGraphCounters;
constexpr uint32_t kTableRing = 1u << 22;  // int32 entries of device scratch for host block tables
constexpr uint64_t kNvlinkInFlightBytes = 8ull << 20;  // 6 MiB in flight saturate one NVLink direction (tools/footprint.py); margin for a busy switch
constexpr uint32_t kMemMagic = 0x4d58564bu;   // "KVXM"
constexpr uint32_t kPoolMagic = 0x5058564bu;  // "KVXP"

struct MemRecord {  // wire format of one exported device buffer
  uint32_t magic;
  int32_t device;
  int64_t pid;
  uint64_t raw_ptr;
  uint64_t offset;  // raw_ptr - allocation base
  uint64_t bytes;
  cudaIpcMemHandle_t handle;
  unsigned char device_uuid[16];  // the GPU the bytes live on: ordinals differ between processes, UUIDs do not
  uint32_t flags;                 // kMemHasIpcHandle: `handle` can be opened by another process
  uint32_t reserved;
};
constexpr uint32_t kMemHasIpcHandle = 1u;

struct PoolRecordHeader {  // wire format of an exported pool: header + regions x MemRecord
  uint32_t magic;
  uint32_t regions;
  uint32_t block_tokens, heads, head_dim, elem_bytes, layout, tp_rank, tp_size, head_stride;
  uint64_t num_blocks, block_stride;
};

// Peer allocations opened in this process, by handle bytes: a handle may be opened once.
struct IpcMapping {
  void* base;
  int refs;
};
std::mutex g_ipc_mu;
std::map<std::string, IpcMapping> g_ipc_open;
// A kernel's dynamic shared-memory limit is per process and device: every path that raises it (contexts on any
// host thread, kvx_handoff_static's set-up) reads and writes it under this one lock, so "raise, never lower" holds.
std::mutex g_func_attr_mu;

}  // namespace

struct kvx_ctx {
  int device = 0;
  int sms = 0;
  int64_t opt[KVX_OPT__COUNT] = {0};
  uint64_t launches = 0;
  uint64_t* d_flags = nullptr;
  unsigned* d_counters = nullptr;
  unsigned* d_status = nullptr;
  int* d_table_ring = nullptr;
  int* h_table_ring = nullptr;  // pinned mirror of d_table_ring, allocated on first use (long block-id lists only)
  uint32_t table_cursor = 0;
  struct RingUse {  // a span of the table ring and the event that marks its reader's completion
    uint32_t begin, end;
    cudaEvent_t done;
  };
  std::vector<RingUse> ring_uses;
  std::map<cudaStream_t, uint32_t> stream_slot;  // stream → its pair of counters
  std::vector<uint8_t> slot_phase;               // which counter of the pair the stream's next launch takes
  uint32_t graph_counters = 0;                   // counters handed to launches recorded into graphs
  uint32_t priv_cursor = 0;
  cudaStream_t aux = nullptr;
  bool force_sys_fence = false;  // KVX_FORCE_SYS_FENCE=1: system-scope fences everywhere (debugging aid)
  std::string last_error;
  unsigned char uuid[16] = {0};  // this device's UUID
  struct Imported {
    std::string key;  // IPC handle bytes ("" = same process)
    uint64_t bytes;
    bool local;       // the memory lives on this context's GPU
  };
  std::map<void*, Imported> imported;  // pointers handed out by import_record
  std::map<const void*, size_t> smem_limit;  // kernel → dynamic shared memory limit already set
  std::vector<struct kvx_queue*> queues;      // resident receivers started on this context (stopped with it)
  std::map<void*, uint64_t> vmm_allocs;       // kvx_mem_alloc_vmm ranges
};

struct kvx_pool {
  kvx_ctx* ctx = nullptr;
  kvx_pool_desc desc{};
  std::vector<void*> base;
  std::vector<void*> imported_ptrs;  // to unmap on release
  unsigned long long* d_bases = nullptr;
  mutable void* d_tmaps = nullptr;  // tensor maps of an NHD pool's regions (built on first use by the tensor-map engine)
  struct VmmRange {
    CUdeviceptr va;
    uint64_t size;
  };
  std::vector<VmmRange> vmm_ranges;  // address ranges mapped by kvx_pool_import_fds (unmapped on release)
  uint64_t chunk_bytes = 0;
  bool local = true;  // the pool's memory lives on the context's own GPU (whoever allocated or exported it)
};

namespace {

void pool_unmap_vmm(kvx_pool* pool);  // kvx_vmm_host.inl

int32_t cuda_fail(kvx_ctx* ctx, cudaError_t e, const char* what) {
  if (ctx) {
    ctx->last_error = std::string(what) + ": " + cudaGetErrorName(e) + " (" + cudaGetErrorString(e) + ")";
  }
  (void)cudaGetLastError();
  return KVX_ECUDA;
}
#define KVX_CUDA(ctx, call)                                    \
  do {                                                         \
    cudaError_t e__ = (call);                                  \
    if (e__ != cudaSuccess) return cuda_fail(ctx, e__, #call); \
  } while (0)

// Nothing may leave the C ABI by exception: the containers used on the host side can throw bad_alloc.
#define KVX_CATCH_ALL                               \
  catch (const std::bad_alloc&) { return KVX_ENOMEM; } \
  catch (...) { return KVX_EINVAL; }

struct DeviceGuard {
  int prev = -1;
  bool ok = false;
  explicit DeviceGuard(int dev) {
    if (cudaGetDevice(&prev) == cudaSuccess && cudaSetDevice(dev) == cudaSuccess) ok = true;
  }
  ~DeviceGuard() {
    if (ok && prev >= 0) cudaSetDevice(prev);
  }
};

int64_t env_or(const char* name, int64_t dflt) {
  const char* v = getenv(name);
  if (!v || !*v) return dflt;
  return strtoll(v, nullptr, 0);
}

// ---- geometry ---------------------------------------------------------------------------------
struct ChunkGeom {  // one side's chunk: H heads in `layout`, slice starts at head `off`, nb tokens per block
  uint32_t layout, heads, off, nb;
  uint32_t hs = 0;  // HND only: bytes between consecutive heads of a chunk; 0 = nb * row (heads back to back)
};

struct Plan {
  uint32_t ppc, piece_bytes;
  uint32_t s_runs, s_run_bytes, s_stride, s_slice_off, s_piece_stride, s_sub_stride;
  uint32_t d_runs, d_run_bytes, d_n1, d_stride_a, d_stride_b, d_slice_off, d_piece_stride, d_sub_stride;
};

// Split a slice of Hc heads x N tokens x R bytes (one GRANULE: N is the smaller of the two sides'
// block sizes) into pieces of at most `cap` bytes and describe each piece as source runs /
// destination runs (see XferParams).  Pieces split along the source's major dimension (tokens for
// NHD, heads for HND) so that both sides stay a 2-level run pattern.  A side whose blocks hold
// nb = k*N tokens keeps k granules per block: NHD granules are contiguous token ranges, HND
// granules are N-row runs inside each head's nb rows (head stride nb*R).
int32_t make_plan(uint32_t N, uint32_t R, uint32_t Hc, ChunkGeom s, ChunkGeom d, uint32_t cap, Plan* out) {
  if (Hc == 0 || N == 0 || R == 0 || (R % 16u) != 0) return KVX_EUNSUPPORTED;
  if (s.off + Hc > s.heads || d.off + Hc > d.heads) return KVX_EINVAL;
  if (s.nb < N || d.nb < N || s.nb % N || d.nb % N) return KVX_EINVAL;
  const bool src_hnd = s.layout == KVX_LAYOUT_HND;
  const bool dst_hnd = d.layout == KVX_LAYOUT_HND;
  // distance between heads of an HND chunk: nb rows when heads lie back to back, more in a cross-layer
  // pool whose heads are the outermost dimension of a block ((nb, H, L, 2, N, D), [vllm] flash_attn.py:162-165)
  const uint32_t shs = s.hs ? s.hs : s.nb * R, dhs = d.hs ? d.hs : d.nb * R;
  if (src_hnd && (shs < s.nb * R || shs % 16u)) return KVX_EINVAL;
  if (dst_hnd && (dhs < d.nb * R || dhs % 16u)) return KVX_EINVAL;
  const uint32_t major = src_hnd ? Hc : N;
  const uint32_t unit_bytes = src_hnd ? N * R : Hc * R;
  if (unit_bytes > cap) return KVX_EUNSUPPORTED;
  uint32_t g = 1;
  for (uint32_t c = major; c >= 1; --c) {
    if (major % c == 0 && static_cast<uint64_t>(c) * unit_bytes <= cap) {
      g = c;
      break;
    }
  }
  Plan p{};
  p.ppc = major / g;
  p.piece_bytes = g * unit_bytes;
  p.s_sub_stride = src_hnd ? N * R : N * s.heads * R;
  p.d_sub_stride = dst_hnd ? N * R : N * d.heads * R;
  if (src_hnd) {
    // image order (h, t); head h of the granule is N*R contiguous bytes, heads are shs apart
    if (shs == N * R) { p.s_runs = 1; p.s_run_bytes = p.piece_bytes; p.s_stride = 0; }
    else              { p.s_runs = g; p.s_run_bytes = N * R; p.s_stride = shs; }
    p.s_slice_off = s.off * shs; p.s_piece_stride = g * shs;
    if (dst_hnd) {
      if (dhs == N * R) { p.d_runs = 1; p.d_run_bytes = p.piece_bytes; p.d_n1 = 1; p.d_stride_a = 0; }
      else              { p.d_runs = g; p.d_run_bytes = N * R; p.d_n1 = 1; p.d_stride_a = dhs; }
      p.d_stride_b = 0;
      p.d_slice_off = d.off * dhs; p.d_piece_stride = g * dhs;
    } else {
      // image order (h, t) → NHD row (t, off+h): run x = h*N + t
      p.d_runs = g * N; p.d_run_bytes = R; p.d_n1 = N; p.d_stride_a = R; p.d_stride_b = d.heads * R;
      p.d_slice_off = d.off * R; p.d_piece_stride = g * R;
    }
  } else {
    // image order (t, h); token t of the slice is Hc*R contiguous bytes, tokens are s.heads*R apart
    if (Hc == s.heads) { p.s_runs = 1; p.s_run_bytes = p.piece_bytes; p.s_stride = 0; }
    else               { p.s_runs = g; p.s_run_bytes = Hc * R; p.s_stride = s.heads * R; }
    p.s_slice_off = s.off * R; p.s_piece_stride = g * s.heads * R;
    if (!dst_hnd) {
      if (Hc == d.heads) { p.d_runs = 1; p.d_run_bytes = p.piece_bytes; p.d_n1 = 1; p.d_stride_a = 0; }
      else               { p.d_runs = g; p.d_run_bytes = Hc * R; p.d_n1 = 1; p.d_stride_a = d.heads * R; }
      p.d_stride_b = 0;
      p.d_slice_off = d.off * R; p.d_piece_stride = g * d.heads * R;
    } else {
      // image order (t, h) → HND row (off+h, t): run x = t*Hc + h
      p.d_runs = g * Hc; p.d_run_bytes = R; p.d_n1 = Hc; p.d_stride_a = R; p.d_stride_b = dhs;
      p.d_slice_off = d.off * dhs; p.d_piece_stride = g * R;
    }
  }
  *out = p;
  return KVX_OK;
}

enum Engine { kEngineAuto = 0, kEngineLdg = 1, kEngineTma = 2 };

struct LaunchSpec {  // what a transfer needs besides geometry
  XferSide src{}, dst{};
  uint32_t n = 0, regions = 0;
  uint32_t N = 0, R = 0, Hc = 0;
  ChunkGeom sg{}, dg{};
  uint64_t* flag = nullptr;
  uint64_t flag_val = 0;
  bool copy_mode = false;  // contiguous→contiguous bytes (kvx_copy)
  bool dst_is_peer = true;  // unknown destinations are treated as remote (fence scope)
  bool dst_is_peer_known = false;  // the destination is KNOWN to be a peer's memory (launch geometry)
  bool src_is_peer = false;        // the source is known to be a peer's memory
  uint64_t copy_bytes = 0;
  // kvx_handoff_host: block ids still in host memory (ns source ids, nd destination ids) that may
  // ride in the kernel's parameter block, and the caller's pinned completion word (device alias).
  const int32_t* h_src_tab = nullptr;
  const int32_t* h_dst_tab = nullptr;
  uint32_t ns = 0, nd = 0;
  uint64_t* flag2 = nullptr;
  // kvx_handoff_batch: (address, value) pairs the last CTA stores, carried next to the ids
  const uint64_t* notify = nullptr;
  uint32_t notify_count = 0;
  bool block_major = false;  // visit items block by block (cross-layer pools)
  const kvx_pool* dst_pool = nullptr;  // paged destination as a pool (tensor-map engine)
  const kvx_pool* src_pool = nullptr;  // paged source as a pool (tensor-map engine, load side)
};

// int32 entries of the parameter table area a launch needs: both sides' ids, then the notifications
uint32_t param_entries(const LaunchSpec& ls) {
  return ls.notify_count ? ((ls.ns + ls.nd + 1u) & ~1u) + 4u * ls.notify_count : ls.ns + ls.nd;
}

// launch_xfer's answer when the host tables cannot ride in the parameters (engine / geometry / size):
// the caller uploads them and launches with device tables instead.
constexpr int32_t kNoParamTables = 1;
// Four size classes of the parameter block (ids of both sides together).  Measured on B200
// (tools/e2e_host.py): with the ids read by generic loads the host-table entry costs the same as the
// device-table entry at every size up to the 32 KiB parameter limit, 5-6 us less per request than
// uploading the lists first.  (Indexed constant-bank reads were tried first: the 32 lanes of a lookup
// want 32 different entries and the constant cache serves one address at a time, +40 us at 4096 ids.)
constexpr uint32_t kParamTabSmall = 128, kParamTabMid = 512, kParamTabBig = 2048, kParamTabMax = 8000;
static_assert(sizeof(kvx::XferParamsTab<kParamTabMax>) <= 32764, "kernel parameter block limit (CUDA 12.1+, sm_70+)");

bool all_aligned(uint64_t a, const Plan& pl, const XferSide& s, const XferSide& d) {
  auto m = [a](uint64_t v) { return (v % a) == 0; };
  return m(pl.s_run_bytes) && m(pl.d_run_bytes) && m(pl.s_stride) && m(pl.d_stride_a) && m(pl.d_stride_b) &&
         m(pl.s_slice_off) && m(pl.d_slice_off) && m(pl.s_piece_stride) && m(pl.d_piece_stride) &&
         m(pl.s_sub_stride) && m(pl.d_sub_stride) &&
         m(s.block_stride) && m(d.block_stride) && m(s.base0) && m(d.base0);
}

// Raise a kernel's dynamic shared-memory limit once per (context, kernel): later launches — including
// ones issued while a stream is being captured into a CUDA graph — make no attribute call.
template <typename K>
int32_t set_smem(kvx_ctx* ctx, K kernel, size_t bytes) {
  size_t& have = ctx->smem_limit[reinterpret_cast<const void*>(kernel)];
  if (have >= bytes) return KVX_OK;
  std::lock_guard<std::mutex> lk(g_func_attr_mu);
  cudaFuncAttributes fa{};
  KVX_CUDA(ctx, cudaFuncGetAttributes(&fa, kernel));  // another context or kvx_handoff_static may have raised it already
  if (static_cast<size_t>(fa.maxDynamicSharedSizeBytes) >= bytes) {
    have = static_cast<size_t>(fa.maxDynamicSharedSizeBytes);
    return KVX_OK;
  }
  KVX_CUDA(ctx, cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(bytes)));
  have = bytes;
  return KVX_OK;
}

// One launch.  With `pdl` the kernel is allowed to become resident while the previous kernel on the
// stream is still draining (programmatic dependent launch); the kernels wait for that kernel's
// completion themselves before they touch global memory (grid_dep_wait in kvx_kernels.cuh).
template <typename K, typename P>
cudaError_t launch_kernel(K kernel, uint32_t grid, uint32_t block, size_t smem, cudaStream_t stream, bool pdl,
                          const P& p) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(block);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl ? 1u : 0u;
  return cudaLaunchKernelEx(&cfg, kernel, p);
}

// The finished-CTA counter of one flagged launch (see kStreamSlots above).
int32_t take_counter(kvx_ctx* ctx, cudaStream_t stream, unsigned** out) {
  cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
  if (cudaStreamIsCapturing(stream, &cs) != cudaSuccess) {
    (void)cudaGetLastError();
    cs = cudaStreamCaptureStatusNone;
  }
  if (cs == cudaStreamCaptureStatusActive) {
    if (ctx->graph_counters == kGraphCounters) {
      ctx->last_error = "more than 128 flagged transfers recorded into CUDA graphs by one context";
      return KVX_ERANGE;
    }
    *out = ctx->d_counters + 2 * kStreamSlots + ctx->graph_counters++;
    return KVX_OK;
  }
  auto it = ctx->stream_slot.find(stream);
  if (it == ctx->stream_slot.end()) {
    if (ctx->stream_slot.size() == kStreamSlots) {
      // Every pair is taken by a stream seen earlier (an engine that keeps creating streams): once the device
      // is idle no counter is in use and all of them can be handed out again.
      KVX_CUDA(ctx, cudaDeviceSynchronize());
      ctx->stream_slot.clear();
    }
    if (ctx->slot_phase.size() != kStreamSlots) ctx->slot_phase.assign(kStreamSlots, 0);
    // after a clear() the pairs are reassigned from 0 upwards
    it = ctx->stream_slot.emplace(stream, static_cast<uint32_t>(ctx->stream_slot.size())).first;
  }
  const uint32_t slot = it->second;
  *out = ctx->d_counters + 2 * slot + (ctx->slot_phase[slot] ^= 1u);
  return KVX_OK;
}

// ---- launch geometry ----------------------------------------------------------------------------
// Every KVX_OPT_* geometry value of 0 means "choose here", from what the sweeps under tools/ and
// profiles/ measured on B200:
//   on-GPU transfers are HBM-bound and want every SM: 4 x 16 KiB stages, 16 consumer warps, ONE CTA per
//     SM — with dependent launch the next handoff's CTA then fits beside the draining one (2 x 64 KiB,
//     2 x 544 threads) and its set-up and pipeline fill overlap the tail completely
//     (tools/small_requests.py: 15.5 -> 11.6 us at 8 blocks, 47.3 -> 44.1 at 64, 296.6 -> 294.1 at 512);
//   pieces of 8 KiB or less (chunks of TP shards: 8 KiB at 2 heads, 4 KiB at 1): an item then costs the
//     hand-shake between the producer and every consumer warp, not its bytes, so few fat consumer warps
//     and more CTAs win (tools/small_pieces.py: 8 KiB chunks 4.9 -> 5.9 TB/s, 4 KiB chunks 2.7 -> 6.0 TB/s);
//   transfers with one side in a peer's memory are NVLink-bound: ~6 MiB in flight saturate the link
//     (tools/footprint.py), so they run deep rings (6 x 32 KiB, 8 x 8 KiB for small pieces) and, when the
//     PEER is the source (pull), in just enough CTAs for kNvlinkInFlightBytes, which leaves the other
//     SMs to the work running beside them;
//   the register engine has no shared-memory ring to keep bytes in flight and keeps two CTAs per SM.
struct Geometry {
  uint32_t stages, stage_bytes, ctas_per_sm, cons_warps, max_ctas;
};

uint32_t plan_stage_bytes(const kvx_ctx* ctx, bool remote) {
  const uint32_t opt = static_cast<uint32_t>(ctx->opt[KVX_OPT_STAGE_BYTES]);
  return opt ? opt : (remote ? 32768u : 16384u);
}

Geometry resolve_geometry(const kvx_ctx* ctx, bool remote, bool pull, int engine, bool copy_mode, uint32_t piece_bytes) {
  Geometry g{};
  g.stages = static_cast<uint32_t>(ctx->opt[KVX_OPT_STAGES]);
  g.stage_bytes = static_cast<uint32_t>(ctx->opt[KVX_OPT_STAGE_BYTES]);
  g.ctas_per_sm = static_cast<uint32_t>(ctx->opt[KVX_OPT_CTAS_PER_SM]);
  g.cons_warps = static_cast<uint32_t>(ctx->opt[KVX_OPT_CONSUMER_WARPS]);
  const bool small = engine == kEngineTma && !copy_mode && piece_bytes <= 8192u && g.stages == 0 && g.stage_bytes == 0;
  if (g.stage_bytes == 0) g.stage_bytes = small ? 8192u : plan_stage_bytes(ctx, remote);
  if (g.stages == 0) g.stages = small ? (remote ? 8u : 4u) : (remote ? 6u : 4u);
  if (g.ctas_per_sm == 0) g.ctas_per_sm = small ? 3u : (engine == kEngineLdg ? 2u : 1u);
  if (g.cons_warps == 0) g.cons_warps = small ? 4u : 16u;
  g.max_ctas = static_cast<uint32_t>(ctx->sms) * g.ctas_per_sm;
  if (ctx->opt[KVX_OPT_MAX_CTAS] > 0) {  // the caller bounds the SMs a transfer occupies
    g.max_ctas = std::min<uint32_t>(g.max_ctas, static_cast<uint32_t>(ctx->opt[KVX_OPT_MAX_CTAS]));
  } else if (pull && engine == kEngineTma) {
    // loads from the peer are bounded by the bytes the rings keep in flight; pushed stores are posted and
    // gain ~1.3 % from every SM (699 -> 708 GB/s), so only the pull direction runs lean
    const uint64_t in_flight_per_cta = static_cast<uint64_t>(g.stages) * piece_bytes;
    const uint32_t need = static_cast<uint32_t>((kNvlinkInFlightBytes + in_flight_per_cta - 1) / in_flight_per_cta);
    g.max_ctas = std::min<uint32_t>(g.max_ctas, std::max<uint32_t>(16u, need));
  }
  return g;
}

// ---- tensor maps ----------------------------------------------------------------------------------
// One CUtensorMap per region of an NHD pool: the region as a 4-d tensor (row elements, heads, tokens of a block,
// blocks) with the pool's own strides; the box is one head's rows of one block — (row, 1, N, 1) — so that a
// shared-memory slab [N][row] (a head of an HND chunk) lands as N rows, heads*row apart.
typedef CUresult (*PFN_tensorMapEncodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                             const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                             CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int32_t pool_tensor_maps(kvx_ctx* ctx, const kvx_pool* pool, const void** out) {
  if (pool->d_tmaps) {
    *out = pool->d_tmaps;
    return KVX_OK;
  }
  static PFN_tensorMapEncodeTiled encode = nullptr;
  if (!encode) {
    void* f = nullptr;
    cudaDriverEntryPointQueryResult q;
    KVX_CUDA(ctx, cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q));
    if (!f || q != cudaDriverEntryPointSuccess) return KVX_EUNSUPPORTED;
    encode = reinterpret_cast<PFN_tensorMapEncodeTiled>(f);
  }
  const kvx_pool_desc& d = pool->desc;
  if (d.layout != KVX_LAYOUT_NHD) return KVX_EUNSUPPORTED;
  const uint64_t row = static_cast<uint64_t>(d.head_dim) * d.elem_bytes;
  // element type: the widest integer that divides a row and keeps the row within 256 box elements
  uint32_t es = 0;
  CUtensorMapDataType dt = CU_TENSOR_MAP_DATA_TYPE_UINT8;
  if (row % 8 == 0 && row / 8 <= 256) { es = 8; dt = CU_TENSOR_MAP_DATA_TYPE_UINT64; }
  else if (row % 4 == 0 && row / 4 <= 256) { es = 4; dt = CU_TENSOR_MAP_DATA_TYPE_UINT32; }
  else if (row % 2 == 0 && row / 2 <= 256) { es = 2; dt = CU_TENSOR_MAP_DATA_TYPE_UINT16; }
  else if (row <= 256) { es = 1; }
  if (es == 0 || d.block_tokens > 256 || (row * d.block_tokens) % 128 != 0) return KVX_EUNSUPPORTED;
  std::vector<CUtensorMap> maps(pool->base.size());
  const cuuint64_t dims[4] = {row / es, d.heads, d.block_tokens, d.num_blocks};
  const cuuint64_t strides[3] = {row, row * d.heads, d.block_stride};  // bytes, dimensions 1..3
  const cuuint32_t box[4] = {static_cast<cuuint32_t>(row / es), 1u, d.block_tokens, 1u};
  const cuuint32_t estr[4] = {1u, 1u, 1u, 1u};
  for (size_t r = 0; r < pool->base.size(); ++r) {
    CUresult cr = encode(&maps[r], dt, 4, pool->base[r], dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                         CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (cr != CUDA_SUCCESS) {
      ctx->last_error = "cuTensorMapEncodeTiled failed with CUresult " + std::to_string(static_cast<int>(cr));
      return KVX_EUNSUPPORTED;
    }
  }
  static_assert(sizeof(CUtensorMap) == 128, "kernels index the map array in 128-byte steps");
  void* dev = nullptr;
  KVX_CUDA(ctx, cudaMalloc(&dev, maps.size() * sizeof(CUtensorMap)));
  cudaError_t e = cudaMemcpy(dev, maps.data(), maps.size() * sizeof(CUtensorMap), cudaMemcpyHostToDevice);
  if (e != cudaSuccess) {
    cudaFree(dev);
    return cuda_fail(ctx, e, "tensor map upload");
  }
  pool->d_tmaps = dev;
  *out = dev;
  return KVX_OK;
}

// The default kernel with the request's block ids copied into its parameter block: no H2D copy, no
// copy-engine hand-over before the launch.
template <int NT>
int32_t launch_ptab(kvx_ctx* ctx, const XferParams& p, const LaunchSpec& ls, uint32_t grid, uint32_t block, size_t smem,
                    cudaStream_t stream, bool pdl) {
  auto k = kvx::kvx_xfer_tma_ptab_kernel<16, 8, NT>;
  int32_t rc = set_smem(ctx, k, smem);
  if (rc) return rc;
  kvx::XferParamsTab<NT> pt;
  pt.p = p;
  pt.p.src.table = nullptr;
  pt.p.dst.table = nullptr;
  pt.p.src.tab_off = 0;
  pt.p.dst.tab_off = ls.ns;
  memcpy(pt.tab, ls.h_src_tab, ls.ns * sizeof(int));
  memcpy(pt.tab + ls.ns, ls.h_dst_tab, ls.nd * sizeof(int));
  if (ls.notify_count != 0) {
    const uint32_t off = (ls.ns + ls.nd + 1u) & ~1u;  // 8-byte aligned inside the int32 array
    pt.p.notify_off = off;
    pt.p.notify_count = ls.notify_count;
    memcpy(pt.tab + off, ls.notify, static_cast<size_t>(ls.notify_count) * 2 * sizeof(uint64_t));
  }
  KVX_CUDA(ctx, launch_kernel(k, grid, block, smem, stream, pdl, pt));
  return KVX_OK;
}

int32_t launch_xfer(kvx_ctx* ctx, const LaunchSpec& ls, const std::vector<void*>* src_bases,
                    const std::vector<void*>* dst_bases, cudaStream_t stream) {
  DeviceGuard guard(ctx->device);
  if (!guard.ok) return cuda_fail(ctx, cudaGetLastError(), "cudaSetDevice");

  int engine = static_cast<int>(ctx->opt[KVX_OPT_ENGINE]);
  if (engine == kEngineAuto) engine = kEngineTma;
  const bool remote = ls.src_is_peer || ls.dst_is_peer_known;  // one side in a peer's memory: NVLink-bound
  const uint32_t stage_cap = plan_stage_bytes(ctx, remote);     // bytes one piece may fill
  const int store_mode = static_cast<int>(ctx->opt[KVX_OPT_STORE_MODE]);
  int vec_opt = static_cast<int>(ctx->opt[KVX_OPT_VEC_BYTES]);
  const bool pdl = ctx->opt[KVX_OPT_PDL] != 0;

  constexpr uint32_t kLdgThreads = 256;
  // KMAX by vector width: 8 units of 16 B or 4 units of 32 B per thread.
  const uint32_t ldg_cap16 = kLdgThreads * 8u * 16u, ldg_cap32 = kLdgThreads * 4u * 32u;

  XferParams p{};
  p.src = ls.src;
  p.dst = ls.dst;
  p.n = ls.n;
  p.regions = ls.regions;
  Plan pl{};
  const void* tmap_load_maps = nullptr;
  bool tmap_load_auto = false;
  if (ls.copy_mode) {
    uint32_t piece = engine == kEngineTma ? stage_cap : ldg_cap16;
    if (engine == kEngineTma && store_mode == 1 && ctx->opt[KVX_OPT_CONSUMER_WARPS] > 0)
      piece = std::min<uint32_t>(piece, static_cast<uint32_t>(ctx->opt[KVX_OPT_CONSUMER_WARPS]) * 32u * 8u * 16u);
    const uint64_t items = (ls.copy_bytes + piece - 1) / piece;
    if (items == 0) return KVX_OK;
    if (items >= (1ull << 31)) return KVX_EUNSUPPORTED;
    pl.ppc = 1; pl.piece_bytes = piece;
    pl.s_runs = 1; pl.s_run_bytes = piece; pl.d_runs = 1; pl.d_run_bytes = piece; pl.d_n1 = 1;
    p.n = static_cast<uint32_t>(items);
    p.regions = 1;
    p.src.block_stride = piece;
    p.dst.block_stride = piece;
    const uint64_t tail = ls.copy_bytes - (items - 1) * piece;
    p.tail_bytes = tail == piece ? 0u : static_cast<uint32_t>(tail);
  } else {
    // store_mode 3 with an NHD source and an HND destination: the permutation moves to the LOAD side (one tensor-map load
    // per head lands the piece head-major in shared memory), so the piece is planned as if the source were HND
    ChunkGeom sg = ls.sg;
    // (store_mode 0 takes this path by itself for the case it measured faster in: whole 8 KiB chunks of two or more heads on
    // one GPU, 6.35 against 5.95 TB/s — the mirror image of the store-side rule further down)
    const bool load_auto = store_mode == 0 && !remote && static_cast<uint64_t>(ls.Hc) * ls.N * ls.R == 8192u && ls.N * ls.R <= 4096u &&
                           ctx->opt[KVX_OPT_STAGES] == 0 && ctx->opt[KVX_OPT_STAGE_BYTES] == 0 && ctx->opt[KVX_OPT_CONSUMER_WARPS] == 0 &&
                           ctx->opt[KVX_OPT_VEC_BYTES] == 0;
    tmap_load_auto = load_auto;
    if (engine == kEngineTma && (store_mode == 3 || load_auto) && ls.src_pool != nullptr && ls.h_src_tab == nullptr && ls.notify_count == 0 &&
        ls.sg.layout == KVX_LAYOUT_NHD && ls.dg.layout == KVX_LAYOUT_HND && ls.sg.nb == ls.N && ls.dg.nb == ls.N &&
        pool_tensor_maps(ctx, ls.src_pool, &tmap_load_maps) == KVX_OK) {
      sg.layout = KVX_LAYOUT_HND;
      sg.hs = 0;
    } else {
      tmap_load_maps = nullptr;
    }
    int32_t rc = make_plan(ls.N, ls.R, ls.Hc, sg, ls.dg, engine == kEngineTma ? stage_cap : ldg_cap16, &pl);
    if (rc == KVX_EUNSUPPORTED && engine == kEngineTma) {  // unit larger than a stage: register path
      engine = kEngineLdg;
      tmap_load_maps = nullptr;
      rc = make_plan(ls.N, ls.R, ls.Hc, ls.sg, ls.dg, ldg_cap16, &pl);
    }
    if (rc != KVX_OK) return rc;
    p.src.slice_off = pl.s_slice_off; p.src.piece_stride = pl.s_piece_stride;
    p.dst.slice_off = pl.d_slice_off; p.dst.piece_stride = pl.d_piece_stride;
    p.src.ratio = ls.sg.nb / ls.N; p.src.sub_stride = pl.s_sub_stride;
    p.dst.ratio = ls.dg.nb / ls.N; p.dst.sub_stride = pl.d_sub_stride;
  }
  if (p.src.ratio == 0) p.src.ratio = 1;
  if (p.dst.ratio == 0) p.dst.ratio = 1;
  const uint64_t total = static_cast<uint64_t>(p.n) * p.regions * pl.ppc;
  if (total == 0) {
    if (ls.flag) {  // empty request (full prefix hit): notify only
      KVX_CUDA(ctx, cudaMemcpyAsync(ls.flag, &ls.flag_val, sizeof(uint64_t), cudaMemcpyHostToDevice, stream));
    }
    if (ls.flag2) KVX_CUDA(ctx, cudaMemcpyAsync(ls.flag2, &ls.flag_val, sizeof(uint64_t), cudaMemcpyDefault, stream));
    return KVX_OK;
  }
  if (total >= (1ull << 31)) return KVX_EUNSUPPORTED;
  p.total_items = static_cast<uint32_t>(total);
  p.ppc = pl.ppc; p.piece_bytes = pl.piece_bytes;
  p.s_runs = pl.s_runs; p.s_run_bytes = pl.s_run_bytes; p.s_stride = pl.s_stride;
  p.d_runs = pl.d_runs; p.d_run_bytes = pl.d_run_bytes; p.d_n1 = pl.d_n1;
  p.d_stride_a = pl.d_stride_a; p.d_stride_b = pl.d_stride_b;
  const Geometry g = resolve_geometry(ctx, remote, ls.src_is_peer, engine, ls.copy_mode, pl.piece_bytes);
  const uint32_t stages = g.stages, stage_bytes = g.stage_bytes, cons_warps = g.cons_warps;
  p.stages = stages; p.stage_bytes = stage_bytes;
  p.block_major = (ls.block_major && !ls.copy_mode && env_or("KVX_BLOCK_MAJOR", 1) != 0) ? 1u : 0u;
  p.l2_hint = static_cast<uint32_t>(ctx->opt[KVX_OPT_L2_HINT]);
  p.flag = reinterpret_cast<unsigned long long*>(ls.flag);
  // Fence scopes (finish_and_flag): per CTA device-wide when the data stays on this GPU, and to the
  // end device-wide when the flag is one of this context's own slots and no second word is written.
  p.flag2 = reinterpret_cast<unsigned long long*>(ls.flag2);
  p.data_local = (!ls.dst_is_peer && !ctx->force_sys_fence) ? 1u : 0u;
  p.flag_local = (p.data_local && ls.flag2 == nullptr && ls.notify_count == 0 &&
                  (ls.flag == nullptr || (ls.flag >= ctx->d_flags && ls.flag < ctx->d_flags + kNumFlags + kPrivFlags))) ? 1u : 0u;
  p.flag_val = ls.flag_val;
  p.status = ctx->d_status;
  if (ls.flag != nullptr || ls.flag2 != nullptr || ls.notify_count != 0) {  // only flagged launches count their CTAs
    int32_t rc = take_counter(ctx, stream, &p.done_counter);
    if (rc) return rc;
  }

  // 32-byte accesses need every address component 32-byte aligned (region bases included).
  bool ok32 = all_aligned(32, pl, p.src, p.dst);
  if (ok32 && src_bases) for (void* b : *src_bases) ok32 = ok32 && (reinterpret_cast<uintptr_t>(b) % 32 == 0);
  if (ok32 && dst_bases) for (void* b : *dst_bases) ok32 = ok32 && (reinterpret_cast<uintptr_t>(b) % 32 == 0);
  if (p.tail_bytes % 32 != 0) ok32 = false;
  if (vec_opt == 0) vec_opt = 16;  // measured: 16-byte stores beat 32-byte ones on B200 for this path
  if (vec_opt == 32 && !ok32) vec_opt = 16;

  const uint32_t grid = std::min<uint32_t>(g.max_ctas, p.total_items);

  if (ls.h_src_tab != nullptr) {
    // Host tables ride in the parameters only with the default kernel (TMA engine, 16-byte thread stores).
    const bool dflt_kernel = engine == kEngineTma && store_mode != 2 && vec_opt == 16 &&
                             static_cast<uint64_t>(cons_warps) * 32u * 8u * 16u >= pl.piece_bytes;
    if (!dflt_kernel || param_entries(ls) > std::min<int64_t>(kParamTabMax, ctx->opt[KVX_OPT_PARAM_TABLES])) return kNoParamTables;
  }

  if (engine == kEngineTma) {
    // store_mode 0 = thread stores: they match or beat bulk stores at these stage sizes (profiles/r01_sweep_*)
    const size_t smem = static_cast<size_t>(stages) * stage_bytes;
    // store_mode 3 = tensor-map stores, where the case allows it: HND chunks into an NHD paged pool of the same block
    // size (the fused handoff's transform and the TP slot); everything else falls through to thread stores
    // store_mode 0 picks them by itself where they measured faster: 8 KiB pieces of two or more heads on one GPU (chunks
    // of TP4 shards of an 8-head model: 6.30 against 5.95 TB/s, tools/tmap_variants.py), two CTAs per SM
    const bool tmap_shape = !ls.copy_mode && ls.dst_pool != nullptr && ls.h_src_tab == nullptr && ls.sg.layout == KVX_LAYOUT_HND &&
                            ls.dg.layout == KVX_LAYOUT_NHD && p.src.ratio == 1 && p.dst.ratio == 1 && ls.notify_count == 0;
    const bool tmap_auto = store_mode == 0 && tmap_shape && !remote && pl.piece_bytes == 8192u && ls.N * ls.R <= 4096u &&
                           ctx->opt[KVX_OPT_STAGES] == 0 && ctx->opt[KVX_OPT_STAGE_BYTES] == 0 && ctx->opt[KVX_OPT_CONSUMER_WARPS] == 0 &&
                           ctx->opt[KVX_OPT_VEC_BYTES] == 0;
    const bool tmap_case = (store_mode == 3 && tmap_shape) || tmap_auto;
    if (tmap_load_maps != nullptr) {
      p.src_tmaps = tmap_load_maps;
      p.tm_head_bytes = ls.N * ls.R;
      p.tm_head0 = ls.sg.off;
      auto k = kvx::kvx_xfer_tmapld_kernel;
      int32_t rc = set_smem(ctx, k, smem);
      if (rc) return rc;
      uint32_t tgrid = grid;
      if (tmap_load_auto && ctx->opt[KVX_OPT_CTAS_PER_SM] == 0 && ctx->opt[KVX_OPT_MAX_CTAS] == 0)
        tgrid = std::min<uint32_t>(static_cast<uint32_t>(ctx->sms) * 2u, p.total_items);
      KVX_CUDA(ctx, launch_kernel(k, tgrid, 64u, smem, stream, pdl, p));
      KVX_CUDA(ctx, cudaGetLastError());
      ctx->launches++;
      return KVX_OK;
    }
    if (tmap_case) {
      const void* maps = nullptr;
      int32_t rc = pool_tensor_maps(ctx, ls.dst_pool, &maps);
      if (rc == KVX_OK) {
        p.dst_tmaps = maps;
        p.tm_head_bytes = ls.N * ls.R;
        p.tm_head0 = ls.dg.off;
        auto k = kvx::kvx_xfer_tmap_kernel;
        rc = set_smem(ctx, k, smem);
        if (rc) return rc;
        uint32_t tgrid = grid;
        if (tmap_auto && ctx->opt[KVX_OPT_CTAS_PER_SM] == 0 && ctx->opt[KVX_OPT_MAX_CTAS] == 0)
          tgrid = std::min<uint32_t>(static_cast<uint32_t>(ctx->sms) * 2u, p.total_items);
        KVX_CUDA(ctx, launch_kernel(k, tgrid, 64u, smem, stream, pdl, p));
        KVX_CUDA(ctx, cudaGetLastError());
        ctx->launches++;
        return KVX_OK;
      }
      if (rc != KVX_EUNSUPPORTED) return rc;
    }
    if (store_mode == 2) {
      if (stages < 2) return KVX_EINVAL;
      auto k = kvx::kvx_xfer_tma_kernel<0, 1>;
      int32_t rc = set_smem(ctx, k, smem);
      if (rc) return rc;
      KVX_CUDA(ctx, launch_kernel(k, grid, 32u * (1u + cons_warps), smem, stream, pdl, p));
    } else {
      const uint32_t T = cons_warps * 32u;
      if (vec_opt == 32 && static_cast<uint64_t>(T) * 4u * 32u >= pl.piece_bytes) {
        auto k = kvx::kvx_xfer_tma_kernel<32, 4>;
        int32_t rc = set_smem(ctx, k, smem);
        if (rc) return rc;
        KVX_CUDA(ctx, launch_kernel(k, grid, 32u + T, smem, stream, pdl, p));
      } else if (static_cast<uint64_t>(T) * 8u * 16u >= pl.piece_bytes) {
        if (ls.h_src_tab != nullptr) {
          const uint32_t entries = param_entries(ls);
          int32_t rc = entries <= kParamTabSmall ? launch_ptab<kParamTabSmall>(ctx, p, ls, grid, 32u + T, smem, stream, pdl)
                       : entries <= kParamTabMid ? launch_ptab<kParamTabMid>(ctx, p, ls, grid, 32u + T, smem, stream, pdl)
                       : entries <= kParamTabBig ? launch_ptab<kParamTabBig>(ctx, p, ls, grid, 32u + T, smem, stream, pdl)
                                                 : launch_ptab<kParamTabMax>(ctx, p, ls, grid, 32u + T, smem, stream, pdl);
          if (rc) return rc;
        } else {
          auto k = kvx::kvx_xfer_tma_kernel<16, 8>;
          int32_t rc = set_smem(ctx, k, smem);
          if (rc) return rc;
          KVX_CUDA(ctx, launch_kernel(k, grid, 32u + T, smem, stream, pdl, p));
        }
      } else {
        return KVX_EINVAL;  // stage larger than the consumer threads can drain: raise CONSUMER_WARPS
      }
    }
  } else {
    if (vec_opt == 32 && pl.piece_bytes <= ldg_cap32) {
      KVX_CUDA(ctx, launch_kernel(kvx::kvx_xfer_ldg_kernel<32, 4>, grid, kLdgThreads, 0, stream, pdl, p));
    } else {
      KVX_CUDA(ctx, launch_kernel(kvx::kvx_xfer_ldg_kernel<16, 8>, grid, kLdgThreads, 0, stream, pdl, p));
    }
  }
  KVX_CUDA(ctx, cudaGetLastError());
  ctx->launches++;
  return KVX_OK;
}

int32_t resolve_xform(const kvx_pool_desc* src, const kvx_pool_desc* dst, const kvx_xform* xf, uint32_t* s_off,
                      uint32_t* d_off, uint32_t* heads, uint32_t* stage_layout, uint32_t dflt_layout) {
  uint32_t so = 0, dof = 0, hc = 0, sl = dflt_layout;
  if (xf) {
    so = xf->src_head_off; dof = xf->dst_head_off; hc = xf->heads; sl = xf->stage_layout;
  }
  if (sl != KVX_LAYOUT_NHD && sl != KVX_LAYOUT_HND) return KVX_EINVAL;
  if (hc == 0) {
    uint32_t a = src ? src->heads - std::min(so, src->heads) : UINT32_MAX;
    uint32_t b = dst ? dst->heads - std::min(dof, dst->heads) : UINT32_MAX;
    hc = std::min(a, b);
  }
  if (hc == 0) return KVX_EINVAL;
  if (src && so + hc > src->heads) return KVX_EINVAL;
  if (dst && dof + hc > dst->heads) return KVX_EINVAL;
  *s_off = so; *d_off = dof; *heads = hc; *stage_layout = sl;
  return KVX_OK;
}

XferSide paged_side(const kvx_pool* pool, const int32_t* table) {
  XferSide s{};
  s.bases = pool->d_bases;
  s.table = table;
  s.block_stride = pool->desc.block_stride;
  s.num_blocks = pool->desc.num_blocks;
  return s;
}

XferSide contiguous_side(const void* base, uint64_t chunk_bytes) {
  XferSide s{};
  s.bases = nullptr;
  s.base0 = reinterpret_cast<uint64_t>(base);
  s.block_stride = chunk_bytes;
  return s;
}

// Pools that can be joined by one transfer: same regions and row size; block sizes may differ by an
// integer ratio (regrouped in flight).
bool same_shape(const kvx_pool_desc& a, const kvx_pool_desc& b) {
  const uint32_t lo = std::min(a.block_tokens, b.block_tokens), hi = std::max(a.block_tokens, b.block_tokens);
  return a.regions == b.regions && lo != 0 && hi % lo == 0 && a.head_dim == b.head_dim && a.elem_bytes == b.elem_bytes;
}

uint32_t table_entries(uint32_t n, const kvx_pool_desc& side, const kvx_pool_desc& other) {
  const uint32_t k = side.block_tokens / std::min(side.block_tokens, other.block_tokens);
  return (n + k - 1) / k;
}

typedef CUresult (*PFN_cuMemGetAddressRange)(CUdeviceptr*, size_t*, CUdeviceptr);

int32_t alloc_base_of(kvx_ctx* ctx, const void* ptr, uint64_t* base, uint64_t* size) {
  static PFN_cuMemGetAddressRange fn = nullptr;
  if (!fn) {
    void* f = nullptr;
    cudaDriverEntryPointQueryResult q;
    KVX_CUDA(ctx, cudaGetDriverEntryPoint("cuMemGetAddressRange", &f, cudaEnableDefault, &q));
    if (!f || q != cudaDriverEntryPointSuccess) return KVX_EIPC;
    fn = reinterpret_cast<PFN_cuMemGetAddressRange>(f);
  }
  CUdeviceptr b = 0;
  size_t sz = 0;
  if (fn(&b, &sz, reinterpret_cast<CUdeviceptr>(ptr)) != CUDA_SUCCESS) return KVX_EIPC;
  *base = static_cast<uint64_t>(b);
  *size = sz;
  return KVX_OK;
}

int32_t export_record(kvx_ctx* ctx, const void* ptr, uint64_t bytes, MemRecord* rec) {
  DeviceGuard guard(ctx->device);
  memset(rec, 0, sizeof(*rec));
  rec->magic = kMemMagic;
  rec->device = ctx->device;
  rec->pid = static_cast<int64_t>(getpid());
  rec->raw_ptr = reinterpret_cast<uint64_t>(ptr);
  rec->bytes = bytes;
  memcpy(rec->device_uuid, ctx->uuid, 16);  // exported memory is memory of this context's device
  uint64_t base = 0, size = 0;
  int32_t rc = alloc_base_of(ctx, ptr, &base, &size);
  if (rc) return rc;
  rec->offset = rec->raw_ptr - base;
  if (rec->offset + bytes > size) return KVX_EINVAL;
  cudaError_t e = cudaIpcGetMemHandle(&rec->handle, reinterpret_cast<void*>(base));
  if (e == cudaSuccess) {
    rec->flags |= kMemHasIpcHandle;
  } else {
    // Not a cudaMalloc allocation (VMM / expandable segments, a memory pool): the record still serves a context
    // in THIS process, and says so — another process gets KVX_EIPC with this reason instead of a zero handle.
    (void)cudaGetLastError();
    memset(&rec->handle, 0, sizeof(rec->handle));
  }
  return KVX_OK;
}

int32_t import_record(kvx_ctx* ctx, const MemRecord* rec, void** out, bool* local = nullptr) {
  if (rec->magic != kMemMagic) return KVX_EINVAL;
  DeviceGuard guard(ctx->device);
  // cudaPointerGetAttributes reports the IMPORTING device for memory mapped through an IPC handle, so where the
  // bytes live is taken from the exporter's record
  const bool is_local = memcmp(rec->device_uuid, ctx->uuid, 16) == 0;
  if (local) *local = is_local;
  if (rec->pid == static_cast<int64_t>(getpid())) {
    if (rec->device != ctx->device) {
      int can = 0;
      KVX_CUDA(ctx, cudaDeviceCanAccessPeer(&can, ctx->device, rec->device));
      if (!can) return KVX_EIPC;
      cudaError_t e = cudaDeviceEnablePeerAccess(rec->device, 0);
      if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) return cuda_fail(ctx, e, "cudaDeviceEnablePeerAccess");
      (void)cudaGetLastError();
    }
    *out = reinterpret_cast<void*>(rec->raw_ptr);
    ctx->imported[*out] = kvx_ctx::Imported{std::string(), rec->bytes, is_local};
    return KVX_OK;
  }
  if (!(rec->flags & kMemHasIpcHandle)) {
    ctx->last_error = "the exporter's memory has no CUDA IPC handle (not a plain cudaMalloc allocation): allocate the "
                      "pool with kvx_mem_alloc / cudaMalloc, or switch off expandable segments";
    return KVX_EIPC;
  }
  const std::string key(reinterpret_cast<const char*>(&rec->handle), sizeof(rec->handle));
  std::lock_guard<std::mutex> lk(g_ipc_mu);
  auto it = g_ipc_open.find(key);
  void* base = nullptr;
  if (it != g_ipc_open.end()) {
    base = it->second.base;
    it->second.refs++;
  } else {
    cudaError_t e = cudaIpcOpenMemHandle(&base, rec->handle, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) {
      cuda_fail(ctx, e, "cudaIpcOpenMemHandle");
      return KVX_EIPC;
    }
    g_ipc_open[key] = IpcMapping{base, 1};
  }
  *out = static_cast<char*>(base) + rec->offset;
  ctx->imported[*out] = kvx_ctx::Imported{key, rec->bytes, is_local};
  return KVX_OK;
}

int32_t unmap_ptr(kvx_ctx* ctx, void* ptr) {
  auto it = ctx->imported.find(ptr);
  if (it == ctx->imported.end()) return KVX_EINVAL;
  const std::string key = it->second.key;
  ctx->imported.erase(it);
  if (key.empty()) return KVX_OK;
  std::lock_guard<std::mutex> lk(g_ipc_mu);
  auto m = g_ipc_open.find(key);
  if (m != g_ipc_open.end() && --m->second.refs == 0) {
    DeviceGuard guard(ctx->device);
    cudaIpcCloseMemHandle(m->second.base);
    g_ipc_open.erase(m);
  }
  return KVX_OK;
}

// Does this address live on the context's own GPU?  Imported ranges answer from the exporter's record; for
// memory allocated in this process the pointer attributes are reliable.
bool pointer_is_local(kvx_ctx* ctx, const void* ptr) {
  const uintptr_t a = reinterpret_cast<uintptr_t>(ptr);
  for (const auto& kv : ctx->imported) {
    const uintptr_t b = reinterpret_cast<uintptr_t>(kv.first);
    if (a >= b && a < b + kv.second.bytes) return kv.second.local;
  }
  cudaPointerAttributes at{};
  const bool remote = cudaPointerGetAttributes(&at, ptr) == cudaSuccess && at.type == cudaMemoryTypeDevice && at.device != ctx->device;
  (void)cudaGetLastError();
  return !remote;
}

// A cross-layer pool keeps every region of a block inside that block (region bases less than one block stride
// apart, (nb, L, 2, N, H, D) or (nb, H, L, 2, N, D)): walking the items block by block then touches one
// contiguous block of memory at a time instead of one chunk per block.
bool regions_inside_one_block(const kvx_pool& pool) {
  if (pool.base.size() < 2) return false;
  uintptr_t lo = UINTPTR_MAX, hi = 0;
  for (void* b : pool.base) {
    lo = std::min(lo, reinterpret_cast<uintptr_t>(b));
    hi = std::max(hi, reinterpret_cast<uintptr_t>(b));
  }
  return hi - lo < pool.desc.block_stride;
}

struct HostTables {
  const int32_t* src;
  const int32_t* dst;
  uint32_t ns, nd;
};
struct NotifyList {
  const uint64_t* pairs;  // (address, value) x count
  uint32_t count;
};

// ---- table ring: device scratch (+ pinned mirror) for block-id lists too long for the parameter block ----------
// Spans are handed out front to back and wrap; each span in use is remembered with an event recorded behind the
// kernel that reads it, and a span overlapping it is not handed out before that event has fired.
int32_t table_ring_take(kvx_ctx* ctx, uint32_t entries, uint32_t* at) {
  if (entries > kTableRing) return KVX_ERANGE;
  if (!ctx->h_table_ring) {
    cudaError_t e = cudaHostAlloc(reinterpret_cast<void**>(&ctx->h_table_ring), static_cast<size_t>(kTableRing) * sizeof(int),
                                  cudaHostAllocDefault);
    if (e != cudaSuccess) {
      (void)cudaGetLastError();
      ctx->h_table_ring = nullptr;
      return KVX_ENOMEM;
    }
  }
  if (ctx->table_cursor + static_cast<uint64_t>(entries) > kTableRing) ctx->table_cursor = 0;
  const uint32_t b = ctx->table_cursor, e = b + entries;
  for (size_t i = 0; i < ctx->ring_uses.size();) {
    kvx_ctx::RingUse& u = ctx->ring_uses[i];
    const bool overlaps = u.begin < e && b < u.end;
    if (overlaps || cudaEventQuery(u.done) == cudaSuccess) {
      if (overlaps) KVX_CUDA(ctx, cudaEventSynchronize(u.done));
      cudaEventDestroy(u.done);
      ctx->ring_uses[i] = ctx->ring_uses.back();
      ctx->ring_uses.pop_back();
    } else {
      (void)cudaGetLastError();  // cudaErrorNotReady is not an error here
      ++i;
    }
  }
  ctx->table_cursor = e;
  *at = b;
  return KVX_OK;
}

int32_t table_ring_mark(kvx_ctx* ctx, uint32_t at, uint32_t entries, cudaStream_t st) {
  cudaEvent_t ev = nullptr;
  KVX_CUDA(ctx, cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
  cudaError_t e = cudaEventRecord(ev, st);
  if (e != cudaSuccess) {
    cudaEventDestroy(ev);
    return cuda_fail(ctx, e, "cudaEventRecord");
  }
  ctx->ring_uses.push_back(kvx_ctx::RingUse{at, at + entries, ev});
  return KVX_OK;
}

// The fused handoff behind kvx_handoff_fused (device tables) and kvx_handoff_host (host tables `ht`,
// tried first as kernel parameters; completion word `flag2`).
int32_t handoff_impl(kvx_ctx* ctx, const kvx_pool* src, const int32_t* d_src_table, const kvx_pool* dst,
                     const int32_t* d_dst_table, uint32_t n, const kvx_xform* xf, uint64_t* flag, uint64_t flag_val,
                     void* stream, const HostTables* ht, const NotifyList* notify, uint64_t* flag2) {
  if (!same_shape(src->desc, dst->desc)) return KVX_EINVAL;
  uint32_t so, dof, hc, sl;
  int32_t rc = resolve_xform(&src->desc, &dst->desc, xf, &so, &dof, &hc, &sl, src->desc.layout);
  if (rc) return rc;
  const kvx_pool_desc& s = src->desc;
  const kvx_pool_desc& d = dst->desc;
  LaunchSpec ls;
  ls.src = paged_side(src, d_src_table);
  ls.dst = paged_side(dst, d_dst_table);
  ls.n = n; ls.regions = s.regions;
  ls.N = std::min(s.block_tokens, d.block_tokens);  // granule = the smaller block
  ls.R = s.head_dim * s.elem_bytes; ls.Hc = hc;
  ls.sg = ChunkGeom{s.layout, s.heads, so, s.block_tokens, s.head_stride};
  ls.dg = ChunkGeom{d.layout, d.heads, dof, d.block_tokens, d.head_stride};
  ls.flag = flag;
  ls.flag_val = flag_val;
  ls.flag2 = flag2;
  ls.block_major = regions_inside_one_block(*src) || regions_inside_one_block(*dst);
  ls.dst_pool = dst;
  ls.src_pool = src;
  ls.dst_is_peer = !dst->local;
  ls.dst_is_peer_known = ls.dst_is_peer;
  ls.src_is_peer = !src->local;
  if (ht) {
    ls.h_src_tab = ht->src; ls.h_dst_tab = ht->dst; ls.ns = ht->ns; ls.nd = ht->nd;
  }
  if (notify) {
    ls.notify = notify->pairs; ls.notify_count = notify->count;
  }
  return launch_xfer(ctx, ls, &src->base, &dst->base, static_cast<cudaStream_t>(stream));
}

}  // namespace

// =================================================================================================
// C ABI
// =================================================================================================
extern "C" {

uint32_t kvx_version(void) { return KVX_ABI_VERSION; }

const char* kvx_strerror(int32_t code) {
  switch (code) {
    case KVX_OK: return "ok";
    case KVX_EINVAL: return "invalid argument";
    case KVX_ECUDA: return "CUDA error (see kvx_last_error)";
    case KVX_ENOMEM: return "out of memory";
    case KVX_EUNSUPPORTED: return "unsupported geometry";
    case KVX_ETIMEOUT: return "timed out waiting for a completion flag";
    case KVX_ERANGE: return "size out of range";
    case KVX_EIPC: return "peer memory mapping failed";
    case KVX_ENOTREADY: return "not ready";
    default: return "unknown kvx error";
  }
}

const char* kvx_last_error(const kvx_ctx* ctx) { return ctx ? ctx->last_error.c_str() : ""; }

int32_t kvx_ctx_create(int32_t device, kvx_ctx** out) try {
  if (!out) return KVX_EINVAL;
  *out = nullptr;
  int count = 0;
  if (cudaGetDeviceCount(&count) != cudaSuccess || device < 0 || device >= count) {
    (void)cudaGetLastError();
    return KVX_ECUDA;  // no CPU fallback
  }
  kvx_ctx* ctx = new (std::nothrow) kvx_ctx();
  if (!ctx) return KVX_ENOMEM;
  ctx->device = device;
  DeviceGuard guard(device);
  auto fail = [&](int32_t rc) {
    kvx_ctx_destroy(ctx);
    return rc;
  };
  if (!guard.ok) return fail(KVX_ECUDA);
  if (cudaDeviceGetAttribute(&ctx->sms, cudaDevAttrMultiProcessorCount, device) != cudaSuccess) return fail(KVX_ECUDA);
  {
    cudaDeviceProp prop{};
    if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return fail(KVX_ECUDA);
    static_assert(sizeof(prop.uuid) == 16, "device UUID size");
    memcpy(ctx->uuid, &prop.uuid, 16);
  }
  if (cudaMalloc(&ctx->d_flags, (kNumFlags + kPrivFlags) * sizeof(uint64_t)) != cudaSuccess) return fail(KVX_ENOMEM);
  if (cudaMalloc(&ctx->d_counters, (kNumCounters + 1) * sizeof(unsigned)) != cudaSuccess) return fail(KVX_ENOMEM);
  if (cudaMalloc(&ctx->d_table_ring, static_cast<size_t>(kTableRing) * sizeof(int)) != cudaSuccess) return fail(KVX_ENOMEM);
  if (cudaMemset(ctx->d_flags, 0, (kNumFlags + kPrivFlags) * sizeof(uint64_t)) != cudaSuccess) return fail(KVX_ECUDA);
  if (cudaMemset(ctx->d_counters, 0, (kNumCounters + 1) * sizeof(unsigned)) != cudaSuccess) return fail(KVX_ECUDA);
  ctx->d_status = ctx->d_counters + kNumCounters;
  if (cudaStreamCreateWithFlags(&ctx->aux, cudaStreamNonBlocking) != cudaSuccess) return fail(KVX_ECUDA);
  ctx->opt[KVX_OPT_ENGINE] = env_or("KVX_ENGINE", 0);
  ctx->opt[KVX_OPT_STAGES] = env_or("KVX_STAGES", 0);
  ctx->opt[KVX_OPT_STAGE_BYTES] = env_or("KVX_STAGE_BYTES", 0);
  ctx->opt[KVX_OPT_CTAS_PER_SM] = env_or("KVX_CTAS_PER_SM", 0);
  ctx->opt[KVX_OPT_STORE_MODE] = env_or("KVX_STORE_MODE", 0);
  ctx->opt[KVX_OPT_VEC_BYTES] = env_or("KVX_VEC_BYTES", 0);
  ctx->opt[KVX_OPT_CONSUMER_WARPS] = env_or("KVX_CONSUMER_WARPS", 0);
  ctx->opt[KVX_OPT_L2_HINT] = env_or("KVX_L2_HINT", 0);
  ctx->opt[KVX_OPT_PDL] = env_or("KVX_PDL", 1);
  ctx->opt[KVX_OPT_PARAM_TABLES] = env_or("KVX_PARAM_TABLES", 8000);
  ctx->opt[KVX_OPT_MAX_CTAS] = env_or("KVX_MAX_CTAS", 0);
  ctx->force_sys_fence = env_or("KVX_FORCE_SYS_FENCE", 0) != 0;
  *out = ctx;
  return KVX_OK;
} KVX_CATCH_ALL

int32_t kvx_ctx_destroy(kvx_ctx* ctx) try {
  if (!ctx) return KVX_EINVAL;
  while (!ctx->queues.empty()) kvx_queue_stop(ctx->queues.back());  // a resident kernel would block the sync below
  {
    DeviceGuard guard(ctx->device);
    if (guard.ok) {
      cudaDeviceSynchronize();
      while (!ctx->imported.empty()) unmap_ptr(ctx, ctx->imported.begin()->first);
      if (ctx->aux) cudaStreamDestroy(ctx->aux);
      cudaFree(ctx->d_flags);
      cudaFree(ctx->d_counters);
      cudaFree(ctx->d_table_ring);
      while (!ctx->vmm_allocs.empty()) kvx_mem_free_vmm(ctx, ctx->vmm_allocs.begin()->first);
      for (auto& u : ctx->ring_uses) cudaEventDestroy(u.done);
      if (ctx->h_table_ring) cudaFreeHost(ctx->h_table_ring);
      (void)cudaGetLastError();
    }
  }
  delete ctx;
  return KVX_OK;
} KVX_CATCH_ALL

int32_t kvx_ctx_set_option(kvx_ctx* ctx, int32_t option, int64_t value) try {
  if (!ctx || option < 0 || option >= KVX_OPT__COUNT) return KVX_EINVAL;
  switch (option) {
    case KVX_OPT_ENGINE: if (value < 0 || value > 2) return KVX_EINVAL; break;
    case KVX_OPT_STAGES: if (value < 0 || value > kvx::kMaxStages) return KVX_EINVAL; break;
    case KVX_OPT_STAGE_BYTES: if (value != 0 && (value < 1024 || value > 200 * 1024 || value % 1024)) return KVX_EINVAL; break;
    case KVX_OPT_CTAS_PER_SM: if (value < 0 || value > 32) return KVX_EINVAL; break;
    case KVX_OPT_STORE_MODE: if (value < 0 || value > 3) return KVX_EINVAL; break;
    case KVX_OPT_VEC_BYTES: if (value != 0 && value != 16 && value != 32) return KVX_EINVAL; break;
    case KVX_OPT_CONSUMER_WARPS: if (value < 0 || value > 16) return KVX_EINVAL; break;
    case KVX_OPT_L2_HINT: if (value < 0 || value > 1) return KVX_EINVAL; break;
    case KVX_OPT_PDL: if (value < 0 || value > 1) return KVX_EINVAL; break;
    case KVX_OPT_PARAM_TABLES: if (value < 0 || value > kParamTabMax) return KVX_EINVAL; break;
    case KVX_OPT_MAX_CTAS: if (value < 0 || value > 65535) return KVX_EINVAL; break;
    default: break;
  }
  ctx->opt[option] = value;
  return KVX_OK;
} KVX_CATCH_ALL

int32_t kvx_ctx_get_option(const kvx_ctx* ctx, int32_t option, int64_t* value) try {
  if (!ctx || !value || option < 0 || option >= KVX_OPT__COUNT) return KVX_EINVAL;
  *value = ctx->opt[option];
  return KVX_OK;
} KVX_CATCH_ALL

int32_t kvx_ctx_launch_count(const kvx_ctx* ctx, uint64_t* count) try {
  if (!ctx || !count) return KVX_EINVAL;
  *count = ctx->launches;
  return KVX_OK;
} KVX_CATCH_ALL

int32_t kvx_ctx_sm_count(const kvx_ctx* ctx, int32_t* sms) try {
  if (!ctx || !sms) return KVX_EINVAL;
  *sms = ctx->sms;
  return KVX_OK;
} KVX_CATCH_ALL

int32_t kvx_mem_alloc(kvx_ctx* ctx, uint64_t bytes, void** out) try {
  if (!ctx || !out || bytes == 0) return KVX_EINVAL;
  DeviceGuard guard(ctx->device);
  cudaError_t e = cudaMalloc(out, bytes);
  if (e == cudaErrorMemoryAllocation) {
    (void)cudaGetLastError();
    return KVX_ENOMEM;
  }
  if (e != cudaSuccess) return cuda_fail(ctx, e, "cudaMalloc");
  return KVX_OK;
} KVX_CATCH_ALL

int32_t kvx_mem_free(kvx_ctx* ctx, void* ptr) try {
  if (!ctx) return KVX_EINVAL;
  DeviceGuard guard(ctx->device);
  KVX_CUDA(ctx, cudaFree(ptr));
  return KVX_OK;
} KVX_CATCH_ALL

int32_t kvx_mem_export(kvx_ctx* ctx, const void* ptr, uint64_t bytes, void* buf, size_t* len) try {
  if (!ctx || !ptr || !len) return KVX_EINVAL;
  if (!buf || *len < sizeof(MemRecord)) {
    *len = sizeof(MemRecord);
    return KVX_ERANGE;
  }
  MemRecord rec;
  int32_t rc = export_record(ctx, ptr, bytes, &rec);
  if (rc) return rc;
  memcpy(buf, &rec, sizeof(rec));
  *len = sizeof(rec);
  return KVX_OK;
} KVX_CATCH_ALL

int32_t kvx_mem_import(kvx_ctx* ctx, const void* buf, size_t len, void** out) try {
  if (!ctx || !buf || !out || len < sizeof(MemRecord)) return KVX_EINVAL;
  MemRecord rec;
  memcpy(&rec, buf, sizeof(rec));
  return import_record(ctx, &rec, out);
} KVX_CATCH_ALL

int32_t kvx_mem_unmap(kvx_ctx* ctx, void* imported) try {
  if (!ctx || !imported) return KVX_EINVAL;
  return unmap_ptr(ctx, imported);
} KVX_CATCH_ALL

int32_t kvx_pool_register(kvx_ctx* ctx, const kvx_pool_desc* desc, kvx_pool** out) try {
  if (!ctx || !desc || !out || !desc->base) return KVX_EINVAL;
  if (desc->struct_size != sizeof(kvx_pool_desc)) return KVX_EINVAL;
  if (desc->regions == 0 || desc->block_tokens == 0 || desc->heads == 0 || desc->head_dim == 0 ||
      desc->elem_bytes == 0 || desc->num_blocks == 0)
    return KVX_EINVAL;
  if (desc->layout != KVX_LAYOUT_NHD && desc->layout != KVX_LAYOUT_HND) return KVX_EINVAL;
  const uint64_t row = static_cast<uint64_t>(desc->head_dim) * desc->elem_bytes;
  if (row % 16 != 0) return KVX_EUNSUPPORTED;  // 16-byte vector / bulk-copy granularity
  const uint64_t chunk = row * desc->heads * desc->block_tokens;
  if (chunk >= (1ull << 31)) return KVX_EUNSUPPORTED;
  const uint64_t stride = desc->block_stride ? desc->block_stride : chunk;
  if (stride < chunk || stride % 16 != 0) return KVX_EINVAL;
  if (desc->head_stride != 0) {  // heads of a chunk spread apart (cross-layer HND pools)
    if (desc->layout != KVX_LAYOUT_HND || desc->head_stride % 16 != 0 || desc->head_stride < row * desc->block_tokens ||
        static_cast<uint64_t>(desc->head_stride) * desc->heads > stride)
      return KVX_EINVAL;
  }
  for (uint32_t r = 0; r < desc->regions; ++r) {
    if (!desc->base[r] || reinterpret_cast<uintptr_t>(desc->base[r]) % 16 != 0) return KVX_EINVAL;
  }
  kvx_pool* pool = new (std::nothrow) kvx_pool();
  if (!pool) return KVX_ENOMEM;
  pool->ctx = ctx;
  {
    // Where the bytes live decides the launch geometry and the fence scope, not how the pointer was obtained:
    // a pool imported from another process on the SAME GPU is local memory, a peer's pool is not.
    DeviceGuard g(ctx->device);
    pool->local = pointer_is_local(ctx, desc->base[0]);
  }
  pool->desc = *desc;
  pool->desc.block_stride = stride;
  pool->base.assign(desc->base, desc->base + desc->regions);
  pool->desc.base = pool->base.data();
  pool->chunk_bytes = chunk;
  DeviceGuard guard(ctx->device);
  cudaError_t e = cudaMalloc(&pool->d_bases, desc->regions * sizeof(unsigned long long));
  if (e == cudaSuccess)
    e = cudaMemcpy(pool->d_bases, pool->base.data(), desc->regions * sizeof(unsigned long long), cudaMemcpyHostToDevice);
  if (e != cudaSuccess) {
    int32_t rc = cuda_fail(ctx, e, "pool base upload");
    cudaFree(pool->d_bases);
    delete pool;
    return rc;
  }
  *out = pool;
  return KVX_OK;
} KVX_CATCH_ALL

int32_t kvx_pool_release(kvx_pool* pool) try {
  if (!pool) return KVX_EINVAL;
  DeviceGuard guard(pool->ctx->device);
  for (void* p : pool->imported_ptrs) unmap_ptr(pool->ctx, p);
  cudaFree(pool->d_bases);
  if (pool->d_tmaps) cudaFree(pool->d_tmaps);
  if (!pool->vmm_ranges.empty()) {
    cudaDeviceSynchronize();
    pool_unmap_vmm(pool);
  }
  (void)cudaGetLastError();
  delete pool;
  return KVX_OK;
} KVX_CATCH_ALL

int32_t kvx_pool_chunk_bytes(const kvx_pool* pool, uint64_t* bytes) try {
  if (!pool || !bytes) return KVX_EINVAL;
  *bytes = pool->chunk_bytes;
  return KVX_OK;
} KVX_CATCH_ALL

int32_t kvx_pool_export(kvx_pool* pool, void* buf, size_t* len) try {
  if (!pool || !len) return KVX_EINVAL;
  const size_t need = sizeof(PoolRecordHeader) + pool->base.size() * sizeof(MemRecord);
  if (!buf || *len < need) {
    *len = need;
    return KVX_ERANGE;
  }
  PoolRecordHeader h{};
  const kvx_pool_desc& d = pool->desc;
  h.magic = kPoolMagic; h.regions = d.regions; h.block_tokens = d.block_tokens; h.heads = d.heads;
  h.head_dim = d.head_dim; h.elem_bytes = d.elem_bytes; h.layout = d.layout; h.tp_rank = d.tp_rank;
  h.tp_size = d.tp_size; h.head_stride = d.head_stride; h.num_blocks = d.num_blocks; h.block_stride = d.block_stride;
  memcpy(buf, &h, sizeof(h));
  char* w = static_cast<char*>(buf) + sizeof(h);
  const uint64_t row_bytes = static_cast<uint64_t>(d.head_dim) * d.elem_bytes;
  const uint64_t chunk_extent = d.head_stride ? static_cast<uint64_t>(d.heads - 1) * d.head_stride + row_bytes * d.block_tokens
                                              : pool->chunk_bytes;
  const uint64_t span = (d.num_blocks - 1) * d.block_stride + chunk_extent;
  for (size_t r = 0; r < pool->base.size(); ++r) {
    MemRecord rec;
    int32_t rc = export_record(pool->ctx, pool->base[r], span, &rec);
    if (rc) return rc;
    memcpy(w + r * sizeof(MemRecord), &rec, sizeof(rec));
  }
  *len = need;
  return KVX_OK;
} KVX_CATCH_ALL

int32_t kvx_pool_region_base(const kvx_pool* pool, uint32_t region, void** base) try {
  if (!pool || !base || region >= pool->base.size()) return KVX_EINVAL;
  *base = pool->base[region];
  return KVX_OK;
} KVX_CATCH_ALL

int32_t kvx_pool_slice(const kvx_pool* pool, uint32_t first_region, uint32_t count, kvx_pool** out) try {
  if (!pool || !out || count == 0 || static_cast<uint64_t>(first_region) + count > pool->base.size()) return KVX_EINVAL;
  kvx_pool_desc d = pool->desc;
  d.regions = count;
  d.base = pool->base.data() + first_region;
  kvx_pool* view = nullptr;
  int32_t rc = kvx_pool_register(pool->ctx, &d, &view);
  if (rc) return rc;
  view->local = pool->local;  // same memory as the parent, whoever mapped it
  *out = view;
  return KVX_OK;
} KVX_CATCH_ALL

int32_t kvx_pool_exportable(kvx_pool* pool, int32_t* across_processes) try {
  if (!pool || !across_processes) return KVX_EINVAL;
  *across_processes = 1;
  for (void* b : pool->base) {
    MemRecord rec;
    int32_t rc = export_record(pool->ctx, b, 16, &rec);
    if (rc) return rc;
    if (!(rec.flags & kMemHasIpcHandle)) *across_processes = 0;
  }
  return KVX_OK;
} KVX_CATCH_ALL

int32_t kvx_pool_import(kvx_ctx* ctx, const void* buf, size_t len, kvx_pool** out) try {
  if (!ctx || !buf || !out || len < sizeof(PoolRecordHeader)) return KVX_EINVAL;
  PoolRecordHeader h;
  memcpy(&h, buf, sizeof(h));
  if (h.magic != kPoolMagic) return KVX_EINVAL;
  if (len < sizeof(h) + static_cast<size_t>(h.regions) * sizeof(MemRecord)) return KVX_EINVAL;
  std::vector<void*> bases(h.regions, nullptr);
  bool all_local = true;
  const char* r = static_cast<const char*>(buf) + sizeof(h);
  int32_t rc = KVX_OK;
  uint32_t done = 0;
  for (; done < h.regions; ++done) {
    MemRecord rec;
    memcpy(&rec, r + done * sizeof(MemRecord), sizeof(rec));
    bool rec_local = true;
    rc = import_record(ctx, &rec, &bases[done], &rec_local);
    if (rc) break;
    all_local = all_local && rec_local;
  }
  if (rc) {
    for (uint32_t i = 0; i < done; ++i) unmap_ptr(ctx, bases[i]);
    return rc;
  }
  kvx_pool_desc d{};
  d.struct_size = sizeof(d); d.regions = h.regions; d.block_tokens = h.block_tokens; d.heads = h.heads;
  d.head_dim = h.head_dim; d.elem_bytes = h.elem_bytes; d.layout = h.layout; d.tp_rank = h.tp_rank;
  d.tp_size = h.tp_size; d.head_stride = h.head_stride; d.num_blocks = h.num_blocks; d.block_stride = h.block_stride; d.base = bases.data();
  kvx_pool* pool = nullptr;
  rc = kvx_pool_register(ctx, &d, &pool);
  if (rc) {
    for (void* p : bases) unmap_ptr(ctx, p);
    return rc;
  }
  pool->imported_ptrs = bases;
  pool->local = all_local;  // from the exporter's records (pointer attributes cannot tell for IPC mappings)
  *out = pool;
  return KVX_OK;
} KVX_CATCH_ALL

int32_t kvx_ctx_flags(kvx_ctx* ctx, uint64_t** flags, uint32_t* count) try {
  if (!ctx || !flags || !count) return KVX_EINVAL;
  *flags = ctx->d_flags;
  *count = kNumFlags;
  return KVX_OK;
} KVX_CATCH_ALL

int32_t kvx_flag_reset(kvx_ctx* ctx, uint64_t* flag, uint64_t value, void* stream) try {
  if (!ctx || !flag) return KVX_EINVAL;
  DeviceGuard guard(ctx->device);
  KVX_CUDA(ctx, cudaMemcpyAsync(flag, &value, sizeof(value), cudaMemcpyHostToDevice, static_cast<cudaStream_t>(stream)));
  return KVX_OK;
} KVX_CATCH_ALL

int32_t kvx_flag_wait(kvx_ctx* ctx, const uint64_t* flag, uint64_t value, uint32_t timeout_ms, void* stream) try {
  if (!ctx || !flag) return KVX_EINVAL;
  DeviceGuard guard(ctx->device);
  const unsigned long long ns = static_cast<unsigned long long>(timeout_ms ? timeout_ms : 1000u) * 1000000ull;
  kvx::kvx_flag_wait_kernel<<<1, 1, 0, static_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const unsigned long long*>(flag), value, ns, ctx->d_status);
  KVX_CUDA(ctx, cudaGetLastError());
  ctx->launches++;
  return KVX_OK;
} KVX_CATCH_ALL

int32_t kvx_flag_poll(kvx_ctx* ctx, const uint64_t* flag, uint64_t value) try {
  if (!ctx || !flag) return KVX_EINVAL;
  DeviceGuard guard(ctx->device);
  uint64_t cur = 0;
  KVX_CUDA(ctx, cudaMemcpyAsync(&cur, flag, sizeof(cur), cudaMemcpyDeviceToHost, ctx->aux));
  KVX_CUDA(ctx, cudaStreamSynchronize(ctx->aux));
  return cur >= value ? KVX_OK : KVX_ENOTREADY;
} KVX_CATCH_ALL

int32_t kvx_ctx_check(kvx_ctx* ctx) try {
  if (!ctx) return KVX_EINVAL;
  DeviceGuard guard(ctx->device);
  unsigned st = 0;
  KVX_CUDA(ctx, cudaMemcpyAsync(&st, ctx->d_status, sizeof(st), cudaMemcpyDeviceToHost, ctx->aux));
  KVX_CUDA(ctx, cudaStreamSynchronize(ctx->aux));
  if (st != 0) {
    KVX_CUDA(ctx, cudaMemsetAsync(ctx->d_status, 0, sizeof(unsigned), ctx->aux));
    KVX_CUDA(ctx, cudaStreamSynchronize(ctx->aux));
  }
  if (st & kvx::kStatusWaitTimeout) return KVX_ETIMEOUT;
  if (st & kvx::kStatusBadBlockId) return KVX_ERANGE;
  return KVX_OK;
} KVX_CATCH_ALL

int32_t kvx_gather(kvx_ctx* ctx, const kvx_pool* src, const int32_t* d_table, uint32_t n, void* staging,
                   const kvx_xform* xf, void* stream) try {
  if (!ctx || !src) return KVX_EINVAL;
  if (n == 0) return KVX_OK;
  if (!d_table || !staging || reinterpret_cast<uintptr_t>(staging) % 16 != 0) return KVX_EINVAL;
  uint32_t so, dof, hc, sl;
  int32_t rc = resolve_xform(&src->desc, nullptr, xf, &so, &dof, &hc, &sl, src->desc.layout);
  if (rc) return rc;
  const kvx_pool_desc& d = src->desc;
  LaunchSpec ls;
  ls.src = paged_side(src, d_table);
  const uint64_t stage_chunk = static_cast<uint64_t>(d.block_tokens) * hc * d.head_dim * d.elem_bytes;
  ls.dst = contiguous_side(staging, stage_chunk);
  ls.n = n; ls.regions = d.regions;
  ls.N = d.block_tokens; ls.R = d.head_dim * d.elem_bytes; ls.Hc = hc;
  ls.sg = ChunkGeom{d.layout, d.heads, so, d.block_tokens, d.head_stride};
  ls.dg = ChunkGeom{sl, hc, 0, d.block_tokens};
  ls.src_is_peer = !src->local;
  return launch_xfer(ctx, ls, &src->base, nullptr, static_cast<cudaStream_t>(stream));
} KVX_CATCH_ALL

int32_t kvx_scatter(kvx_ctx* ctx, const void* staging, const kvx_pool* dst, const int32_t* d_table, uint32_t n,
                    const kvx_xform* xf, void* stream) try {
  if (!ctx || !dst) return KVX_EINVAL;
  if (n == 0) return KVX_OK;
  if (!d_table || !staging || reinterpret_cast<uintptr_t>(staging) % 16 != 0) return KVX_EINVAL;
  uint32_t so, dof, hc, sl;
  int32_t rc = resolve_xform(nullptr, &dst->desc, xf, &so, &dof, &hc, &sl, dst->desc.layout);
  if (rc) return rc;
  const kvx_pool_desc& d = dst->desc;
  LaunchSpec ls;
  const uint64_t stage_chunk = static_cast<uint64_t>(d.block_tokens) * hc * d.head_dim * d.elem_bytes;
  ls.src = contiguous_side(staging, stage_chunk);
  ls.dst = paged_side(dst, d_table);
  ls.n = n; ls.regions = d.regions;
  ls.N = d.block_tokens; ls.R = d.head_dim * d.elem_bytes; ls.Hc = hc;
  ls.sg = ChunkGeom{sl, hc, 0, d.block_tokens};
  ls.dg = ChunkGeom{d.layout, d.heads, dof, d.block_tokens, d.head_stride};
  ls.dst_is_peer_known = !dst->local;
  return launch_xfer(ctx, ls, nullptr, &dst->base, static_cast<cudaStream_t>(stream));
} KVX_CATCH_ALL

int32_t kvx_copy(kvx_ctx* ctx, const void* src, void* dst, uint64_t bytes, uint64_t* flag, uint64_t flag_val,
                 void* stream) try {
  if (!ctx) return KVX_EINVAL;
  if (bytes != 0 && (!src || !dst)) return KVX_EINVAL;
  if (bytes % 16 != 0 || reinterpret_cast<uintptr_t>(src) % 16 != 0 || reinterpret_cast<uintptr_t>(dst) % 16 != 0)
    return KVX_EINVAL;
  LaunchSpec ls;
  ls.src = contiguous_side(src, 0);
  ls.dst = contiguous_side(dst, 0);
  ls.copy_mode = true;
  ls.copy_bytes = bytes;
  ls.flag = flag;
  ls.flag_val = flag_val;
  if (bytes != 0) {  // a side that lives on another GPU makes this an NVLink transfer (launch geometry only)
    DeviceGuard guard(ctx->device);
    ls.src_is_peer = !pointer_is_local(ctx, src);
    ls.dst_is_peer_known = !pointer_is_local(ctx, dst);
  }
  if (bytes == 0) {
    if (flag) return kvx_flag_reset(ctx, flag, flag_val, stream);
    return KVX_OK;
  }
  return launch_xfer(ctx, ls, nullptr, nullptr, static_cast<cudaStream_t>(stream));
} KVX_CATCH_ALL

int32_t kvx_handoff_fused(kvx_ctx* ctx, const kvx_pool* src, const int32_t* d_src_table, const kvx_pool* dst,
                          const int32_t* d_dst_table, uint32_t n, const kvx_xform* xf, uint64_t* flag,
                          uint64_t flag_val, void* stream) try {
  if (!ctx || !src || !dst) return KVX_EINVAL;
  if (n != 0 && (!d_src_table || !d_dst_table)) return KVX_EINVAL;
  return handoff_impl(ctx, src, d_src_table, dst, d_dst_table, n, xf, flag, flag_val, stream, nullptr, nullptr, nullptr);
} KVX_CATCH_ALL

int32_t kvx_handoff_host(kvx_ctx* ctx, const kvx_pool* src, const int32_t* h_src_table, const kvx_pool* dst,
                         const int32_t* h_dst_table, uint32_t n, const kvx_xform* xf, uint64_t* flag,
                         uint64_t flag_val, uint64_t* h_done, void* stream) try {
  if (!ctx || !src || !dst) return KVX_EINVAL;
  if (n > KVX_MAX_HOST_TABLE) return KVX_ERANGE;
  if (n != 0 && (!h_src_table || !h_dst_table)) return KVX_EINVAL;
  if (!same_shape(src->desc, dst->desc)) return KVX_EINVAL;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  DeviceGuard guard(ctx->device);
  // The peer may own the destination blocks: ids are validated here against each pool's size so a
  // bad list never reaches the kernel (the kernel re-checks device-resident tables itself).
  const uint32_t ns = table_entries(n, src->desc, dst->desc), nd = table_entries(n, dst->desc, src->desc);
  for (uint32_t i = 0; i < ns; ++i)
    if (h_src_table[i] < 0 || static_cast<uint64_t>(h_src_table[i]) >= src->desc.num_blocks) return KVX_ERANGE;
  for (uint32_t i = 0; i < nd; ++i)
    if (h_dst_table[i] < 0 || static_cast<uint64_t>(h_dst_table[i]) >= dst->desc.num_blocks) return KVX_ERANGE;

  // Completion word: pinned (or registered) host memory is written by the kernel itself through its
  // device alias; pageable memory gets a D2H copy of the flag after the kernel.
  uint64_t* done_dev = nullptr;
  if (h_done) {
    cudaPointerAttributes at{};
    if (cudaPointerGetAttributes(&at, h_done) == cudaSuccess && at.type == cudaMemoryTypeHost && at.devicePointer)
      done_dev = static_cast<uint64_t*>(at.devicePointer);
    else
      (void)cudaGetLastError();
  }
  // a private word behind the exported flags (peers never see it), rotated so that calls on different streams
  // do not share one
  uint64_t* f = flag ? flag : ((h_done && !done_dev) ? ctx->d_flags + kNumFlags + (ctx->priv_cursor++ % kPrivFlags) : nullptr);

  int32_t rc = kNoParamTables;
  if (n != 0 && static_cast<int64_t>(ns) + nd <= std::min<int64_t>(kParamTabMax, ctx->opt[KVX_OPT_PARAM_TABLES])) {
    // Short lists ride in the kernel's parameter block: one launch, no copy-engine work.
    HostTables ht{h_src_table, h_dst_table, ns, nd};
    rc = handoff_impl(ctx, src, nullptr, dst, nullptr, n, xf, f, flag_val, stream, &ht, nullptr, done_dev);
  }
  if (rc == kNoParamTables) {
    int* d_src = nullptr;
    int* d_dst = nullptr;
    uint32_t at = 0;
    if (n != 0) {
      // Long lists (more than the parameter block holds) go through the table ring.  They are copied into the
      // ring's pinned mirror HERE, so the caller may reuse its lists as soon as the call returns whatever kind
      // of host memory they are in, and a span of the ring is handed out again only after the kernel that
      // read it has finished (an event per span; waiting for it is the back-pressure of a very deep backlog).
      rc = table_ring_take(ctx, ns + nd, &at);
      if (rc) return rc;
      memcpy(ctx->h_table_ring + at, h_src_table, ns * sizeof(int));
      memcpy(ctx->h_table_ring + at + ns, h_dst_table, nd * sizeof(int));
      d_src = ctx->d_table_ring + at;
      d_dst = d_src + ns;
      KVX_CUDA(ctx, cudaMemcpyAsync(d_src, ctx->h_table_ring + at, (static_cast<size_t>(ns) + nd) * sizeof(int),
                                    cudaMemcpyHostToDevice, st));
    }
    rc = handoff_impl(ctx, src, d_src, dst, d_dst, n, xf, f, flag_val, stream, nullptr, nullptr, done_dev);
    if (n != 0) {
      const int32_t rc2 = table_ring_mark(ctx, at, ns + nd, st);
      if (rc == KVX_OK) rc = rc2;
    }
  }
  if (rc) return rc;
  if (h_done && !done_dev) {
    // A peer flag is readable through the same mapping the kernel wrote it through.
    KVX_CUDA(ctx, cudaMemcpyAsync(h_done, f, sizeof(uint64_t), cudaMemcpyDeviceToHost, st));
  }
  return KVX_OK;
} KVX_CATCH_ALL

int32_t kvx_handoff_batch(kvx_ctx* ctx, const kvx_pool* src, const kvx_pool* dst, const kvx_batch_item* items,
                          uint32_t count, const kvx_xform* xf, void* stream) try {
  if (!ctx || !src || !dst || (count != 0 && !items)) return KVX_EINVAL;
  if (!same_shape(src->desc, dst->desc)) return KVX_EINVAL;
  if (count == 0) return KVX_OK;
  uint64_t total = 0, notes = 0;
  bool pinned_only = true;
  DeviceGuard guard(ctx->device);
  std::vector<uint64_t> pairs;
  pairs.reserve(4 * static_cast<size_t>(count));
  for (uint32_t k = 0; k < count; ++k) {
    const kvx_batch_item& it = items[k];
    if (it.n > KVX_MAX_HOST_TABLE) return KVX_ERANGE;
    if (it.n != 0 && (!it.h_src_table || !it.h_dst_table)) return KVX_EINVAL;
    for (uint32_t i = 0; i < it.n; ++i) {
      if (it.h_src_table[i] < 0 || static_cast<uint64_t>(it.h_src_table[i]) >= src->desc.num_blocks) return KVX_ERANGE;
      if (it.h_dst_table[i] < 0 || static_cast<uint64_t>(it.h_dst_table[i]) >= dst->desc.num_blocks) return KVX_ERANGE;
    }
    total += it.n;
    if (it.flag) {
      pairs.push_back(reinterpret_cast<uint64_t>(it.flag));
      pairs.push_back(it.flag_val);
    }
    if (it.h_done) {
      cudaPointerAttributes at{};
      if (cudaPointerGetAttributes(&at, it.h_done) == cudaSuccess && at.type == cudaMemoryTypeHost && at.devicePointer) {
        pairs.push_back(reinterpret_cast<uint64_t>(at.devicePointer));
        pairs.push_back(it.flag_val);
      } else {
        (void)cudaGetLastError();
        pinned_only = false;
      }
    }
  }
  notes = pairs.size() / 2;
  const bool one_launch = count > 1 && total != 0 && pinned_only && src->desc.block_tokens == dst->desc.block_tokens &&
                          2 * total + 4 * notes + 1 <= static_cast<uint64_t>(std::min<int64_t>(kParamTabMax, ctx->opt[KVX_OPT_PARAM_TABLES]));
  if (one_launch) {
    std::vector<int32_t> ids(2 * static_cast<size_t>(total));
    size_t at = 0;
    for (uint32_t k = 0; k < count; ++k) {
      if (items[k].n) memcpy(ids.data() + at, items[k].h_src_table, items[k].n * sizeof(int32_t));
      if (items[k].n) memcpy(ids.data() + total + at, items[k].h_dst_table, items[k].n * sizeof(int32_t));
      at += items[k].n;
    }
    HostTables ht{ids.data(), ids.data() + total, static_cast<uint32_t>(total), static_cast<uint32_t>(total)};
    NotifyList nl{pairs.data(), static_cast<uint32_t>(notes)};
    int32_t rc = handoff_impl(ctx, src, nullptr, dst, nullptr, static_cast<uint32_t>(total), xf, nullptr, 0, stream, &ht,
                              notes ? &nl : nullptr, nullptr);
    if (rc != kNoParamTables) return rc;
  }
  // one request per launch: same bytes, same notifications
  for (uint32_t k = 0; k < count; ++k) {
    const kvx_batch_item& it = items[k];
    int32_t rc = kvx_handoff_host(ctx, src, it.h_src_table, dst, it.h_dst_table, it.n, xf, it.flag, it.flag_val, it.h_done, stream);
    if (rc) return rc;
  }
  return KVX_OK;
} KVX_CATCH_ALL

}  // extern "C"

// =================================================================================================
// kvx_handoff_static: the fused handoff with all scratch in one fixed device array (see kvx.h).
// =================================================================================================
namespace {

// Layout of the array for one call, in bytes: [counter, status, padding: 256][src region bases][dst region bases]
// [src table][dst table], packed for that call's region and block counts so that ONE upload carries everything.
constexpr size_t kStaticWordsBytes = 256;  // counter at +0, status at +4
static_assert(kStaticWordsBytes + 2 * static_cast<size_t>(KVX_STATIC_MAX_REGIONS) * sizeof(unsigned long long) +
                      2 * static_cast<size_t>(KVX_STATIC_MAX_BLOCKS) * sizeof(int) == KVX_STATIC_SCRATCH_BYTES,
              "scratch layout and header limit disagree");

// In host code this name is only a handle for cudaGetSymbolAddress, never a device address.
static __device__ __align__(256) unsigned char g_static_scratch[KVX_STATIC_SCRATCH_BYTES];

constexpr int kStaticMaxDevices = 64;
constexpr uint32_t kStaticStages = 4, kStaticStageBytes = 16384, kStaticConsumerWarps = 16, kStaticCtasPerSm = 2;

struct StaticDevice {
  bool ready = false;
  unsigned char* scratch = nullptr;  // device address of g_static_scratch on this device
  int sms = 0;
};
StaticDevice g_static_dev[kStaticMaxDevices];
std::mutex g_static_mu;
// Host-side image of the array for the call being issued (ordinary pageable memory, no allocation).  The caller's
// lists and base addresses are copied and range-checked HERE, and the upload reads this copy: an H2D copy from
// pageable memory has taken its source by the time cudaMemcpyAsync returns, so the caller's own buffers — pinned or
// not — are free again on return, and the ids the kernel reads are exactly the ids that were checked.
alignas(16) unsigned char g_static_stage[KVX_STATIC_SCRATCH_BYTES];

// Once per device, with that device current: resolve the array's device address, size the kernel's
// shared memory, read the SM count.  None of these calls allocates.
int32_t static_setup(int device, StaticDevice** out) {
  StaticDevice& d = g_static_dev[device];  // the caller holds g_static_mu
  if (!d.ready) {
    std::lock_guard<std::mutex> attr_lk(g_func_attr_mu);
    void* ptr = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&ptr, g_static_scratch);
    if (e != cudaSuccess || ptr == nullptr) {
      (void)cudaGetLastError();
      return KVX_ECUDA;
    }
    // Raise the kernel's dynamic shared-memory limit to what this entry needs, never lower it: a
    // context on the same device may already have raised it further for a deeper pipeline.
    cudaFuncAttributes fa{};
    e = cudaFuncGetAttributes(&fa, kvx::kvx_xfer_tma_kernel<16, 8>);
    if (e == cudaSuccess && fa.maxDynamicSharedSizeBytes < static_cast<int>(kStaticStages * kStaticStageBytes))
      e = cudaFuncSetAttribute(kvx::kvx_xfer_tma_kernel<16, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                               static_cast<int>(kStaticStages * kStaticStageBytes));
    if (e == cudaSuccess) e = cudaDeviceGetAttribute(&d.sms, cudaDevAttrMultiProcessorCount, device);
    if (e != cudaSuccess) {
      (void)cudaGetLastError();
      return KVX_ECUDA;
    }
    d.scratch = static_cast<unsigned char*>(ptr);
    d.ready = true;
  }
  *out = &d;
  return KVX_OK;
}

int32_t static_check_desc(const kvx_pool_desc* d, uint64_t* chunk, uint64_t* stride) {
  if (!d || !d->base || d->struct_size != sizeof(kvx_pool_desc)) return KVX_EINVAL;
  if (d->regions == 0 || d->block_tokens == 0 || d->heads == 0 || d->head_dim == 0 || d->elem_bytes == 0 ||
      d->num_blocks == 0)
    return KVX_EINVAL;
  if (d->layout != KVX_LAYOUT_NHD && d->layout != KVX_LAYOUT_HND) return KVX_EINVAL;
  if (d->regions > KVX_STATIC_MAX_REGIONS) return KVX_ERANGE;
  const uint64_t row = static_cast<uint64_t>(d->head_dim) * d->elem_bytes;
  if (row % 16 != 0) return KVX_EUNSUPPORTED;
  *chunk = row * d->heads * d->block_tokens;
  if (*chunk >= (1ull << 31)) return KVX_EUNSUPPORTED;
  *stride = d->block_stride ? d->block_stride : *chunk;
  if (*stride < *chunk || *stride % 16 != 0) return KVX_EINVAL;
  if (d->head_stride != 0 && (d->layout != KVX_LAYOUT_HND || d->head_stride % 16 != 0 || d->head_stride < row * d->block_tokens ||
                              static_cast<uint64_t>(d->head_stride) * d->heads > *stride))
    return KVX_EINVAL;
  for (uint32_t r = 0; r < d->regions; ++r)
    if (!d->base[r] || reinterpret_cast<uintptr_t>(d->base[r]) % 16 != 0) return KVX_EINVAL;
  return KVX_OK;
}

}  // namespace

extern "C" int32_t kvx_handoff_static(int32_t device, const kvx_pool_desc* src, const int32_t* h_src_table,
                                      const kvx_pool_desc* dst, const int32_t* h_dst_table, uint32_t n,
                                      const kvx_xform* xf, uint64_t* flag, uint64_t flag_val, void* stream) try {
  if (device < 0 || device >= kStaticMaxDevices) return KVX_EINVAL;
  if (n > KVX_STATIC_MAX_BLOCKS) return KVX_ERANGE;  // beyond the fixed table capacity
  uint64_t s_chunk = 0, s_stride = 0, d_chunk = 0, d_stride = 0;
  int32_t rc = static_check_desc(src, &s_chunk, &s_stride);
  if (rc) return rc;
  rc = static_check_desc(dst, &d_chunk, &d_stride);
  if (rc) return rc;
  if (!same_shape(*src, *dst)) return KVX_EINVAL;
  if (src->block_tokens != dst->block_tokens) return KVX_EUNSUPPORTED;  // no block-size regrouping here
  if (n != 0 && (!h_src_table || !h_dst_table)) return KVX_EINVAL;
  uint32_t so, dof, hc, sl;
  rc = resolve_xform(src, dst, xf, &so, &dof, &hc, &sl, src->layout);
  if (rc) return rc;

  DeviceGuard guard(device);
  if (!guard.ok) {
    (void)cudaGetLastError();
    return KVX_ECUDA;
  }
  std::lock_guard<std::mutex> lk(g_static_mu);  // one host thread at a time builds and uploads the image
  StaticDevice* dev = nullptr;
  rc = static_setup(device, &dev);  // first call on this device: do it at start-up, before any stream capture
  if (rc) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  auto bad = [](cudaError_t e) {
    if (e == cudaSuccess) return false;
    (void)cudaGetLastError();
    return true;
  };
  if (n == 0) {  // nothing to move: notify only
    if (flag && bad(cudaMemcpyAsync(flag, &flag_val, sizeof(flag_val), cudaMemcpyHostToDevice, st))) return KVX_ECUDA;
    return KVX_OK;
  }

  Plan pl{};
  rc = make_plan(src->block_tokens, src->head_dim * src->elem_bytes, hc,
                 ChunkGeom{src->layout, src->heads, so, src->block_tokens, src->head_stride},
                 ChunkGeom{dst->layout, dst->heads, dof, dst->block_tokens, dst->head_stride}, kStaticStageBytes, &pl);
  if (rc) return rc;  // a head/token unit larger than one 16 KiB stage is not supported here
  const uint64_t total = static_cast<uint64_t>(n) * src->regions * pl.ppc;
  if (total >= (1ull << 31)) return KVX_EUNSUPPORTED;

  unsigned char* sc = dev->scratch;
  const uint32_t regions = src->regions;  // == dst->regions (same_shape)
  const size_t off_sbases = kStaticWordsBytes, off_dbases = off_sbases + regions * sizeof(unsigned long long);
  const size_t off_stable = off_dbases + regions * sizeof(unsigned long long), off_dtable = off_stable + n * sizeof(int);
  const size_t image_bytes = off_dtable + n * sizeof(int);
  // The array keeps whatever the previous call left: build this call's image — zeroed counter and status word,
  // base addresses, both tables — in host memory, check the ids on that copy, and upload it in one piece,
  // ordered on the caller's stream.
  memset(g_static_stage, 0, kStaticWordsBytes);
  memcpy(g_static_stage + off_sbases, src->base, regions * sizeof(void*));
  memcpy(g_static_stage + off_dbases, dst->base, regions * sizeof(void*));
  memcpy(g_static_stage + off_stable, h_src_table, n * sizeof(int));
  memcpy(g_static_stage + off_dtable, h_dst_table, n * sizeof(int));
  {
    const int* ts = reinterpret_cast<const int*>(g_static_stage + off_stable);
    const int* td = reinterpret_cast<const int*>(g_static_stage + off_dtable);
    for (uint32_t i = 0; i < n; ++i) {
      if (ts[i] < 0 || static_cast<uint64_t>(ts[i]) >= src->num_blocks) return KVX_ERANGE;
      if (td[i] < 0 || static_cast<uint64_t>(td[i]) >= dst->num_blocks) return KVX_ERANGE;
    }
  }
  if (bad(cudaMemcpyAsync(sc, g_static_stage, image_bytes, cudaMemcpyHostToDevice, st))) return KVX_ECUDA;

  XferParams p{};
  p.src.bases = reinterpret_cast<const unsigned long long*>(sc + off_sbases);
  p.src.table = reinterpret_cast<const int*>(sc + off_stable);
  p.src.block_stride = s_stride;
  p.src.num_blocks = src->num_blocks;
  p.src.slice_off = pl.s_slice_off;
  p.src.piece_stride = pl.s_piece_stride;
  p.src.ratio = 1;
  p.dst.ratio = 1;
  p.dst.bases = reinterpret_cast<const unsigned long long*>(sc + off_dbases);
  p.dst.table = reinterpret_cast<const int*>(sc + off_dtable);
  p.dst.block_stride = d_stride;
  p.dst.num_blocks = dst->num_blocks;
  p.dst.slice_off = pl.d_slice_off;
  p.dst.piece_stride = pl.d_piece_stride;
  p.total_items = static_cast<uint32_t>(total);
  p.n = n; p.regions = src->regions; p.ppc = pl.ppc; p.piece_bytes = pl.piece_bytes;
  p.s_runs = pl.s_runs; p.s_run_bytes = pl.s_run_bytes; p.s_stride = pl.s_stride;
  p.d_runs = pl.d_runs; p.d_run_bytes = pl.d_run_bytes; p.d_n1 = pl.d_n1;
  p.d_stride_a = pl.d_stride_a; p.d_stride_b = pl.d_stride_b;
  p.stages = kStaticStages; p.stage_bytes = kStaticStageBytes;
  p.flag = reinterpret_cast<unsigned long long*>(flag);
  p.flag_val = flag_val;
  p.done_counter = reinterpret_cast<unsigned int*>(sc);
  p.status = reinterpret_cast<unsigned int*>(sc + 4);

  const uint32_t grid = std::min<uint32_t>(static_cast<uint32_t>(dev->sms) * kStaticCtasPerSm, p.total_items);
  kvx::kvx_xfer_tma_kernel<16, 8><<<grid, 32u * (1u + kStaticConsumerWarps), kStaticStages * kStaticStageBytes, st>>>(p);
  if (bad(cudaGetLastError())) return KVX_ECUDA;
  return KVX_OK;
} KVX_CATCH_ALL

#include "kvx_queue_host.inl"
#include "kvx_vmm_host.inl"
