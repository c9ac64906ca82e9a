/*
 * kvx.h — C ABI of libkvx.so: B200 (sm_100a) prefill→decode paged-KV handoff.
 *
 * The reference repository (emolinaro/dynamo-k8s-llm-inference) has no FFI for this path: the
 * handoff lives inside the `nvcr.io/nvidia/ai-dynamo/{vllm,sglang,tensorrtllm}-runtime:0.8.1` images that its manifests
 * select (examples/deploy/vllm/disagg.yaml:28,37,57; examples/deploy/sglang/disagg.yaml:45-50,80-85;
 * examples/deploy/trtllm/disagg.yaml:39-40,62-63).  Each entry point below therefore cites the
 * manifest line that switches the behaviour on, and the verb of the transfer agent the images use
 * for it, as seen in the on-box vLLM 0.22 NIXL connector ("[vllm]" =
 * vllm/distributed/kv_transfer/kv_connector/v1/nixl/, a layout authority, not the reference).
 *
 * Conventions
 *   - every function returns 0 (KVX_OK) or a negative KVX_E* code; nothing throws or exits;
 *   - the library is asynchronous with respect to the host and ordered on the caller's stream
 *     (`stream` is a cudaStream_t passed as void*; NULL = the legacy default stream);
 *   - the caller owns every KV pool, staging buffer and block table; the library owns contexts,
 *     pool records and its own small scratch;
 *   - block tables are DEVICE pointers to int32 block ids unless the name says `_host`;
 *   - a context is used by one host thread at a time; different contexts are independent;
 *   - there is no CPU fallback: without a CUDA device every compute entry returns KVX_ECUDA.
 */
#ifndef KVX_H_
#define KVX_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define KVX_ABI_VERSION 0x00010000u

enum {
  KVX_OK = 0,
  KVX_EINVAL = -1,       /* bad argument (null, misaligned, out of range)        */
  KVX_ECUDA = -2,        /* a CUDA call failed; text in kvx_last_error()         */
  KVX_ENOMEM = -3,       /* host or device allocation failed                     */
  KVX_EUNSUPPORTED = -4, /* geometry the kernels cannot express                  */
  KVX_ETIMEOUT = -5,     /* flag did not reach the value in time                 */
  KVX_ERANGE = -6,       /* caller buffer too small / size beyond a fixed limit  */
  KVX_EIPC = -7,         /* peer mapping failed                                  */
  KVX_ENOTREADY = -8     /* kvx_flag_poll: not there yet                         */
};

enum { KVX_LAYOUT_NHD = 0, KVX_LAYOUT_HND = 1 };

/* Tunables for kvx_ctx_set_option (also read once from the environment as KVX_<NAME>). */
enum {
  KVX_OPT_ENGINE = 0,      /* 0 auto, 1 "ldg" (register path), 2 "tma" (bulk-copy pipeline) */
  KVX_OPT_STAGES = 1,      /* tma: pipeline stages (1..8); 0 (default) = 4 for transfers that stay
                              on this GPU, 6..8 when one side is a peer's memory               */
  KVX_OPT_STAGE_BYTES = 2, /* tma: bytes per stage (multiple of 1024); 0 (default) = 16 KiB
                              on-GPU, 32 KiB over NVLink                                       */
  KVX_OPT_CTAS_PER_SM = 3, /* persistent grid = SMs x this; 0 (default) = 1 (3 for pieces of 8 KiB or
                              less, 2 for the register engine)                                 */
  KVX_OPT_STORE_MODE = 4,  /* tma: 0 auto, 1 thread stores, 2 bulk stores (one per destination run), 3 tensor-map
                              engine (cp.async.bulk.tensor: the TMA unit does the permutation, 2 warps per
                              CTA): one box STORE per head for HND chunks into an NHD pool, one box LOAD per
                              head for NHD chunks into an HND pool; same block size on both sides, device
                              tables — other transfers use thread stores.  0 picks it by itself for 8 KiB
                              chunks of two or more heads on one GPU, where it measured ~6 % faster        */
  KVX_OPT_VEC_BYTES = 5,   /* thread path: 16 or 32 byte accesses (0 auto)                   */
  KVX_OPT_CONSUMER_WARPS = 6, /* tma: consumer warps per CTA (1..16); 0 (default) = 16, or 4 when a
                                 piece is 8 KiB or less (chunks of TP shards)                    */
  KVX_OPT_L2_HINT = 7,     /* 1: evict_first on loads                                        */
  KVX_OPT_PDL = 8,         /* 1 (default): programmatic dependent launch — a transfer kernel may
                              become resident while the previous kernel on the stream drains; it
                              still waits for that kernel to complete before touching memory    */
  KVX_OPT_PARAM_TABLES = 9, /* kvx_handoff_host carries lists of up to this many ids (both sides
                               together; default and maximum 8000, 0 = never) in the kernel's
                               parameter block instead of uploading them first                   */
  KVX_OPT_MAX_CTAS = 10,    /* upper bound on the CTAs (hence SMs) one transfer occupies.  0 (default):
                               every SM for transfers that stay on this GPU (HBM-bound); for a
                               transfer over NVLink just enough CTAs to keep 8 MiB in flight
                               (43 of 148 SMs at 32 KiB chunks; 6 MiB saturate the link), so that
                               the rest of the GPU stays free for work running beside it         */
  KVX_OPT__COUNT = 11
};

typedef struct kvx_ctx kvx_ctx;
typedef struct kvx_pool kvx_pool;

/*
 * One worker's paged KV pool = what [vllm] register_kv_caches records (worker.py:788-1012) and
 * NixlAgentMetadata publishes (metadata.py:41-53): a base address per region (K and V of every
 * layer are separate regions), the number of blocks, and the byte size of one block of one region.
 * chunk_bytes = block_tokens * heads * head_dim * elem_bytes.  `block_stride` is the distance in
 * bytes between block id b and b+1 inside a region: chunk_bytes for split K/V pools
 * (FlashAttention (2, nb, N, H, D)), 2*chunk_bytes for blocks-first joint pools (FlashInfer
 * (nb, 2, N, H, D), where V's region base is K's base + chunk_bytes).  0 means chunk_bytes.
 */
typedef struct kvx_pool_desc {
  uint32_t struct_size;  /* sizeof(kvx_pool_desc), for ABI growth */
  uint32_t regions;      /* number of entries in base[]            */
  uint32_t block_tokens; /* N (manifests pin 16: sglang/disagg.yaml:39-40) */
  uint32_t heads;        /* KV heads held by THIS rank             */
  uint32_t head_dim;     /* D                                      */
  uint32_t elem_bytes;   /* E (2 = bf16)                           */
  uint32_t layout;       /* KVX_LAYOUT_NHD: [N][H][D]; KVX_LAYOUT_HND: [H][N][D] */
  uint32_t tp_rank;
  uint32_t tp_size;      /* sglang/disagg.yaml:41-42 (--tp)        */
  uint32_t head_stride;  /* HND only: bytes between consecutive heads of one chunk; 0 = heads back to back
                            (block_tokens * head_dim * elem_bytes).  Larger in cross-layer block pools
                            whose heads are the outermost dimension of a block, (nb, H, L, 2, N, D)
                            ([vllm] flash_attn.py:162-165): there head_stride = 2L * N * D * E, region
                            (l, kv) starts (2l + kv) * N * D * E into the block, block_stride is the block */
  uint64_t num_blocks;
  uint64_t block_stride;
  void* const* base;     /* host array [regions] of device addresses */
} kvx_pool_desc;

/*
 * In-flight transform.  Moves `heads` KV heads starting at local head `src_head_off` of the source
 * chunk to head slot `dst_head_off` of the destination chunk; heads == 0 means
 * min(src heads, dst heads) and offsets as given.  This is the TP reshard of [vllm]
 * tp_mapping.py:65-142 (rank_offset_factor → src_head_off, rank_to_attention_slot → dst_head_off).
 * `stage_layout` is the row order of the contiguous staging image that kvx_gather writes and
 * kvx_scatter reads ([regions][n][N*heads*D*E] bytes, heads compacted).  A NULL xform means: all
 * heads, no offsets, staging in the pool's own layout.
 */
typedef struct kvx_xform {
  uint32_t src_head_off;
  uint32_t dst_head_off;
  uint32_t heads;
  uint32_t stage_layout;
} kvx_xform;

/* ABI check / messages (static strings; do not free). */
uint32_t kvx_version(void);
const char* kvx_strerror(int32_t code);
/* Text of the last CUDA failure seen by this context ("" if none). */
const char* kvx_last_error(const kvx_ctx* ctx);

/* Per-worker context: device, scratch, completion flags.  Replaces the NIXL agent a worker
 * started with --is-prefill-worker / --is-decode-worker creates (vllm/disagg.yaml:37,57). */
int32_t kvx_ctx_create(int32_t device, kvx_ctx** out);
int32_t kvx_ctx_destroy(kvx_ctx* ctx);
int32_t kvx_ctx_set_option(kvx_ctx* ctx, int32_t option, int64_t value);
int32_t kvx_ctx_get_option(const kvx_ctx* ctx, int32_t option, int64_t* value);
/* Kernels launched by this context so far (bench.py's gpu_launches). */
int32_t kvx_ctx_launch_count(const kvx_ctx* ctx, uint64_t* count);
/* SM count of the context's device. */
int32_t kvx_ctx_sm_count(const kvx_ctx* ctx, int32_t* sms);

/* Raw device memory that can be exported to a peer process (cudaMalloc, not a caching allocator). */
int32_t kvx_mem_alloc(kvx_ctx* ctx, uint64_t bytes, void** out);
int32_t kvx_mem_free(kvx_ctx* ctx, void* ptr);
/* Serialise / map one device buffer across processes (handshake payload; [vllm]
 * get_agent_metadata / add_remote_agent, worker.py:991,1309).  *len in: capacity, out: used. */
int32_t kvx_mem_export(kvx_ctx* ctx, const void* ptr, uint64_t bytes, void* buf, size_t* len);
int32_t kvx_mem_import(kvx_ctx* ctx, const void* buf, size_t len, void** out);
int32_t kvx_mem_unmap(kvx_ctx* ctx, void* imported);

/* = register_kv_caches ([vllm] worker.py:788): records addresses, copies nothing. */
int32_t kvx_pool_register(kvx_ctx* ctx, const kvx_pool_desc* desc, kvx_pool** out);
int32_t kvx_pool_release(kvx_pool* pool);
int32_t kvx_pool_chunk_bytes(const kvx_pool* pool, uint64_t* bytes);
/* = agent metadata + add_remote_agent: descriptor + peer handles of every region. */
int32_t kvx_pool_export(kvx_pool* pool, void* buf, size_t* len);
int32_t kvx_pool_import(kvx_ctx* ctx, const void* buf, size_t len, kvx_pool** out);
/* Device address of one region (as registered, or as mapped by kvx_pool_import). */
int32_t kvx_pool_region_base(const kvx_pool* pool, uint32_t region, void** base);
/* A pool made of `count` consecutive regions of another (registered or imported) pool: the K and V of a group
 * of layers.  Layer-wise handoff — sending a layer group as soon as prefill has produced it, so that the
 * transfer overlaps the tail of prefill — is a handoff between two such slices; the block tables are the
 * request's own.  The slice maps nothing and owns nothing but its record: release it before its parent. */
int32_t kvx_pool_slice(const kvx_pool* pool, uint32_t first_region, uint32_t count, kvx_pool** out);
/* Can another PROCESS import this pool?  *across_processes = 1 when every region lies in a plain
 * cudaMalloc allocation (CUDA IPC handle available), 0 for memory without one (VMM / expandable
 * segments, stream-ordered pools): such a pool still exports to contexts of the same process, and
 * kvx_pool_import in another process fails with KVX_EIPC and the reason in kvx_last_error().  The
 * pods of the reference's manifests are separate processes (vllm/disagg.yaml:23-25,43-45), so a
 * worker should check this once after registering its pool. */
int32_t kvx_pool_exportable(kvx_pool* pool, int32_t* across_processes);

/*
 * Pools in memory that has no CUDA IPC handle: CUDA virtual memory management (cuMemCreate + cuMemMap), which is what an
 * engine running with expandable segments holds its KV cache in.  Such memory crosses process boundaries as POSIX file
 * descriptors (one per physical allocation under the pool), which the caller passes over a Unix-domain socket with
 * SCM_RIGHTS — the pods of the reference's manifests are separate processes (vllm/disagg.yaml:23-25,43-45).
 * kvx_pool_export_fds writes the record into buf and one fd per allocation into fds (*len / *count: capacity in, size
 * out; KVX_ERANGE with the needed sizes when too small); the caller sends them and closes its copies.  The memory must
 * have been created with CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR (torch's expandable segments are; KVX_EIPC otherwise).
 * kvx_pool_import_fds maps the allocations into this process — allocations that were contiguous in the exporter's
 * address space stay contiguous — with read/write access for the context's device, and returns a pool like
 * kvx_pool_import does; kvx_pool_release unmaps.  The fds can be closed after the import.
 * kvx_mem_alloc_vmm / kvx_mem_free_vmm allocate such memory (in pieces of piece_bytes, 0 = one piece): for engines
 * that want an exportable VMM pool, and for the tests.
 */
int32_t kvx_mem_alloc_vmm(kvx_ctx* ctx, uint64_t bytes, uint64_t piece_bytes, void** out);
int32_t kvx_mem_free_vmm(kvx_ctx* ctx, void* ptr);
int32_t kvx_pool_export_fds(kvx_pool* pool, void* buf, size_t* len, int32_t* fds, uint32_t* count);
int32_t kvx_pool_import_fds(kvx_ctx* ctx, const void* buf, size_t len, const int32_t* fds, uint32_t count, kvx_pool** out);

/* Completion flags: an array of uint64 in this context's device memory that a peer may write
 * (= NIXL notifications, [vllm] worker.py:2164,1792-1845).
 * Limits of flagged transfers (those given a flag, an h_done word or batch notifications): one
 * context serves up to 448 distinct streams at a time (beyond that a call waits for the device to
 * go idle once and starts over) and up to 128 flagged transfers recorded into CUDA graphs
 * (KVX_ERANGE afterwards); the backlog on each stream is unbounded. */
int32_t kvx_ctx_flags(kvx_ctx* ctx, uint64_t** flags, uint32_t* count);
int32_t kvx_flag_reset(kvx_ctx* ctx, uint64_t* flag, uint64_t value, void* stream);
/* Stream-ordered wait until *flag >= value (bounded: timeout_ms, then the context is marked failed
 * and the next kvx_flag_poll / kvx_ctx_check returns KVX_ETIMEOUT).  = check_xfer_state polling
 * ([vllm] worker.py:1868-1913) moved onto the GPU. */
int32_t kvx_flag_wait(kvx_ctx* ctx, const uint64_t* flag, uint64_t value, uint32_t timeout_ms,
                      void* stream);
/* Host-side check: KVX_OK if *flag >= value, else KVX_ENOTREADY. Synchronous 8-byte read. */
int32_t kvx_flag_poll(kvx_ctx* ctx, const uint64_t* flag, uint64_t value);
/* KVX_OK unless a stream-ordered wait timed out since the last call. */
int32_t kvx_ctx_check(kvx_ctx* ctx);

/* H1+H2(+H3 source slice): gather the request's n blocks of every region from the paged pool into
 * the contiguous staging image, transforming the layout on the way.  Replaces the descriptor list
 * of [vllm] worker.py:2202-2213 and the receive-side permute of kv_connector/utils.py:255-277. */
int32_t kvx_gather(kvx_ctx* ctx, const kvx_pool* src, const int32_t* d_table, uint32_t n,
                   void* staging, const kvx_xform* xf, void* stream);

/* H4+H6: contiguous copy into (peer) memory, then release-store flag_val to *flag (flag may be
 * NULL).  = make_prepped_xfer("READ") + transfer + notif ([vllm] worker.py:2220-2233). */
int32_t kvx_copy(kvx_ctx* ctx, const void* src, void* dst, uint64_t bytes, uint64_t* flag,
                 uint64_t flag_val, void* stream);

/* H5(+H3 destination slot): scatter a staging image into the pool's freshly allocated blocks. */
int32_t kvx_scatter(kvx_ctx* ctx, const void* staging, const kvx_pool* dst, const int32_t* d_table,
                    uint32_t n, const kvx_xform* xf, void* stream);

/* H1-H6 in one kernel: source blocks → shared memory → destination blocks (local or peer pool),
 * then the completion flag.  `flag` may be NULL.  Either pool may be a peer's (imported): run it on
 * the decode GPU with an imported source to PULL (reads cross NVLink), or on the prefill GPU with an
 * imported destination to PUSH.
 * Block-size regrouping: the two pools' block_tokens may differ by an integer ratio k (P 16-token
 * blocks into D 32/64-token blocks, [vllm] kv_connector/utils.py:224-252,280-299, or the reverse).
 * `n` then counts GRANULES = blocks of the smaller size; the other side's table holds ceil(n/k)
 * ids and granule i occupies tokens [(i % k)*N, +N) of its block table[i / k]. */
int32_t kvx_handoff_fused(kvx_ctx* ctx, const kvx_pool* src, const int32_t* d_src_table,
                          const kvx_pool* dst, const int32_t* d_dst_table, uint32_t n,
                          const kvx_xform* xf, uint64_t* flag, uint64_t flag_val, void* stream);

/* The call a serving engine makes per request ([vllm] kv_transfer_params, scheduler.py:664-674):
 * block-id lists arrive in HOST memory.  Lists of up to 8000 ids (both sides together, i.e. prompts
 * up to 64k tokens at 16-token blocks) travel in the kernel's parameter block — one launch, no copy
 * before it; longer ones are uploaded into context scratch first.  When `h_done` is not NULL it receives flag_val once every byte has landed: pinned
 * (cudaHostAlloc / cudaHostRegister) memory is written by the kernel itself, pageable memory by a
 * D2H copy queued behind it on the same stream.  The lists are consumed before the call returns,
 * whatever kind of host memory they are in (long lists are copied into a pinned ring of the
 * context first; a span of that ring is reused only after the kernel that read it has finished).
 * n is bounded by KVX_MAX_HOST_TABLE. */
#define KVX_MAX_HOST_TABLE (1u << 20)
int32_t kvx_handoff_host(kvx_ctx* ctx, const kvx_pool* src, const int32_t* h_src_table,
                         const kvx_pool* dst, const int32_t* h_dst_table, uint32_t n,
                         const kvx_xform* xf, uint64_t* flag, uint64_t flag_val, uint64_t* h_done,
                         void* stream);

/*
 * Several requests between the same two pools in ONE launch: what a decode worker does once per
 * scheduler step for every request it has to receive ([vllm] start_load_kv → _read_blocks_for_req for
 * each entry of reqs_to_recv, worker.py:1933-1968).  Short prompts are launch-bound (a handoff of 8
 * blocks costs ~15 us, of which ~9 are launch, pipeline fill and the completion flag), so batching
 * them pays.  Every item's `flag` (device or peer memory, may be NULL) and `h_done` (PINNED host
 * memory, may be NULL) receive that item's flag_val once every byte of the WHOLE batch has landed.
 * Items with n == 0 only notify.  Both pools must use the same block size.  When the ids and
 * notifications of the batch do not fit one kernel parameter block (8000 int32 entries: the ids of
 * both sides plus 4 per notification), or an h_done is pageable, the call degrades to one
 * kvx_handoff_host per item — same bytes, same notifications, more launches.
 */
typedef struct kvx_batch_item {
  const int32_t* h_src_table; /* host block ids on the source pool      */
  const int32_t* h_dst_table; /* host block ids on the destination pool */
  uint32_t n;                 /* blocks of this request                 */
  uint32_t reserved;
  uint64_t* flag;             /* completion flag in device / peer memory, or NULL */
  uint64_t flag_val;
  uint64_t* h_done;           /* completion word in pinned host memory, or NULL   */
} kvx_batch_item;
int32_t kvx_handoff_batch(kvx_ctx* ctx, const kvx_pool* src, const kvx_pool* dst, const kvx_batch_item* items,
                          uint32_t count, const kvx_xform* xf, void* stream);

/*
 * kvx_handoff_static — the fused handoff for callers that must not allocate.  It has no
 * counterpart in the reference or in the transfer agent its images use; it exists because some
 * users of this library cannot call cudaMalloc on the request path:
 *   - a serving engine that gave every free byte of HBM to its KV pool at start-up (the usual
 *     "GPU memory utilisation" budget), so a later cudaMalloc can fail or push it over budget;
 *   - a scheduler thread for which cudaMalloc/cudaFree are off limits because they synchronise the
 *     whole device and would stall decode steps running on other streams;
 *   - a process whose allocator has been frozen or is metered by a quota layer.
 * Such a caller has no kvx_ctx either (creating one allocates), so this entry takes the device
 * ordinal and the two pool descriptors directly.  Everything the kernel needs beyond the caller's
 * pools, tables and flag — device copies of the region base addresses and of both block tables,
 * the finished-CTA counter and a status word — lives in one fixed array inside the library
 * (KVX_STATIC_SCRATCH_BYTES of device memory, held by every GPU that loads libkvx.so until the
 * process exits).  The call performs no cudaMalloc, cudaFree or other allocating call.
 *
 * Limits: n <= KVX_STATIC_MAX_BLOCKS and src/dst regions <= KVX_STATIC_MAX_REGIONS, otherwise the
 * call returns KVX_ERANGE and launches nothing.  Tables are HOST pointers; `flag` (device memory,
 * may be NULL) receives flag_val when every byte has landed.
 *
 * Lifetime of the arguments: the block-id lists and both descriptors' base arrays are copied into
 * a buffer inside the library before anything is queued, the ids are checked against num_blocks on
 * that copy (KVX_ERANGE, nothing launched), and the upload reads that copy.  The caller may
 * therefore reuse or free them — pinned or pageable — as soon as the call returns, and the kernel
 * can never see an id that was not checked.
 *
 * One-time set-up: the first call on each device looks up the array's device address and prepares
 * the kernel; loading the kernel image and the array may take device memory once.  Make that first
 * call (n = 0 is enough) at start-up: before the KV pool claims the rest of HBM, and before
 * starting any stream capture on that device.  Later calls leave free device memory unchanged.
 *
 * Sharing: every call on a device uses the same array, which still holds the previous call's
 * tables; the call re-uploads what it reads and clears the counter and status word itself.  Two
 * calls that can overlap on the GPU (different streams, or different host threads) race on it:
 * issue them on one stream, or order them with events.
 */
#define KVX_STATIC_MAX_BLOCKS 4096u
#define KVX_STATIC_MAX_REGIONS 512u
#define KVX_STATIC_SCRATCH_BYTES (2u * KVX_STATIC_MAX_REGIONS * 8u + 2u * KVX_STATIC_MAX_BLOCKS * 4u + 256u)
int32_t kvx_handoff_static(int32_t device, const kvx_pool_desc* src, const int32_t* h_src_table,
                           const kvx_pool_desc* dst, const int32_t* h_dst_table, uint32_t n,
                           const kvx_xform* xf, uint64_t* flag, uint64_t flag_val, void* stream);

/*
 * kvx_queue_* — the resident receiver: the fused handoff of one P→D edge as a persistent kernel fed from a queue,
 * instead of one launch per request.  It has no counterpart in the transfer agent of the reference's images
 * (there a request is a descriptor list posted to the transport, [vllm] worker.py:2220-2230); it serves the same
 * call site — a decode worker receiving request after request from one prefill worker (vllm/disagg.yaml:18-37)
 * — when prompts are short and the launch, pipeline fill and completion round trip of a per-request kernel
 * dominate (ISL 1k: ~45 us per request against ~36 us of bytes on one GPU).
 *
 * kvx_queue_start binds a queue to (src pool, dst pool, xform): one edge, one geometry.  kvx_queue_submit hands
 * over a request — host block-id lists (copied and range-checked before the call returns), an optional flag in
 * device / peer memory and an optional completion word in PINNED host memory, both receiving flag_val once every
 * byte of THAT request has landed — and returns a ticket; it never launches anything while the kernel is
 * resident (cost: a few hundred nanoseconds plus the copy of the ids).  Requests are served in order; the CTAs
 * run from one request into the next without draining their pipelines.  kvx_queue_wait / kvx_queue_poll follow a
 * ticket from the host; to order GPU work behind a request use its `flag` with kvx_flag_wait on your stream.
 *
 * The kernel is resident, not immortal: after `idle_us` (0 = 1000) without a request it parks (exits), and the
 * next submit launches it again (one ordinary launch latency).  A resident kernel blocks cudaDeviceSynchronize,
 * cudaFree and everything else that waits for the whole device until it has parked; kvx_queue_park makes it leave
 * at once (after the submitted requests have landed).  The kernel occupies `max_ctas` CTAs (0 = the context's
 * geometry for this kind of transfer: every SM on one GPU, ~43 CTAs for a pull over NVLink), all of which must
 * be co-resident; it runs on a stream of its own, so it is NOT ordered with the caller's streams: the source
 * blocks must be complete when a request is submitted.  Both pools must use the same block size.  Up to 256
 * requests can be in flight (submit waits for a free slot, KVX_ETIMEOUT after 10 s); n <= KVX_QUEUE_MAX_BLOCKS.
 * A queue is used by one host thread at a time; kvx_ctx_destroy stops the context's queues.
 */
typedef struct kvx_queue kvx_queue;
#define KVX_QUEUE_MAX_BLOCKS (1u << 17)
int32_t kvx_queue_start(kvx_ctx* ctx, const kvx_pool* src, const kvx_pool* dst, const kvx_xform* xf,
                        uint32_t max_ctas, uint32_t idle_us, kvx_queue** out);
int32_t kvx_queue_submit(kvx_queue* q, const int32_t* h_src_table, const int32_t* h_dst_table, uint32_t n,
                         uint64_t* flag, uint64_t flag_val, uint64_t* h_done, uint64_t* ticket);
/* Every request of a scheduler step in one call (the items of kvx_handoff_batch: lists, flag, value, pinned word
 * per request).  Unlike kvx_handoff_batch each request is complete, and notified, on its own; *last_ticket is the
 * ticket of the last one (tickets are consecutive).  On an error the requests before the failing one stay queued. */
int32_t kvx_queue_submit_batch(kvx_queue* q, const kvx_batch_item* items, uint32_t count, uint64_t* last_ticket);
/* Host-side wait (bounded; KVX_ETIMEOUT) / check (KVX_ENOTREADY) for the request with this ticket. */
int32_t kvx_queue_wait(kvx_queue* q, uint64_t ticket, uint32_t timeout_ms);
int32_t kvx_queue_poll(kvx_queue* q, uint64_t ticket);
/* Wait for everything submitted, then make the kernel leave the GPU now (the next submit brings it back). */
int32_t kvx_queue_park(kvx_queue* q);
/* Kernel launches this queue has made so far (1 per resident period) and the CTAs it occupies. */
int32_t kvx_queue_launch_count(const kvx_queue* q, uint64_t* launches, uint32_t* ctas);
/* Drain, park and free. */
int32_t kvx_queue_stop(kvx_queue* q);

#ifdef __cplusplus
}
#endif
#endif /* KVX_H_ */
