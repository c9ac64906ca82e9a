/*
 * kvx_c_client.c — a host program in plain C (gcc, no CUDA headers, no Python, no torch) that drives
 * libkvx.so through include/kvx.h only, the way a non-Python serving engine would bind it: register
 * two KV pools, hand a request over with host block-id lists (kvx_handoff_host), repeat it through
 * the allocation-free entry (kvx_handoff_static) and, in three parts, through the resident receiver
 * (kvx_queue_start / submit / wait / stop), and compare every byte of the three destination pools
 * with the oracle's C restatement.  TEST CODE: it links the oracle, the library does not.
 *
 *   gcc -I include kvx_c_client.c -L <dir of libkvx.so> -lkvx -ldl
 *   kvx_c_client <libkvx_oracle.so> <libcudart.so>
 * exit 0 = bit-exact; 3 = no CUDA device (kvx_ctx_create failed loudly, as it must without a GPU).
 */
#include <dlfcn.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "kvx.h"

typedef struct kvxo_side {  /* mirrors oracle/kvx_oracle.c */
  uint8_t* const* base;
  const int32_t* table;
  uint8_t* base0;
  uint64_t block_stride, num_blocks;
  uint32_t layout, heads, head_off, block_tokens;
} kvxo_side;

#define SYM(lib, type, name)                                   \
  type name = (type)dlsym(lib, #name);                         \
  if (!name) { fprintf(stderr, "missing symbol %s\n", #name); return 2; }
/* libkvx.so itself is LINKED (-lkvx): its entry points are the prototypes of include/kvx.h. */

typedef int (*cudaMemcpy_t)(void*, const void*, size_t, int);
typedef int (*cudaDeviceSynchronize_t)(void);
typedef int (*kvxo_xfer_t)(const kvxo_side*, const kvxo_side*, uint32_t, uint32_t, uint32_t, uint32_t, uint32_t, uint32_t);

static uint64_t rng_state = 0x9E3779B97F4A7C15ull;
static uint64_t rng(void) {
  rng_state ^= rng_state << 13; rng_state ^= rng_state >> 7; rng_state ^= rng_state << 17;
  return rng_state;
}

static void shuffled_ids(int32_t* out, uint32_t n, uint32_t num_blocks) {  /* unique, scattered, never block 0 */
  int32_t* all = (int32_t*)malloc(sizeof(int32_t) * (num_blocks - 1));
  for (uint32_t i = 0; i < num_blocks - 1; ++i) all[i] = (int32_t)(i + 1);
  for (uint32_t i = num_blocks - 2; i > 0; --i) {
    uint32_t j = (uint32_t)(rng() % (i + 1));
    int32_t t = all[i]; all[i] = all[j]; all[j] = t;
  }
  memcpy(out, all, sizeof(int32_t) * n);
  free(all);
}

int main(int argc, char** argv) {
  if (argc < 3) { fprintf(stderr, "usage: %s libkvx_oracle.so libcudart.so\n", argv[0]); return 2; }
  void* lo = dlopen(argv[1], RTLD_NOW);
  if (!lo) { fprintf(stderr, "dlopen: %s\n", dlerror()); return 2; }
  SYM(lo, kvxo_xfer_t, kvxo_xfer)
  if ((kvx_version() >> 16) != (KVX_ABI_VERSION >> 16)) { fprintf(stderr, "ABI mismatch\n"); return 2; }

  kvx_ctx* ctx = NULL;
  int32_t rc = kvx_ctx_create(0, &ctx);
  if (rc != KVX_OK) {  /* no device: the library must say so, not fall back */
    fprintf(stderr, "kvx_ctx_create: %s (%d)\n", kvx_strerror(rc), rc);
    return rc == KVX_ECUDA ? 3 : 2;
  }
  void* lc = dlopen(argv[2], RTLD_NOW);
  if (!lc) { fprintf(stderr, "dlopen cudart: %s\n", dlerror()); return 2; }
  SYM(lc, cudaMemcpy_t, cudaMemcpy)
  SYM(lc, cudaDeviceSynchronize_t, cudaDeviceSynchronize)

  /* Llama-3.1-8B chunk shape, 3 layers: prefill pool HND, decode pools NHD */
  enum { L = 3, H = 8, D = 128, E = 2, N = 16, REG = 2 * L, NB = 96, NREQ = 61 };
  const uint64_t chunk = (uint64_t)N * H * D * E, pool_bytes = (uint64_t)REG * NB * chunk;
  uint8_t* h_src = (uint8_t*)malloc(pool_bytes);
  uint8_t* h_want = (uint8_t*)malloc(pool_bytes);
  uint8_t* h_got = (uint8_t*)malloc(pool_bytes);
  for (uint64_t i = 0; i < pool_bytes; i += 8) { uint64_t v = rng(); memcpy(h_src + i, &v, 8); }
  for (uint64_t i = 0; i < pool_bytes; i += 2) { h_want[i] = 0xDE; h_want[i + 1] = 0xC0; }  /* canary 0xC0DE */

  void *d_src = NULL, *d_dst = NULL, *d_dst2 = NULL, *d_dst3 = NULL;
  if (kvx_mem_alloc(ctx, pool_bytes, &d_src) || kvx_mem_alloc(ctx, pool_bytes, &d_dst) || kvx_mem_alloc(ctx, pool_bytes, &d_dst2) ||
      kvx_mem_alloc(ctx, pool_bytes, &d_dst3))
    return 2;
  if (cudaMemcpy(d_src, h_src, pool_bytes, 1) || cudaMemcpy(d_dst, h_want, pool_bytes, 1) || cudaMemcpy(d_dst2, h_want, pool_bytes, 1) ||
      cudaMemcpy(d_dst3, h_want, pool_bytes, 1))
    return 2;

  void *sb[REG], *db[REG], *db2[REG], *db3[REG];
  uint8_t *hsb[REG], *hdb[REG];
  for (int r = 0; r < REG; ++r) {
    sb[r] = (uint8_t*)d_src + (uint64_t)r * NB * chunk;  db[r] = (uint8_t*)d_dst + (uint64_t)r * NB * chunk;
    db2[r] = (uint8_t*)d_dst2 + (uint64_t)r * NB * chunk;
    db3[r] = (uint8_t*)d_dst3 + (uint64_t)r * NB * chunk;
    hsb[r] = h_src + (uint64_t)r * NB * chunk;            hdb[r] = h_want + (uint64_t)r * NB * chunk;
  }
  kvx_pool_desc sd = {sizeof(kvx_pool_desc), REG, N, H, D, E, KVX_LAYOUT_HND, 0, 1, 0, NB, 0, sb};
  kvx_pool_desc dd = {sizeof(kvx_pool_desc), REG, N, H, D, E, KVX_LAYOUT_NHD, 0, 1, 0, NB, 0, db};
  kvx_pool_desc dd2 = dd, dd3 = dd;
  dd2.base = db2;
  dd3.base = db3;
  kvx_pool *sp = NULL, *dp = NULL, *dp3 = NULL;
  if ((rc = kvx_pool_register(ctx, &sd, &sp)) || (rc = kvx_pool_register(ctx, &dd, &dp)) || (rc = kvx_pool_register(ctx, &dd3, &dp3))) {
    fprintf(stderr, "register: %s\n", kvx_strerror(rc));
    return 2;
  }

  int32_t ts[NREQ], td[NREQ];
  shuffled_ids(ts, NREQ, NB);
  shuffled_ids(td, NREQ, NB);
  uint64_t done = 0;  /* pageable completion word */
  rc = kvx_handoff_host(ctx, sp, ts, dp, td, NREQ, NULL, NULL, 7, &done, NULL);
  if (rc) { fprintf(stderr, "kvx_handoff_host: %s\n", kvx_strerror(rc)); return 2; }
  rc = kvx_handoff_static(0, &sd, ts, &dd2, td, NREQ, NULL, NULL, 0, NULL);
  if (rc) { fprintf(stderr, "kvx_handoff_static: %s\n", kvx_strerror(rc)); return 2; }
  /* the same request in three parts through the resident receiver: no launch per part, tickets instead of streams */
  kvx_queue* q = NULL;
  if ((rc = kvx_queue_start(ctx, sp, dp3, NULL, 0, 0, &q))) { fprintf(stderr, "kvx_queue_start: %s\n", kvx_strerror(rc)); return 2; }
  uint64_t ticket = 0;
  const uint32_t parts[3] = {20, 20, NREQ - 40};
  for (uint32_t k = 0, at = 0; k < 3; at += parts[k], ++k)
    if ((rc = kvx_queue_submit(q, ts + at, td + at, parts[k], NULL, 0, NULL, &ticket))) { fprintf(stderr, "kvx_queue_submit: %s\n", kvx_strerror(rc)); return 2; }
  for (uint64_t t = ticket - 2; t <= ticket; ++t)
    if ((rc = kvx_queue_wait(q, t, 5000))) { fprintf(stderr, "kvx_queue_wait: %s\n", kvx_strerror(rc)); return 2; }
  if ((rc = kvx_queue_stop(q))) { fprintf(stderr, "kvx_queue_stop: %s\n", kvx_strerror(rc)); return 2; }
  if (cudaDeviceSynchronize() || kvx_ctx_check(ctx)) return 2;
  if (done != 7) { fprintf(stderr, "completion word %llu != 7\n", (unsigned long long)done); return 1; }

  kvxo_side os = {hsb, ts, NULL, chunk, NB, KVX_LAYOUT_HND, H, 0, N};
  kvxo_side od = {hdb, td, NULL, chunk, NB, KVX_LAYOUT_NHD, H, 0, N};
  if (kvxo_xfer(&os, &od, REG, NREQ, N, H, D * E, 2)) return 2;

  int bad = 0;
  if (cudaMemcpy(h_got, d_dst, pool_bytes, 2)) return 2;
  if (memcmp(h_got, h_want, pool_bytes)) { fprintf(stderr, "kvx_handoff_host: destination pool differs from the oracle\n"); bad = 1; }
  if (cudaMemcpy(h_got, d_dst2, pool_bytes, 2)) return 2;
  if (memcmp(h_got, h_want, pool_bytes)) { fprintf(stderr, "kvx_handoff_static: destination pool differs from the oracle\n"); bad = 1; }
  if (cudaMemcpy(h_got, d_dst3, pool_bytes, 2)) return 2;
  if (memcmp(h_got, h_want, pool_bytes)) { fprintf(stderr, "kvx_queue_*: destination pool differs from the oracle\n"); bad = 1; }
  if (cudaMemcpy(h_got, d_src, pool_bytes, 2)) return 2;
  if (memcmp(h_got, h_src, pool_bytes)) { fprintf(stderr, "source pool was modified\n"); bad = 1; }

  kvx_pool_release(sp); kvx_pool_release(dp); kvx_pool_release(dp3);
  kvx_mem_free(ctx, d_src); kvx_mem_free(ctx, d_dst); kvx_mem_free(ctx, d_dst2); kvx_mem_free(ctx, d_dst3);
  kvx_ctx_destroy(ctx);
  free(h_src); free(h_want); free(h_got);
  if (bad) return 1;
  printf("C client OK: %d blocks x %d regions x %llu B bit-exact through kvx_handoff_host, kvx_handoff_static and kvx_queue_*\n", NREQ, REG,
         (unsigned long long)chunk);
  return 0;
}
