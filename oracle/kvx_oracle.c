/*
 * kvx_oracle.c — plain-C restatement of the prefill→decode paged-KV handoff on host memory.
 * TEST INFRASTRUCTURE ONLY (checker for tests/, smoke(), and the timed CPU baseline of bench.py);
 * the product library never links or calls it.
 *
 * PARITY UNPINNED BY THE REFERENCE: /root/reference holds no code or golden vectors for this path
 * (SURVEY.md §8c).  The algorithm is the one BASELINE.json's north_star names, a memcpy-per-row
 * gather → layout transform / head slice → scatter; the meaning of block table, region, NHD/HND
 * and head slot follows vLLM 0.22's NIXL connector, the code path the reference's manifests select
 * (examples/deploy/vllm/disagg.yaml:37,57):
 *   address = region base + block id * block stride      (v1/nixl/worker.py:1108-1119)
 *   NHD [N][H][D] <-> HND [H][N][D] row permutation        (kv_connector/utils.py:255-277)
 *   head slice / slot for unequal TP                       (v1/nixl/worker.py:152-192,1159-1168)
 * It is checked against oracle/kvx_oracle.py and the fixtures in tests/golden/.
 *
 * Build: make -C oracle   →  oracle/_build/libkvx_oracle.so
 */
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

enum { KVXO_NHD = 0, KVXO_HND = 1 };

typedef struct kvxo_side {
  uint8_t* const* base;   /* paged: [regions] region base addresses; NULL = contiguous image */
  const int32_t* table;   /* paged: [n] block ids                                            */
  uint8_t* base0;         /* contiguous: image base, chunk (r, i) at (r*n + i) * chunk bytes  */
  uint64_t block_stride;  /* paged: bytes between block ids                                   */
  uint64_t num_blocks;    /* paged: ids must be < num_blocks                                  */
  uint32_t layout;        /* row order of this side's chunk                                   */
  uint32_t heads;         /* heads in this side's chunk                                       */
  uint32_t head_off;      /* first head of the moved slice                                    */
  uint32_t block_tokens;  /* tokens per block on THIS side (0 = the granule N)                */
} kvxo_side;

typedef struct kvxo_job {
  const kvxo_side* src;
  const kvxo_side* dst;
  uint32_t regions, n, N, Hc, row_bytes;
  uint32_t tid, threads;
  int rc;
} kvxo_job;

/* Block holding granule i.  With block_tokens = k*N the granule i is tokens [(i%k)*N, +N) of block
 * table[i/k] (block-size regrouping, vLLM kv_connector/utils.py:224-252,280-299); *sub gets i%k. */
static uint8_t* chunk_addr(const kvxo_side* s, uint32_t r, uint32_t i, uint32_t n, uint32_t N, uint32_t R, int* ok,
                           uint32_t* sub) {
  const uint32_t k = (s->block_tokens > N) ? s->block_tokens / N : 1;
  *sub = i % k;
  if (s->base) {
    int32_t id = s->table ? s->table[i / k] : (int32_t)(i / k);
    if (id < 0 || (uint64_t)id >= s->num_blocks) { *ok = 0; return NULL; }
    return s->base[r] + (uint64_t)id * s->block_stride;
  }
  return s->base0 + ((uint64_t)r * n + i) * ((uint64_t)N * s->heads * R);
}

static inline uint64_t row_off(const kvxo_side* s, uint32_t sub, uint32_t t, uint32_t h, uint32_t N, uint32_t R) {
  const uint32_t hh = s->head_off + h;
  const uint32_t nb = (s->block_tokens > N) ? s->block_tokens : N; /* tokens per block on this side */
  const uint32_t tt = sub * N + t;
  return s->layout == KVXO_NHD ? ((uint64_t)tt * s->heads + hh) * R : ((uint64_t)hh * nb + tt) * R;
}

static void* worker(void* arg) {
  kvxo_job* j = (kvxo_job*)arg;
  const uint64_t items = (uint64_t)j->regions * j->n;
  for (uint64_t w = j->tid; w < items; w += j->threads) {
    const uint32_t r = (uint32_t)(w / j->n), i = (uint32_t)(w % j->n);
    int ok = 1;
    uint32_t ssub = 0, dsub = 0;
    const uint8_t* sc = chunk_addr(j->src, r, i, j->n, j->N, j->row_bytes, &ok, &ssub);
    uint8_t* dc = chunk_addr(j->dst, r, i, j->n, j->N, j->row_bytes, &ok, &dsub);
    if (!ok) { j->rc = -6; continue; }
    for (uint32_t t = 0; t < j->N; ++t)
      for (uint32_t h = 0; h < j->Hc; ++h)
        memcpy(dc + row_off(j->dst, dsub, t, h, j->N, j->row_bytes),
               sc + row_off(j->src, ssub, t, h, j->N, j->row_bytes), j->row_bytes);
  }
  return NULL;
}

/* Move Hc heads of every (region, i) chunk from src to dst.  Returns 0, -1 (bad argument) or -6
 * (a block id out of range; that chunk is skipped). */
int kvxo_xfer(const kvxo_side* src, const kvxo_side* dst, uint32_t regions, uint32_t n, uint32_t N, uint32_t Hc,
              uint32_t row_bytes, uint32_t threads) {
  if (!src || !dst || N == 0 || Hc == 0 || row_bytes == 0) return -1;
  if (src->head_off + Hc > src->heads || dst->head_off + Hc > dst->heads) return -1;
  if (threads == 0) threads = 1;
  if (threads > 256) threads = 256;
  kvxo_job jobs[256];
  pthread_t th[256];
  for (uint32_t k = 0; k < threads; ++k) {
    jobs[k] = (kvxo_job){src, dst, regions, n, N, Hc, row_bytes, k, threads, 0};
  }
  for (uint32_t k = 1; k < threads; ++k) pthread_create(&th[k], NULL, worker, &jobs[k]);
  worker(&jobs[0]);
  int rc = jobs[0].rc;
  for (uint32_t k = 1; k < threads; ++k) {
    pthread_join(th[k], NULL);
    if (jobs[k].rc) rc = jobs[k].rc;
  }
  return rc;
}

uint32_t kvxo_version(void) { return 1; }
