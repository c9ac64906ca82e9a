#!/usr/bin/env python
"""A decode GPU does two things at once: it pulls the next request's KV over NVLink and it runs decode
steps.  Here the decode steps are stood in for by a bf16 GEMM loop (cuBLAS through torch) on a second
stream; the pull runs with the default lean geometry (few CTAs, deep pipelines) and with the full-GPU
geometry of an intra-device handoff.  Reports what each keeps of its stand-alone speed.  2 GPUs, 1 process."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "dynamo-k8s-llm-inference_b200"))

import torch

import kvx
from kvx import synth

shape = kvx.KV_SHAPES["meta-llama/Llama-3.1-8B"]
n = shape.blocks_for(8192)
nb = 2 * n + 1
pctx, dctx = kvx.Context(0), kvx.Context(1)
P = synth.alloc_pool(pctx, shape, nb, kvx.HND)
D = synth.alloc_pool(dctx, shape, nb, kvx.NHD)
synth.fill_random(P, 3)
torch.cuda.set_device(1)
src = dctx.import_pool(P.pool.export())
ts = [torch.from_numpy(synth.make_table(nb, n, 1 + k)).cuda(1) for k in range(2)]
td = [torch.from_numpy(synth.make_table(nb, n, 5 + k)).cuda(1) for k in range(2)]
payload = n * shape.regions * shape.chunk_bytes()
s_pull, s_gemm = torch.cuda.Stream(1), torch.cuda.Stream(1)
# ROWS=4096: a square, compute-bound GEMM (prefill-like, persistent one-CTA-per-SM kernel);
# ROWS=64: a skinny GEMM over a rotating set of weight matrices (decode-like: bound by reading weights from HBM)
ROWS, M = int(os.environ.get("ROWS", "64")), 4096
NW = 1 if ROWS >= 1024 else 64   # 64 x 32 MiB of weights: larger than L2
a = torch.randn(ROWS, M, device="cuda:1", dtype=torch.bfloat16)
b = [torch.randn(M, M, device="cuda:1", dtype=torch.bfloat16) for _ in range(NW)]
c = torch.empty(ROWS, M, device="cuda:1", dtype=torch.bfloat16)
PULLS = 12
GEMMS = 150 if ROWS >= 1024 else 2400


def pulls():
    for k in range(PULLS):
        dctx.handoff_fused(src, ts[k % 2], D.pool, td[k % 2], n, None, None, 0, s_pull)


def gemms():
    with torch.cuda.stream(s_gemm):
        for _ in range(GEMMS):
            torch.matmul(a, b[_ % NW], out=c)


def run(do_pull, do_gemm):
    torch.cuda.synchronize(1)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
    ev[0].record(s_pull); ev[2].record(s_gemm)
    if do_pull:   # the 12 pull launches go first: the GEMM loop is thousands of launches and would
        pulls()   # otherwise keep the host busy until most of it had already run alone
    if do_gemm:
        gemms()
    ev[1].record(s_pull); ev[3].record(s_gemm)
    torch.cuda.synchronize(1)
    return ev[0].elapsed_time(ev[1]), ev[2].elapsed_time(ev[3])


CONFIGS = (
    ("lean_default", {}),
    ("full_gpu", dict(stages=4, stage_bytes=16384, ctas_per_sm=2, max_ctas=296)),
    ("lean16", dict(stages=6, stage_bytes=32768, ctas_per_sm=1, max_ctas=16)),
    ("tiny_everywhere_2x8k", dict(stages=2, stage_bytes=8192, ctas_per_sm=1, max_ctas=148, consumer_warps=4)),
    ("tiny_everywhere_3x8k_x2", dict(stages=3, stage_bytes=8192, ctas_per_sm=2, max_ctas=296, consumer_warps=4)),
    ("small_4x8k_x1", dict(stages=4, stage_bytes=8192, ctas_per_sm=1, max_ctas=148, consumer_warps=4)),
)
if os.environ.get("CORUN_SET") == "max_ctas":   # the SM footprint of the pull as the caller bounds it (KVX_OPT_MAX_CTAS)
    CONFIGS = tuple((f"max_ctas_{k}", dict(stages=6, stage_bytes=32768, ctas_per_sm=1, max_ctas=k)) for k in (16, 24, 43, 148))
for name, opts in CONFIGS:
    for k in ("stages", "stage_bytes", "ctas_per_sm", "max_ctas"):
        dctx.set_option(k, opts.get(k, 0))
    dctx.set_option("consumer_warps", opts.get("consumer_warps", 16))
    for _ in range(2):
        run(True, True)
    t_pull, _ = run(True, False)
    _, t_gemm = run(False, True)
    tp, tg = run(True, True)
    flops = 2 * ROWS * M * M * GEMMS
    print(json.dumps({"pull_geometry": name,
                      "pull_alone_GBps": round(PULLS * payload / t_pull / 1e6, 1), "pull_corun_GBps": round(PULLS * payload / tp / 1e6, 1),
                      "gemm_alone_TFLOPs": round(flops / t_gemm / 1e9, 1), "gemm_corun_TFLOPs": round(flops / tg / 1e9, 1),
                      "gemm_alone_weight_GBps": round(GEMMS * M * M * 2 / t_gemm / 1e6, 1), "gemm_rows": ROWS,
                      "pull_keeps": round(t_pull / tp, 3), "gemm_keeps": round(t_gemm / tg, 3)}), flush=True)
dctx.check()
