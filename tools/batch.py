#!/usr/bin/env python
"""R short requests between the same pools: one kvx_handoff_host per request against one
kvx_handoff_batch for all of them (Qwen3-0.6B KV shape, one GPU; wall clock per scheduler step)."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "dynamo-k8s-llm-inference_b200"))

import numpy as np
import torch

import kvx
from kvx import synth

shape = kvx.KV_SHAPES["Qwen/Qwen3-0.6B"]
ctx = kvx.Context(0)
nb = 4097
P = synth.alloc_pool(ctx, shape, nb, kvx.HND)
D = synth.alloc_pool(ctx, shape, nb, kvx.NHD)
synth.fill_random(P, 1)
stream = torch.cuda.current_stream()
flags, _ = ctx.flags()
rng = np.random.Generator(np.random.PCG64(3))
for isl, R in ((128, 16), (256, 16), (512, 8), (1024, 8), (128, 64)):
    n = shape.blocks_for(isl)
    ids = [(np.ascontiguousarray((1 + rng.permutation(nb - 1)[:n]).astype(np.int32)),
            np.ascontiguousarray((1 + rng.permutation(nb - 1)[:n]).astype(np.int32))) for _ in range(R)]
    done = torch.zeros(R, dtype=torch.int64).pin_memory()
    reqs = [(s, d, flags + 8 * k, 1, done[k:k + 1]) for k, (s, d) in enumerate(ids)]

    def per_request():
        for k, (s, d) in enumerate(ids):
            ctx.handoff_host(P.pool, s, D.pool, d, n, None, flags + 8 * k, 1, done[k:k + 1], stream)

    def batched():
        ctx.handoff_batch(P.pool, D.pool, reqs, None, stream)

    out = {"isl": isl, "requests": R, "payload_MiB": round(R * n * shape.regions * shape.chunk_bytes() / 2**20, 1)}
    for name, fn in (("per_request", per_request), ("batched", batched)):
        for _ in range(5):
            fn()
        torch.cuda.synchronize()
        best = 1e9
        for _ in range(5):
            t0 = time.perf_counter()
            for _ in range(20):
                fn()
            torch.cuda.synchronize()
            best = min(best, (time.perf_counter() - t0) / 20)
        out[name + "_us"] = round(1e6 * best, 1)
    out["speedup"] = round(out["per_request_us"] / out["batched_us"], 2)
    print(json.dumps(out), flush=True)
ctx.check()
