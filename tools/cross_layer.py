#!/usr/bin/env python
"""Cross-layer block pools ((nb, L, 2, N, H, D): every region of a block inside that block) on one GPU: items
walked block by block (default for such pools) against region by region (KVX_BLOCK_MAJOR=0), Qwen3-0.6B shape."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "dynamo-k8s-llm-inference_b200"))

import numpy as np
import torch

import kvx
from kvx import synth

shape = kvx.KV_SHAPES["Qwen/Qwen3-0.6B"]
ctx = kvx.Context(0)
stream = torch.cuda.current_stream()
chunk = shape.chunk_bytes()
R = shape.regions
for isl in (1024, 8192):
    n = shape.blocks_for(isl)
    nb = 4 * n + 1
    pools = []
    for _ in range(2):
        ptr = ctx.mem_alloc(nb * R * chunk)
        pools.append((ptr, ctx.register_pool([ptr + r * chunk for r in range(R)], num_blocks=nb, heads=shape.kv_heads,
                                             head_dim=shape.head_dim, layout=kvx.NHD, block_stride=R * chunk)))
    synth.as_torch(pools[0][0], nb * R * chunk, 0).view(torch.int16).random_(-32768, 32768)
    tp, td = synth.make_table(nb, 4 * n, 1), synth.make_table(nb, 4 * n, 2)
    tabs = [(torch.from_numpy(np.ascontiguousarray(tp[j * n:(j + 1) * n])).cuda(),
             torch.from_numpy(np.ascontiguousarray(td[j * n:(j + 1) * n])).cuda()) for j in range(4)]
    f = lambda k: ctx.handoff_fused(pools[0][1], tabs[k % 4][0], pools[1][1], tabs[k % 4][1], n, None, None, 0, stream)
    for k in range(3):
        f(k)
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(3):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        for k in range(20):
            f(k)
        b.record(stream)
        b.synchronize()
        best = min(best, a.elapsed_time(b) / 20)
    payload = n * R * chunk
    print(json.dumps({"isl": isl, "block_major": os.environ.get("KVX_BLOCK_MAJOR", "1"), "us": round(1e3 * best, 1),
                      "hbm_GBps": round(2 * payload / best / 1e6, 1)}), flush=True)
    for ptr, pool in pools:
        pool.release()
        ctx.mem_free(ptr)
ctx.check()
