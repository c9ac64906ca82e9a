#!/usr/bin/env python
"""Launches for an ncu capture of the two store paths side by side (REPS launches each, in this order):
  Qwen3-0.6B, 8 heads, 32 KiB chunks, ISL 8192:  thread stores | tensor-map stores
  Llama-3.1-70B TP4 shard, 2 heads, 8 KiB chunks, ISL 8192:  thread stores | tensor-map stores (what store_mode 0 picks)
  Llama-3.1-70B TP8 shard, 1 head, 4 KiB chunks:  thread stores
ncu --set full -k regex:kvx_xfer -c 10 ... python tools/ncu_variants.py"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "dynamo-k8s-llm-inference_b200"))
import numpy as np
import torch

import kvx
from kvx import synth

REPS = int(os.environ.get("REPS", "2"))
stream = torch.cuda.current_stream()
out = []
for model, heads, modes in (("Qwen/Qwen3-0.6B", 8, (1, 3)), ("meta-llama/Llama-3.1-70B", 2, (1, 0)), ("meta-llama/Llama-3.1-70B", 1, (1,))):
    shape = kvx.KV_SHAPES[model]
    n = shape.blocks_for(8192)
    nb = 2 * n + 1
    for mode in modes:
        ctx = kvx.Context(0, store_mode=mode)
        P = synth.alloc_pool(ctx, shape, nb, kvx.HND, heads=heads)
        D = synth.alloc_pool(ctx, shape, nb, kvx.NHD, heads=heads)
        synth.fill_random(P, 1)
        ts = [torch.from_numpy(synth.make_table(nb, n, 1 + k)).cuda() for k in range(2)]
        td = [torch.from_numpy(synth.make_table(nb, n, 5 + k)).cuda() for k in range(2)]
        payload = n * shape.regions * shape.block_tokens * heads * shape.head_dim * 2
        torch.cuda.synchronize()
        evs = [torch.cuda.Event(enable_timing=True) for _ in range(REPS + 1)]
        evs[0].record(stream)
        for k in range(REPS):
            ctx.handoff_fused(P.pool, ts[k % 2], D.pool, td[k % 2], n, None, None, 0, stream)
            evs[k + 1].record(stream)
        torch.cuda.synchronize()
        ctx.check()
        out.append({"model": model, "heads": heads, "store_mode": mode, "algorithmic_bytes": 2 * payload,
                    "us": [round(1e3 * evs[k].elapsed_time(evs[k + 1]), 1) for k in range(REPS)]})
        P.free(); D.free(); ctx.close()
print(json.dumps(out))
