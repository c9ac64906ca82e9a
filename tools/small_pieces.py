#!/usr/bin/env python
"""Intra-device handoff of small chunks (chunks of TP shards: 16 / 8 / 4 KiB) under the default launch
geometry; edit `grid` to sweep stages x CTAs per SM x consumer warps (this is the sweep behind the small-piece rule
in launch_xfer)."""
import os, sys, json
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "dynamo-k8s-llm-inference_b200"))
import numpy as np, torch, kvx
from kvx import synth
stream = torch.cuda.current_stream()
def timed(fn, reps=10):
    for k in range(3): fn(k)
    torch.cuda.synchronize()
    best=1e9
    for _ in range(3):
        a,b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        for k in range(reps): fn(k)
        b.record(stream); b.synchronize(); best=min(best,a.elapsed_time(b)/reps)
    return best
for model, heads, isl in (("meta-llama/Llama-3.1-70B", 2, 8192), ("meta-llama/Llama-3.1-70B", 1, 8192), ("meta-llama/Llama-3.1-70B", 4, 8192), ("meta-llama/Llama-3.1-70B", 8, 8192), ("meta-llama/Llama-3.2-1B-Instruct", 8, 2048), ("meta-llama/Llama-3.2-1B-Instruct", 4, 8192)):
    shape = kvx.KV_SHAPES[model]
    n = shape.blocks_for(isl); nb = 4*n+1
    grid = [{}]
    for opts in grid:
        ctx = kvx.Context(0, **opts)
        P = synth.alloc_pool(ctx, shape, nb, kvx.HND, heads=heads); D = synth.alloc_pool(ctx, shape, nb, kvx.NHD, heads=heads)
        tp = synth.make_table(nb, 4*n, 1); td = synth.make_table(nb, 4*n, 2)
        ts=[torch.from_numpy(np.ascontiguousarray(tp[j*n:(j+1)*n])).cuda() for j in range(4)]
        tdd=[torch.from_numpy(np.ascontiguousarray(td[j*n:(j+1)*n])).cuda() for j in range(4)]
        chunk = shape.block_tokens*heads*shape.head_dim*2
        payload = n*shape.regions*chunk
        try:
            t = timed(lambda k: ctx.handoff_fused(P.pool, ts[k%4], D.pool, tdd[k%4], n, None, None, 0, stream))
            print(json.dumps({"model": model, "heads": heads, "isl": isl, "chunk_KiB": chunk//1024, "payload_MiB": payload>>20, "opts": opts, "us": round(1e3*t,1), "hbm_GBps": round(2*payload/t/1e6,1)}), flush=True)
        except kvx.KvxError as e:
            print("skip", opts, e)
        P.free(); D.free(); ctx.close()
