#!/usr/bin/env python
"""All four layout pairs (HND/NHD on either side) between distinct pools, and the same-pool control, at ISL 8192:
one against two CTAs per SM."""
import os, sys, json
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "dynamo-k8s-llm-inference_b200"))
import numpy as np, torch, kvx
from kvx import synth
stream = torch.cuda.current_stream()
shape = kvx.KV_SHAPES["Qwen/Qwen3-0.6B"]
n=512; nb = 4*n+1
def timed(fn, reps=20):
    for k in range(3): fn(k)
    torch.cuda.synchronize()
    best=1e9
    for _ in range(3):
        a,b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        for k in range(reps): fn(k)
        b.record(stream); b.synchronize(); best=min(best,a.elapsed_time(b)/reps)
    return best
rng = np.random.Generator(np.random.PCG64(5))
for opts in ({}, dict(ctas_per_sm=2)):
    ctx = kvx.Context(0, **opts)
    pools = {"HND": synth.alloc_pool(ctx, shape, nb, kvx.HND), "HND2": synth.alloc_pool(ctx, shape, nb, kvx.HND), "NHD": synth.alloc_pool(ctx, shape, nb, kvx.NHD), "NHD2": synth.alloc_pool(ctx, shape, nb, kvx.NHD)}
    tabs=[(torch.from_numpy((1+rng.permutation(nb-1)[:n]).astype(np.int32)).cuda(), torch.from_numpy((1+rng.permutation(nb-1)[:n]).astype(np.int32)).cuda()) for _ in range(4)]
    payload = n*shape.regions*shape.chunk_bytes()
    for a,b in (("HND","HND2"),("HND","NHD"),("NHD","HND"),("NHD","NHD2"),("HND","HND")):
        t=timed(lambda k: ctx.handoff_fused(pools[a].pool, tabs[k%4][0], pools[b].pool, tabs[k%4][1], n, None, None, 0, stream))
        print(json.dumps({"opts":opts,"src":a,"dst":b,"us":round(1e3*t,1),"hbm_GBps":round(2*payload/t/1e6,1)}), flush=True)
    for p in pools.values(): p.free()
    ctx.close()
