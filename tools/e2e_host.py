#!/usr/bin/env python
"""Host-table entry (kvx_handoff_host) against the device-table entry, per request size: parameter-block
tables on/off, completion word in pinned or pageable memory (run under gpurun, one B200)."""
import os, sys, json, time
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "dynamo-k8s-llm-inference_b200"))
import numpy as np, torch, kvx
from kvx import synth
shape = kvx.KV_SHAPES["Qwen/Qwen3-0.6B"]
ctx = kvx.Context(0); stream = torch.cuda.current_stream(); flags,_ = ctx.flags()
nb = 4200
P = synth.alloc_pool(ctx, shape, nb, kvx.HND); D = synth.alloc_pool(ctx, shape, nb, kvx.NHD)
rng = np.random.Generator(np.random.PCG64(5))
def wall(fn, reps=50):
    for k in range(5): fn(k)
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(5):
        t0 = time.perf_counter()
        for k in range(reps): fn(k)
        torch.cuda.synchronize()
        best = min(best, (time.perf_counter()-t0)/reps)
    return best
for n in (8, 64, 256, 512, 1024, 2048):
    if n > nb-1: continue
    hs = [np.ascontiguousarray((1+rng.permutation(nb-1)[:n]).astype(np.int32)) for _ in range(4)]
    hd = [np.ascontiguousarray((1+rng.permutation(nb-1)[:n]).astype(np.int32)) for _ in range(4)]
    ps = [torch.from_numpy(t).pin_memory() for t in hs]; pd = [torch.from_numpy(t).pin_memory() for t in hd]
    ds = [t.cuda() for t in ps]; dd = [t.cuda() for t in pd]
    done = torch.zeros(64, dtype=torch.int64).pin_memory()
    done_pg = np.zeros(64, dtype=np.int64)
    cases = {
      "device_tables_flag": lambda k: ctx.handoff_fused(P.pool, ds[k%4], D.pool, dd[k%4], n, None, flags, k+1, stream),
      "host_pinned_done": lambda k: ctx.handoff_host(P.pool, hs[k%4], D.pool, hd[k%4], n, None, flags, k+1, done[k%64:k%64+1], stream),
      "host_pinned_done_noflag": lambda k: ctx.handoff_host(P.pool, hs[k%4], D.pool, hd[k%4], n, None, None, k+1, done[k%64:k%64+1], stream),
      "host_pageable_done": lambda k: ctx.handoff_host(P.pool, hs[k%4], D.pool, hd[k%4], n, None, flags, k+1, done_pg[k%64:k%64+1], stream),
      "host_no_done": lambda k: ctx.handoff_host(P.pool, hs[k%4], D.pool, hd[k%4], n, None, flags, k+1, None, stream),
    }
    for pt in (8000, 0):
        ctx.set_option("param_tables", pt)
        for name, fn in cases.items():
            if name.startswith("device") and pt == 0: continue
            t = wall(fn)
            print(json.dumps({"n": n, "param_tables": pt, "case": name, "us": round(1e6*t, 2)}), flush=True)
    assert int(done[(49) % 64]) > 0
ctx.check()
