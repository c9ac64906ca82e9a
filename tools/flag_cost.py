#!/usr/bin/env python
"""What the completion flag costs per launch: no flag, a flag in this GPU's memory, a pinned host word.
KVX_LIB_PATH points at another build for A/B runs."""
import os, sys, json
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "dynamo-k8s-llm-inference_b200"))
import numpy as np, torch, kvx
from kvx import synth
shape = kvx.KV_SHAPES["Qwen/Qwen3-0.6B"]
ctx = kvx.Context(0); stream = torch.cuda.current_stream(); flags,_ = ctx.flags()
nb = 2049
P = synth.alloc_pool(ctx, shape, nb, kvx.HND); D = synth.alloc_pool(ctx, shape, nb, kvx.NHD)
rng = np.random.Generator(np.random.PCG64(5))
done = torch.zeros(8, dtype=torch.int64).pin_memory()
def timed(fn, reps=20):
    a,b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(stream)
    for k in range(reps): fn(k)
    b.record(stream); b.synchronize(); return a.elapsed_time(b)/reps
for n in (4, 64, 512):
    hs=[(1+rng.permutation(nb-1)[:n]).astype(np.int32) for _ in range(4)]; hd=[(1+rng.permutation(nb-1)[:n]).astype(np.int32) for _ in range(4)]
    tb = [(torch.from_numpy(a).cuda(), torch.from_numpy(b).cuda()) for a,b in zip(hs,hd)]
    cases = {"noflag": lambda k: ctx.handoff_fused(P.pool, tb[k%4][0], D.pool, tb[k%4][1], n, None, None, 0, stream),
             "local_flag": lambda k: ctx.handoff_fused(P.pool, tb[k%4][0], D.pool, tb[k%4][1], n, None, flags, k+1, stream),
             "host_word": lambda k: ctx.handoff_host(P.pool, hs[k%4], D.pool, hd[k%4], n, None, None, k+1, done[k%8:k%8+1], stream)}
    for name, fn in cases.items():
        for k in range(3): fn(k)
        torch.cuda.synchronize()
        t = min(timed(fn) for _ in range(5))
        print(json.dumps({"lib": os.environ.get("KVX_LIB_PATH","new")[-12:], "n": n, "case": name, "us": round(1e3*t,2)}))
