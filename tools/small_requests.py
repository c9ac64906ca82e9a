#!/usr/bin/env python
"""Short requests (8 .. 2048 blocks, Qwen3-0.6B KV shape, flag on): launch geometry sweep behind the default of
one CTA per SM with 4 x 16 KiB stages."""
import os, sys, json
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "dynamo-k8s-llm-inference_b200"))
import numpy as np, torch, kvx
from kvx import synth
stream = torch.cuda.current_stream()
shape = kvx.KV_SHAPES["Qwen/Qwen3-0.6B"]
nb = 4097
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
grid=[{}]+[dict(stages=st, stage_bytes=sb, ctas_per_sm=cps, consumer_warps=cw) for sb in (16384,) for st in (3,4,6) for cps in (1,2,3) for cw in (4,8,16) if st*sb*cps <= 220*1024 and cw*32*128>=sb]
rng = np.random.Generator(np.random.PCG64(5))
for opts in grid:
    ctx = kvx.Context(0, **opts)
    P = synth.alloc_pool(ctx, shape, nb, kvx.HND); D = synth.alloc_pool(ctx, shape, nb, kvx.NHD)
    flags,_=ctx.flags()
    rec={"opts":opts}
    for n in (8, 64, 512, 2048):
        tabs=[(torch.from_numpy((1+rng.permutation(nb-1)[:n]).astype(np.int32)).cuda(), torch.from_numpy((1+rng.permutation(nb-1)[:n]).astype(np.int32)).cuda()) for _ in range(8)]
        try:
            t=timed(lambda k: ctx.handoff_fused(P.pool, tabs[k%8][0], D.pool, tabs[k%8][1], n, None, flags, k+1, stream))
            rec[f"n{n}_us"]=round(1e3*t,2)
        except kvx.KvxError: rec[f"n{n}_us"]=None
    print(json.dumps(rec), flush=True)
    P.free(); D.free(); ctx.close()
