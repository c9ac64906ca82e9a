#!/usr/bin/env python
"""The kernels of the handoff path, one after the other, for an ncu capture (config #2: Qwen3-0.6B KV
shape, ISL 8192, 896 MiB per request, scattered block tables).  Order of the kvx launches, each op
run REPS times back to back:
  gather HND->NHD image | copy image->image | scatter image->NHD pool | fused HND->NHD (device tables,
  completion flag) | fused through kvx_handoff_host (ids in the parameter block, pinned completion word)

  python tools/ncu_ops.py            # plain run, prints CUDA-event times
  ncu --set full -k regex:kvx_xfer -c 10 ... python tools/ncu_ops.py
"""
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
shape = kvx.KV_SHAPES[os.environ.get("MODEL", "Qwen/Qwen3-0.6B")]
n = shape.blocks_for(int(os.environ.get("ISL", "8192")))
sets = 2
nb = sets * n + 1
ctx = kvx.Context(0)
P = synth.alloc_pool(ctx, shape, nb, kvx.HND)
D = synth.alloc_pool(ctx, shape, nb, kvx.NHD)
synth.fill_random(P, 1)
synth.fill_canary(D)
perm_p, perm_d = synth.make_table(nb, sets * n, 7), synth.make_table(nb, sets * n, 13)
hs = [np.ascontiguousarray(perm_p[s * n:(s + 1) * n]) for s in range(sets)]
hd = [np.ascontiguousarray(perm_d[s * n:(s + 1) * n]) for s in range(sets)]
ts = [torch.from_numpy(t).cuda() for t in hs]
td = [torch.from_numpy(t).cuda() for t in hd]
payload = n * shape.regions * shape.chunk_bytes()
stage = torch.empty(payload, dtype=torch.uint8, device="cuda")
stage2 = torch.empty(payload, dtype=torch.uint8, device="cuda")
done = torch.zeros(8, dtype=torch.int64).pin_memory()
stream = torch.cuda.current_stream()
flags, _ = ctx.flags()
xf = kvx.Xform(0, 0, 0, kvx.NHD)
ops = [
    ("gather_hnd_to_nhd", lambda k: ctx.gather(P.pool, ts[k % sets], n, stage, xf, stream)),
    ("copy", lambda k: ctx.copy(stage, stage2, payload, None, 0, stream)),
    ("scatter_nhd", lambda k: ctx.scatter(stage2, D.pool, td[k % sets], n, xf, stream)),
    ("fused_hnd_to_nhd_flag", lambda k: ctx.handoff_fused(P.pool, ts[k % sets], D.pool, td[k % sets], n, None, flags, k + 1, stream)),
    ("fused_host_tables_pinned_done", lambda k: ctx.handoff_host(P.pool, hs[k % sets], D.pool, hd[k % sets], n, None, None, k + 1, done[k % 8:k % 8 + 1], stream)),
]
torch.cuda.synchronize()
out = {"algorithmic_bytes_per_launch": 2 * payload, "reps": REPS, "ops": {}}
for name, fn in ops:
    evs = [torch.cuda.Event(enable_timing=True) for _ in range(REPS + 1)]
    evs[0].record(stream)
    for k in range(REPS):
        fn(k)
        evs[k + 1].record(stream)
    torch.cuda.synchronize()
    out["ops"][name] = [round(evs[k].elapsed_time(evs[k + 1]) * 1e3, 2) for k in range(REPS)]
ctx.check()
print(json.dumps(out))
