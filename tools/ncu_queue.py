#!/usr/bin/env python
"""A burst of 32 ISL-1k requests (Qwen3-0.6B shape, HND->NHD, 112 MiB each) through the resident receiver, for an ncu
capture of kvx_queue_kernel (one launch serves the whole burst, then idles `idle_us` and parks):
  ncu --replay-mode application --set full -k regex:kvx_queue -c 1 -o ... python tools/ncu_queue.py"""
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
m, R = shape.blocks_for(1024), 32
nb = R * m + 1
ctx = kvx.Context(0)
P = synth.alloc_pool(ctx, shape, nb, kvx.HND)
D = synth.alloc_pool(ctx, shape, nb, kvx.NHD)
synth.fill_random(P, 1)
tp, td = synth.make_table(nb, R * m, 1), synth.make_table(nb, R * m, 2)
torch.cuda.synchronize()
with ctx.queue(P.pool, D.pool, idle_us=100) as q:
    packed = q.pack([(np.ascontiguousarray(tp[j * m:(j + 1) * m]), np.ascontiguousarray(td[j * m:(j + 1) * m]), None, 1, None) for j in range(R)])
    t0 = time.perf_counter()
    q.wait(q.submit_many(packed), 20000)
    dt = time.perf_counter() - t0
    q.park()
    print(json.dumps({"requests": R, "bytes_moved_algorithmic": 2 * R * m * shape.regions * shape.chunk_bytes(), "wall_us": round(1e6 * dt, 1),
                      "launches": q.launches}))
ctx.check()
