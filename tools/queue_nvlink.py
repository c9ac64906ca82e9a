#!/usr/bin/env python
"""Short requests over NVLink through the resident receiver, one process, two GPUs: cuda:1 pulls ISL-1k Llama-3.1-8B
requests (128 MiB) from cuda:0; CTAs of the receiver x {one at a time, stream of 16}.  Host wall clock, best of 5."""
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

shape = kvx.KV_SHAPES["meta-llama/Llama-3.1-8B"]
ISL = int(os.environ.get("ISL", "1024"))
m = shape.blocks_for(ISL)
R = 16
nb = R * m + 1
pctx, dctx = kvx.Context(0), kvx.Context(1)
P = synth.alloc_pool(pctx, shape, nb, kvx.HND)
D = synth.alloc_pool(dctx, shape, nb, kvx.NHD)
synth.fill_random(P, 3)
torch.cuda.set_device(1)
src = dctx.import_pool(P.pool.export())
tp, td = synth.make_table(nb, R * m, 1), synth.make_table(nb, R * m, 2)
ids = [(np.ascontiguousarray(tp[j * m:(j + 1) * m]), np.ascontiguousarray(td[j * m:(j + 1) * m])) for j in range(R)]
b = m * shape.regions * shape.chunk_bytes()


def best(fn):
    fn(); fn()
    t = 1e9
    for _ in range(5):
        t0 = time.perf_counter()
        fn()
        t = min(t, time.perf_counter() - t0)
    return t / R


for ctas in (0, 32, 64, 96, 148):
    for extra in ({}, {"stages": 6, "stage_bytes": 32768}) if ctas else ({},):
        for k in ("stages", "stage_bytes"):
            dctx.set_option(k, extra.get(k, 0))
        with dctx.queue(src, D.pool, max_ctas=ctas) as q:
            packed = q.pack([(a, c, None, 1, None) for a, c in ids])
            t_stream = best(lambda: q.wait(q.submit_many(packed), 10000))

            def single():
                for a, c in ids:
                    q.wait(q.submit(a, c, m), 10000)

            t_single = best(single)
            print(json.dumps({"isl": ISL, "max_ctas": ctas or "default", "ctas": q.ctas, "opts": extra, "stream_us": round(1e6 * t_stream, 1),
                              "stream_GBps": round(b / t_stream / 1e9, 1), "single_us": round(1e6 * t_single, 1),
                              "single_GBps": round(b / t_single / 1e9, 1)}), flush=True)
            q.park()
dctx.check()
