#!/usr/bin/env python
"""Where the time of a short request goes, for the two ways of issuing it on one GPU (Qwen3-0.6B KV shape, HND->NHD):
the same BYTES moved as 1, 2, 4, 8, 16 requests — through the resident receiver (kvx_queue_*) and with one launch per
request (kvx_handoff_host) — host wall clock, best of 7.  The slope over the number of requests at fixed bytes is the
fixed cost of a request; the intercept is the bytes.  One JSON line per total size."""
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
ctx = kvx.Context(0)
nb = 4 * 512 + 1
P = synth.alloc_pool(ctx, shape, nb, kvx.HND)
D = synth.alloc_pool(ctx, shape, nb, kvx.NHD)
synth.fill_random(P, 1)
tp, td = synth.make_table(nb, 4 * 512, 1), synth.make_table(nb, 4 * 512, 2)
stream = torch.cuda.current_stream()


def best(fn, reps=7):
    fn(); fn()
    b = 1e9
    for _ in range(reps):
        t0 = time.perf_counter()
        fn()
        b = min(b, time.perf_counter() - t0)
    return 1e6 * b


for total_blocks in (128, 512):   # 224 MiB (ISL 2048 in all) and 896 MiB
    rec = {"total_blocks": total_blocks, "total_MiB": total_blocks * shape.regions * shape.chunk_bytes() >> 20, "queue_us": {}, "launch_us": {}}
    with ctx.queue(P.pool, D.pool) as q:
        for parts in (1, 2, 4, 8, 16):
            m = total_blocks // parts
            ids = [(np.ascontiguousarray(tp[j * m:(j + 1) * m]), np.ascontiguousarray(td[j * m:(j + 1) * m])) for j in range(parts)]
            packed = q.pack([(a, b, None, 1, None) for a, b in ids])
            rec["queue_us"][parts] = round(best(lambda: q.wait(q.submit_many(packed), 10000)), 1)
        q.park()
        for parts in (1, 2, 4, 8, 16):
            m = total_blocks // parts
            ids = [(np.ascontiguousarray(tp[j * m:(j + 1) * m]), np.ascontiguousarray(td[j * m:(j + 1) * m])) for j in range(parts)]

            def launches():
                for a, b in ids:
                    ctx.handoff_host(P.pool, a, D.pool, b, m, None, None, 1, None, stream)
                stream.synchronize()

            rec["launch_us"][parts] = round(best(launches), 1)
    for k in ("queue_us", "launch_us"):
        xs, ys = np.array(sorted(rec[k])), np.array([rec[k][p_] for p_ in sorted(rec[k])])
        slope, icpt = np.polyfit(xs, ys, 1)
        rec[k + "_fit"] = {"per_request_us": round(float(slope), 2), "bytes_and_once_only_us": round(float(icpt), 1)}
    print(json.dumps(rec), flush=True)
ctx.check()
