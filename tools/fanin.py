#!/usr/bin/env python
"""TP4 -> TP2 fan-in on two GPUs (one process): the decode rank (cuda:1, 4 heads per chunk) pulls one
request from two prefill TP ranks' pools (2 heads per chunk; both placed on cuda:0, whose egress equals
the decode GPU's ingress, the binding link either way).  Compares one stream, two concurrent streams
and the single-edge full-chunk pull of the same bytes."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "dynamo-k8s-llm-inference_b200"))

import numpy as np
import torch

import kvx
from kvx import synth

shape = kvx.KV_SHAPES["meta-llama/Llama-3.1-70B"]
isl = int(os.environ.get("ISL", "8192"))
n = shape.blocks_for(isl)
nb = 2 * n + 1
d_layout = kvx.HND if os.environ.get("DLAYOUT", "hnd") == "hnd" else kvx.NHD
pctx = kvx.Context(0)
dctx = kvx.Context(1)
P = [synth.alloc_pool(pctx, shape, nb, kvx.HND, tp_rank=r, tp_size=4) for r in range(2)]
Pfull = synth.alloc_pool(pctx, shape, nb, kvx.HND, tp_rank=0, tp_size=2)
D = synth.alloc_pool(dctx, shape, nb, d_layout, tp_rank=0, tp_size=2)
for i, p in enumerate(P):
    synth.fill_random(p, 10 + i)
synth.fill_random(Pfull, 20)
torch.cuda.set_device(1)
rp = [dctx.import_pool(p.pool.export()) for p in P]
rfull = dctx.import_pool(Pfull.pool.export())
ts = [torch.from_numpy(synth.make_table(nb, n, 1 + k)).cuda(1) for k in range(2)]
td = [torch.from_numpy(synth.make_table(nb, n, 5 + k)).cuda(1) for k in range(2)]
s0, s1 = torch.cuda.Stream(1), torch.cuda.Stream(1)
payload = n * shape.regions * shape.chunk_bytes(2)  # bytes landing in the decode rank
xf = [kvx.Xform(0, 2 * r, 2, d_layout) for r in range(2)]


def timed(fn, reps=8):
    for k in range(2):
        fn(k)
    torch.cuda.synchronize(1)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(s0)
    for k in range(reps):
        fn(k)
    b.record(s0)
    torch.cuda.synchronize(1)
    return a.elapsed_time(b) / reps


def one_stream(k):
    for r in range(2):
        dctx.handoff_fused(rp[r], ts[k % 2], D.pool, td[k % 2], n, xf[r], None, 0, s0)


def two_streams(k):
    e = torch.cuda.Event()
    e.record(s0)
    s1.wait_event(e)
    dctx.handoff_fused(rp[0], ts[k % 2], D.pool, td[k % 2], n, xf[0], None, 0, s0)
    dctx.handoff_fused(rp[1], ts[k % 2], D.pool, td[k % 2], n, xf[1], None, 0, s1)
    e2 = torch.cuda.Event()
    e2.record(s1)
    s0.wait_event(e2)


def single_edge(k):
    dctx.handoff_fused(rfull, ts[k % 2], D.pool, td[k % 2], n, None, None, 0, s0)


for name, fn in (("single_edge_full_chunk", single_edge), ("fanin2_one_stream", one_stream), ("fanin2_two_streams", two_streams)):
    t = timed(fn)
    print(json.dumps({"case": name, "d_layout": "hnd" if d_layout == kvx.HND else "nhd", "ms": round(t, 4),
                      "ingress_GBps": round(payload / t / 1e6, 1), "opts": os.environ.get("KVX_STAGE_BYTES", "")}), flush=True)
dctx.check()
