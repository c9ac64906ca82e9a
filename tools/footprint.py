#!/usr/bin/env python
"""How many CTAs (SMs) does a handoff need?  Sweeps KVX_OPT_MAX_CTAS for the NVLink pull (2 GPUs, one
process: cuda:1 pulls a Llama-3.1-8B request from cuda:0) and for the intra-device handoff."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "dynamo-k8s-llm-inference_b200"))

import numpy as np
import torch

import kvx
from kvx import synth

shape = kvx.KV_SHAPES["meta-llama/Llama-3.1-8B"]
n = shape.blocks_for(int(os.environ.get("ISL", "8192")))
nb = 2 * n + 1
two = torch.cuda.device_count() >= 2
pctx = kvx.Context(0)
dctx = kvx.Context(1 if two else 0)
dev = dctx.device
P = synth.alloc_pool(pctx, shape, nb, kvx.HND)
D = synth.alloc_pool(dctx, shape, nb, kvx.NHD)
synth.fill_random(P, 3)
torch.cuda.set_device(dev)
src = dctx.import_pool(P.pool.export()) if two else P.pool
contig = os.environ.get("TABLES", "random") == "contig"   # best-case control: ascending block ids on both sides
ts = [torch.from_numpy(np.arange(1 + k * n, 1 + (k + 1) * n, dtype=np.int32) if contig else synth.make_table(nb, n, 1 + k)).cuda(dev) for k in range(2)]
td = [torch.from_numpy(np.arange(1 + k * n, 1 + (k + 1) * n, dtype=np.int32) if contig else synth.make_table(nb, n, 5 + k)).cuda(dev) for k in range(2)]
stream = torch.cuda.current_stream(dev)
payload = n * shape.regions * shape.chunk_bytes()


def timed(reps=6):
    for k in range(2):
        dctx.handoff_fused(src, ts[k % 2], D.pool, td[k % 2], n, None, None, 0, stream)
    torch.cuda.synchronize(dev)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(stream)
    for k in range(reps):
        dctx.handoff_fused(src, ts[k % 2], D.pool, td[k % 2], n, None, None, 0, stream)
    b.record(stream)
    torch.cuda.synchronize(dev)
    return a.elapsed_time(b) / reps


dctx.set_option("max_ctas", 0)
t = timed()
print(json.dumps({"path": "nvlink_pull" if two else "intra_device", "defaults": True, "ms": round(t, 4), "payload_GBps": round(payload / t / 1e6, 1)}), flush=True)
if os.environ.get("DEFAULTS_ONLY"):
    sys.exit(0)
for stages, sb in ((4, 16384), (6, 32768)):
    dctx.set_option("stages", stages)
    dctx.set_option("stage_bytes", sb)
    dctx.set_option("ctas_per_sm", 2 if stages * sb <= 65536 else 1)
    for ctas in (4, 8, 12, 16, 24, 32, 48, 64, 96, 148, 296):
        dctx.set_option("max_ctas", ctas)
        t = timed()
        print(json.dumps({"path": "nvlink_pull" if two else "intra_device", "stages": stages, "stage_bytes": sb, "max_ctas": ctas,
                          "ms": round(t, 4), "payload_GBps": round(payload / t / 1e6, 1)}), flush=True)
dctx.check()
