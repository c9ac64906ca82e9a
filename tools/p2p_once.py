#!/usr/bin/env python
"""One process, two GPUs: the decode GPU (cuda:1) pulls a Llama-3.1-8B request (ISL 8192, 1 GiB) out of
the prefill GPU's pool (cuda:0) a few times.  Small enough to run under ncu for the NVLink capture."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "dynamo-k8s-llm-inference_b200"))

import numpy as np
import torch

import kvx
from kvx import synth
from kvx.workers import DecodeWorker, PrefillWorker

shape = kvx.KV_SHAPES["meta-llama/Llama-3.1-8B"]
n = shape.blocks_for(int(os.environ.get("ISL", "8192")))
nb = 2 * n + 1
p = PrefillWorker(0, shape, nb, kvx.HND)
d = DecodeWorker(1, shape, nb, kvx.NHD)
synth.fill_random(p.kv, 3)
synth.fill_canary(d.kv)
d.add_remote(0, p.metadata())
torch.cuda.set_device(1)
ts = torch.from_numpy(synth.make_table(nb, n, 1)).cuda(1)
td = torch.from_numpy(synth.make_table(nb, n, 2)).cuda(1)
stream = torch.cuda.current_stream(1)
payload = n * shape.regions * shape.chunk_bytes()
reps = int(os.environ.get("REPS", "6"))
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for k in range(reps):
    if k == reps - 3:
        a.record(stream)
    d.pull(0, ts, td, notify_slot=0, value=k + 1, host_tables=False, stream=stream)
b.record(stream)
torch.cuda.synchronize(1)
d.ctx.check()
ms = a.elapsed_time(b) / 3
print(f"pull: {ms:.4f} ms  {payload / ms / 1e6:.1f} GB/s over NVLink")
