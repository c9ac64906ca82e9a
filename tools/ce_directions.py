#!/usr/bin/env python
"""What the copy engines give for 1 GiB between two GPUs, by the GPU that issues the copy (one process, peer access):
issued on the DESTINATION GPU's stream (a pull) and on the SOURCE GPU's stream (a push).  The yardstick for the SM-driven
pull (775-787 GB/s) and push (705-716 GB/s) of the handoff kernel."""
import json

import torch

n = 1 << 30
a = torch.empty(n, dtype=torch.uint8, device="cuda:0")
b = torch.empty(n, dtype=torch.uint8, device="cuda:1")
out = {}
for name, dev in (("issued_on_destination_gpu_pull", 1), ("issued_on_source_gpu_push", 0)):
    torch.cuda.set_device(dev)
    s = torch.cuda.current_stream(dev)
    for _ in range(3):
        b.copy_(a, non_blocking=True)
    torch.cuda.synchronize(0); torch.cuda.synchronize(1)
    best = 1e9
    for _ in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(s)
        b.copy_(a, non_blocking=True)
        e1.record(s)
        e1.synchronize()
        best = min(best, e0.elapsed_time(e1))
    out[name] = {"ms": round(best, 4), "GBps": round(n / best / 1e6, 1)}
print(json.dumps(out))
