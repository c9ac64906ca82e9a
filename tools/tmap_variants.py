#!/usr/bin/env python
"""Tensor-map engine (KVX_OPT_STORE_MODE=3: cp.async.bulk.tensor stores, 2 warps per CTA) against the default engine
(1-D bulk loads + 16 consumer warps of 16-byte stores) on the cases VERDICT r01 names: (i) fused HND->NHD at ISL 8k,
(ii) chunks of TP shards (8 KiB / 4 KiB), and with two GPUs (iii) the PUSH direction over NVLink.  One JSON line per
variant; `bit_exact` compares the whole destination pool of the variant with the default engine's."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "dynamo-k8s-llm-inference_b200"))
import numpy as np
import torch

import kvx
from kvx import synth

stream = torch.cuda.current_stream()


def timed(fn, reps=10):
    for k in range(3):
        fn(k)
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(3):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        for k in range(reps):
            fn(k)
        b.record(stream)
        b.synchronize()
        best = min(best, a.elapsed_time(b) / reps)
    return best


VARIANTS = [
    {},
    {"store_mode": 3},
    {"store_mode": 3, "ctas_per_sm": 2},
    {"store_mode": 3, "ctas_per_sm": 3},
    {"store_mode": 3, "stages": 8},
    {"store_mode": 3, "stages": 6, "stage_bytes": 32768},
    {"store_mode": 3, "stages": 3, "stage_bytes": 32768, "ctas_per_sm": 2},
    {"store_mode": 3, "stages": 8, "stage_bytes": 8192, "ctas_per_sm": 3},
]


def local_cases():
    cases = [("Qwen/Qwen3-0.6B", 8, 8192, kvx.HND, kvx.NHD), ("Qwen/Qwen3-0.6B", 8, 1024, kvx.HND, kvx.NHD),
             ("meta-llama/Llama-3.1-70B", 2, 8192, kvx.HND, kvx.NHD), ("meta-llama/Llama-3.1-70B", 1, 8192, kvx.HND, kvx.NHD),
             # the other direction: tensor-map LOADS (NHD source), plain bulk stores
             ("Qwen/Qwen3-0.6B", 8, 8192, kvx.NHD, kvx.HND), ("meta-llama/Llama-3.1-70B", 2, 8192, kvx.NHD, kvx.HND)]
    if os.environ.get("DIRECTION") == "nhd_to_hnd":
        cases = cases[4:]
    for model, heads, isl, lay_s, lay_d in cases:
        shape = kvx.KV_SHAPES[model]
        n = shape.blocks_for(isl)
        nb = 4 * n + 1
        ref = None
        for opts in VARIANTS:
            ctx = kvx.Context(0, **opts)
            P = synth.alloc_pool(ctx, shape, nb, lay_s, heads=heads)
            D = synth.alloc_pool(ctx, shape, nb, lay_d, heads=heads)
            synth.fill_random(P, 11)
            synth.fill_canary(D)
            tp, td = synth.make_table(nb, 4 * n, 1), synth.make_table(nb, 4 * n, 2)
            ts = [torch.from_numpy(np.ascontiguousarray(tp[j * n:(j + 1) * n])).cuda() for j in range(4)]
            tdd = [torch.from_numpy(np.ascontiguousarray(td[j * n:(j + 1) * n])).cuda() for j in range(4)]
            chunk = shape.block_tokens * heads * shape.head_dim * 2
            payload = n * shape.regions * chunk
            try:
                t = timed(lambda k: ctx.handoff_fused(P.pool, ts[k % 4], D.pool, tdd[k % 4], n, None, None, 0, stream))
                ctx.check()
                got = D.tensor().clone()
                if ref is None:
                    ref = got
                print(json.dumps({"case": "one GPU, fused " + ("HND->NHD" if lay_s == kvx.HND else "NHD->HND"), "model": model, "heads": heads, "isl": isl, "chunk_KiB": chunk // 1024,
                                  "opts": opts or "default (thread stores)", "us": round(1e3 * t, 1), "hbm_GBps": round(2 * payload / t / 1e6, 1),
                                  "bit_exact_vs_default": bool(torch.equal(got, ref))}), flush=True)
            except kvx.KvxError as e:
                print(json.dumps({"case": "one GPU", "opts": opts, "skipped": str(e)}), flush=True)
            P.free(); D.free(); ctx.close()
        del ref


def push_cases():
    """cuda:0 (prefill) stores a Llama-3.1-8B request into cuda:1's NHD pool."""
    from kvx.workers import DecodeWorker, PrefillWorker

    shape = kvx.KV_SHAPES["meta-llama/Llama-3.1-8B"]
    n = shape.blocks_for(8192)
    nb = 2 * n + 1
    payload = n * shape.regions * shape.chunk_bytes()
    ref = None
    for opts in [{}, {"store_mode": 3}, {"store_mode": 3, "ctas_per_sm": 3}, {"store_mode": 3, "stages": 6, "stage_bytes": 32768},
                 {"store_mode": 3, "stages": 6, "stage_bytes": 32768, "ctas_per_sm": 2}, {"store_mode": 3, "max_ctas": 64}]:
        p = PrefillWorker(0, shape, nb, kvx.HND, **opts)
        d = DecodeWorker(1, shape, nb, kvx.NHD)
        synth.fill_random(p.kv, 3)
        synth.fill_canary(d.kv)
        p.add_remote(0, d.metadata())
        torch.cuda.set_device(0)
        ts = torch.from_numpy(synth.make_table(nb, n, 1)).cuda(0)
        td = torch.from_numpy(synth.make_table(nb, n, 2)).cuda(0)
        s0 = torch.cuda.current_stream(0)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        for k in range(6):
            if k == 3:
                a.record(s0)
            p.push(0, ts, td, slot=0, value=k + 1, host_tables=False, stream=s0)
        b.record(s0)
        torch.cuda.synchronize(0)
        p.ctx.check()
        ms = a.elapsed_time(b) / 3
        got = d.kv.tensor().clone()
        if ref is None:
            ref = got
        print(json.dumps({"case": "two GPUs, PUSH over NVLink (prefill GPU stores)", "opts": opts or "default (thread stores)",
                          "ms": round(ms, 4), "GBps": round(payload / ms / 1e6, 1), "pct_of_900": round(100 * payload / ms / 1e6 / 900, 1),
                          "bit_exact_vs_default": bool(torch.equal(got, ref))}), flush=True)
        p.close(); d.close()


if __name__ == "__main__":
    if os.environ.get("CASES", "local") == "local":
        local_cases()
    if torch.cuda.device_count() >= 2 and os.environ.get("CASES", "push") in ("push", "all"):
        push_cases()
