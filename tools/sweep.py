#!/usr/bin/env python
"""Tuning sweep of the transfer kernels on one B200 (run under gpurun; prints one JSON line per point).

  python tools/sweep.py [--isl 8192] [--model Qwen/Qwen3-0.6B] [--quick]
"""
import argparse
import itertools
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "dynamo-k8s-llm-inference_b200"))

import numpy as np
import torch

import kvx
from kvx import synth


def timed(stream, fn, reps):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(stream)
    for k in range(reps):
        fn(k)
    b.record(stream)
    b.synchronize()
    return a.elapsed_time(b) / reps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--isl", type=int, default=8192)
    ap.add_argument("--model", default="Qwen/Qwen3-0.6B")
    ap.add_argument("--reps", type=int, default=10)
    ap.add_argument("--quick", action="store_true")
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "sweep.jsonl"))
    args = ap.parse_args()
    shape = kvx.KV_SHAPES[args.model]
    n = shape.blocks_for(args.isl)
    sets = 4
    nb = sets * n + 1
    ctx = kvx.Context(0)
    P = synth.alloc_pool(ctx, shape, nb, kvx.HND)
    Dn = synth.alloc_pool(ctx, shape, nb, kvx.NHD)
    Dh = synth.alloc_pool(ctx, shape, nb, kvx.HND)
    synth.fill_random(P, 1)
    perm_p, perm_d = synth.make_table(nb, sets * n, 7), synth.make_table(nb, sets * n, 13)
    ts = [torch.from_numpy(np.ascontiguousarray(perm_p[s * n:(s + 1) * n])).cuda() for s in range(sets)]
    td = [torch.from_numpy(np.ascontiguousarray(perm_d[s * n:(s + 1) * n])).cuda() for s in range(sets)]
    payload = n * shape.regions * shape.chunk_bytes()
    stage = torch.empty(payload, dtype=torch.uint8, device="cuda")
    stage2 = torch.empty(payload, dtype=torch.uint8, device="cuda")
    stream = torch.cuda.current_stream()
    xf_n, xf_h = kvx.Xform(0, 0, 0, kvx.NHD), kvx.Xform(0, 0, 0, kvx.HND)

    ops = {
        "fused_H2N": lambda k: ctx.handoff_fused(P.pool, ts[k % sets], Dn.pool, td[k % sets], n, None, None, 0, stream),
        "fused_H2H": lambda k: ctx.handoff_fused(P.pool, ts[k % sets], Dh.pool, td[k % sets], n, None, None, 0, stream),
        "fused_N2H": lambda k: ctx.handoff_fused(Dn.pool, td[k % sets], Dh.pool, ts[k % sets], n, None, None, 0, stream),
        "gather_H2N": lambda k: ctx.gather(P.pool, ts[k % sets], n, stage, xf_n, stream),
        "gather_H2H": lambda k: ctx.gather(P.pool, ts[k % sets], n, stage, xf_h, stream),
        "scatter_N2N": lambda k: ctx.scatter(stage, Dn.pool, td[k % sets], n, xf_n, stream),
        "scatter_H2N": lambda k: ctx.scatter(stage, Dn.pool, td[k % sets], n, xf_h, stream),
        "copy": lambda k: ctx.copy(stage, stage2, payload, None, 0, stream),
    }
    grid = []
    for vec in (16, 32):
        for cps in (2, 4, 8):
            grid.append(dict(engine=1, vec_bytes=vec, ctas_per_sm=cps))
    tma = []
    for stages, sb, cps in ((4, 32768, 1), (6, 32768, 1), (3, 32768, 2), (4, 16384, 2), (8, 16384, 1), (6, 16384, 2), (4, 16384, 3),
                            (2, 65536, 1), (3, 65536, 1), (4, 8192, 4)):
        for store, vec, cw in ((2, 0, 4), (1, 16, 8), (1, 32, 8), (1, 32, 4), (1, 16, 16)):
            tma.append(dict(engine=2, stages=stages, stage_bytes=sb, ctas_per_sm=cps, store_mode=store, vec_bytes=vec, consumer_warps=cw))
    for c in tma:
        grid.append(c)
        if c["store_mode"] == 2:
            grid.append(dict(c, l2_hint=1))
    if args.quick:
        grid = grid[::5]
    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    out = open(args.out, "w")
    # yardstick
    for _ in range(3):
        stage2.copy_(stage)
    torch.cuda.synchronize()
    t = timed(stream, lambda k: stage2.copy_(stage), args.reps)
    rec = {"op": "torch_copy", "ms": t, "hbm_GBps": 2 * payload / t / 1e6}
    print(json.dumps(rec)); out.write(json.dumps(rec) + "\n")
    best = {}
    for cfg in grid:
        full = dict(engine=0, stages=4, stage_bytes=32768, ctas_per_sm=1, store_mode=0, vec_bytes=0, consumer_warps=8, l2_hint=0)
        full.update(cfg)
        try:
            for k, v in full.items():
                ctx.set_option(k, v)
        except kvx.KvxError:
            continue
        for name, fn in ops.items():
            try:
                for k in range(2):
                    fn(k)
                torch.cuda.synchronize()
                t = timed(stream, fn, args.reps)
            except kvx.KvxError as e:
                continue
            rec = {"op": name, "ms": round(t, 4), "hbm_GBps": round(2 * payload / t / 1e6, 1), **full}
            out.write(json.dumps(rec) + "\n")
            if name not in best or rec["hbm_GBps"] > best[name]["hbm_GBps"]:
                best[name] = rec
    out.close()
    ctx.check()
    print("=== best per op ===")
    for name, rec in best.items():
        print(json.dumps(rec))


if __name__ == "__main__":
    main()
