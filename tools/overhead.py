#!/usr/bin/env python
"""Where the fused handoff loses time against a plain copy (run under gpurun, one B200).

  python tools/overhead.py [--out gpurun_out/overhead.jsonl]

Three questions, one JSON line per point:
  size     t(n) for n = 1 .. 2048 blocks: the intercept is the per-launch cost (launch gap, pipeline
           fill and drain, completion flag), the slope the streaming rate;
  order    the same 512-block request with block tables that are random, sorted on one side, sorted
           on both, or contiguous: how much of the gap to a plain copy is address locality;
  pool     the same request drawn from pools of 1x .. 16x its size: the reach of the address
           translation caches.
"""
import argparse
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


def dev(t):
    return torch.from_numpy(np.ascontiguousarray(t.astype(np.int32))).cuda()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="Qwen/Qwen3-0.6B")
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "overhead.jsonl"))
    ap.add_argument("--opts", default="", help="k=v,k=v context options")
    args = ap.parse_args()
    shape = kvx.KV_SHAPES[args.model]
    opts = {k: int(v) for k, v in (kv.split("=") for kv in args.opts.split(",") if kv)}
    ctx = kvx.Context(0, **opts)
    stream = torch.cuda.current_stream()
    flags, _ = ctx.flags()
    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    out = open(args.out, "a")

    def emit(rec):
        rec["opts"] = args.opts
        line = json.dumps(rec)
        print(line, flush=True)
        out.write(line + "\n")

    def bench(fn, reps=args.reps):
        for k in range(3):
            fn(k)
        torch.cuda.synchronize()
        return min(timed(stream, fn, reps) for _ in range(3))

    nmax = 2048
    sets = 4
    nb = sets * 512 + 1
    P = synth.alloc_pool(ctx, shape, nb, kvx.HND)
    D = synth.alloc_pool(ctx, shape, nb, kvx.NHD)
    synth.fill_random(P, 1)
    chunk = shape.chunk_bytes()
    rng = np.random.Generator(np.random.PCG64(5))

    def tables(n, kind, count=sets):
        res = []
        for s in range(count):
            a = 1 + rng.permutation(nb - 1)[:n]
            b = 1 + rng.permutation(nb - 1)[:n]
            if kind in ("sorted_src", "sorted_both"):
                a = np.sort(a)
            if kind in ("sorted_dst", "sorted_both"):
                b = np.sort(b)
            if kind == "contiguous":
                o1 = 1 + (s * n) % max(1, nb - 1 - n)
                a = np.arange(o1, o1 + n)
                b = np.arange(o1, o1 + n)
            if kind == "same_perm":   # both sides use the same scattered ids
                b = a.copy()
            res.append((dev(a), dev(b)))
        return res

    stage = torch.empty(nmax * shape.regions * chunk // 4, dtype=torch.uint8, device="cuda")
    stage2 = torch.empty_like(stage)

    # ---- size ----
    for n in (1, 4, 16, 64, 128, 256, 512, 1024, 2048):
        if n > nb - 1:
            continue
        payload = n * shape.regions * chunk
        tb = tables(n, "random")
        tc = tables(n, "contiguous")
        cases = {
            "fused_H2N_random": lambda k: ctx.handoff_fused(P.pool, tb[k % sets][0], D.pool, tb[k % sets][1], n, None, None, 0, stream),
            "fused_H2N_random_flag": lambda k: ctx.handoff_fused(P.pool, tb[k % sets][0], D.pool, tb[k % sets][1], n, None, flags, k + 1, stream),
            "fused_H2N_contig": lambda k: ctx.handoff_fused(P.pool, tc[k % sets][0], D.pool, tc[k % sets][1], n, None, None, 0, stream),
        }
        if payload <= stage.numel():
            cases["kvx_copy"] = lambda k: ctx.copy(stage, stage2, payload, None, 0, stream)
            cases["torch_copy"] = lambda k: stage2[:payload].copy_(stage[:payload])
        for name, fn in cases.items():
            t = bench(fn)
            emit({"exp": "size", "case": name, "n": n, "payload_MiB": payload / 2**20, "us": round(1e3 * t, 2),
                  "hbm_GBps": round(2 * payload / t / 1e6, 1)})

    # ---- order ----
    n = 512
    payload = n * shape.regions * chunk
    for kind in ("random", "sorted_src", "sorted_dst", "sorted_both", "same_perm", "contiguous"):
        tb = tables(n, kind)
        for lay, Dst in (("H2N", D),):
            t = bench(lambda k: ctx.handoff_fused(P.pool, tb[k % sets][0], Dst.pool, tb[k % sets][1], n, None, None, 0, stream))
            emit({"exp": "order", "case": kind, "layout": lay, "n": n, "us": round(1e3 * t, 2), "hbm_GBps": round(2 * payload / t / 1e6, 1)})
    P.free(); D.free()

    # ---- pool size ----
    for mult in (1, 2, 4, 8, 16):
        nbm = mult * n + 1
        Pm = synth.alloc_pool(ctx, shape, nbm, kvx.HND)
        Dm = synth.alloc_pool(ctx, shape, nbm, kvx.NHD)
        tb = [(dev(1 + rng.permutation(nbm - 1)[:n]), dev(1 + rng.permutation(nbm - 1)[:n])) for _ in range(sets)]
        t = bench(lambda k: ctx.handoff_fused(Pm.pool, tb[k % sets][0], Dm.pool, tb[k % sets][1], n, None, None, 0, stream))
        emit({"exp": "pool", "pool_blocks": nbm, "pool_GiB": round(Pm.nbytes / 2**30, 2), "n": n, "us": round(1e3 * t, 2),
              "hbm_GBps": round(2 * payload / t / 1e6, 1)})
        Pm.free(); Dm.free()
    ctx.check()
    ctx.close()


if __name__ == "__main__":
    main()
