#!/usr/bin/env python
"""NVLink sweep on 2 GPUs of one box (single process, peer access): fused P→D push and kvx_copy to
the peer for a grid of kernel options, next to cudaMemcpyPeerAsync (torch copy_) of the same bytes."""
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
from kvx.workers import DecodeWorker, PrefillWorker


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
    ap.add_argument("--reps", type=int, default=6)
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "sweep_p2p.jsonl"))
    args = ap.parse_args()
    shape = kvx.KV_SHAPES["meta-llama/Llama-3.1-8B"]
    n = shape.blocks_for(args.isl)
    sets = 2
    nb = sets * n + 1
    torch.cuda.set_device(0)
    p = PrefillWorker(0, shape, nb, kvx.HND)
    dn = DecodeWorker(1, shape, nb, kvx.NHD)
    dh = DecodeWorker(1, shape, nb, kvx.HND)
    synth.fill_random(p.kv, 1)
    p.add_remote(0, dn.metadata())
    p.add_remote(1, dh.metadata())
    perm_p, perm_d = synth.make_table(nb, sets * n, 7), synth.make_table(nb, sets * n, 13)
    ts = [torch.from_numpy(np.ascontiguousarray(perm_p[s * n:(s + 1) * n])).cuda(0) for s in range(sets)]
    td = [torch.from_numpy(np.ascontiguousarray(perm_d[s * n:(s + 1) * n])).cuda(0) for s in range(sets)]
    payload = n * shape.regions * shape.chunk_bytes()
    stage = torch.empty(payload, dtype=torch.uint8, device="cuda:0")
    peer_stage_ptr = dn.ctx.mem_alloc(payload)
    peer_stage = p.ctx.mem_import(dn.ctx.mem_export(peer_stage_ptr, payload))
    peer_t = synth.as_torch(peer_stage_ptr, payload, 1)
    stream = torch.cuda.current_stream(0)
    ctx = p.ctx
    # pull direction: the decode GPU runs the kernel and reads the prefill pool over NVLink
    torch.cuda.set_device(1)
    pull_src = dn.ctx.import_pool(p.pool.export())
    ts1 = [t.cuda(1) for t in ts]
    td1 = [t.cuda(1) for t in td]
    stream1 = torch.cuda.current_stream(1)
    local_stage1 = torch.empty(payload, dtype=torch.uint8, device="cuda:1")
    stage_on0 = dn.ctx.mem_import(p.ctx.mem_export(stage.data_ptr(), payload)) if False else None
    torch.cuda.set_device(0)
    pull_src_h = dh.ctx.import_pool(p.pool.export())
    pull_ops = {
        "pull_H2N": (dn.ctx, lambda k: dn.ctx.handoff_fused(pull_src, ts1[k % sets], dn.pool, td1[k % sets], n, None, None, 0, stream1)),
        "pull_H2H": (dh.ctx, lambda k: dh.ctx.handoff_fused(dh.ctx.import_pool(p.pool.export()) if False else pull_src_h, ts1[k % sets], dh.pool, td1[k % sets], n, None, None, 0, stream1)),
    }
    ops = {
        "push_H2N": lambda k: ctx.handoff_fused(p.pool, ts[k % sets], p.remotes[0].pool, td[k % sets], n, None, None, 0, stream),
        "push_H2H": lambda k: ctx.handoff_fused(p.pool, ts[k % sets], p.remotes[1].pool, td[k % sets], n, None, None, 0, stream),
        "copy_peer": lambda k: ctx.copy(stage, peer_stage, payload, None, 0, stream),
    }
    out = open(args.out, "w")
    for _ in range(2):
        peer_t.copy_(stage)
    torch.cuda.synchronize(0)
    t = timed(stream, lambda k: peer_t.copy_(stage), args.reps)
    rec = {"op": "cudaMemcpyPeer(torch copy_)", "ms": round(t, 4), "GBps": round(payload / t / 1e6, 1)}
    print(json.dumps(rec)); out.write(json.dumps(rec) + "\n")
    grid = []
    for vec in (16, 32):
        for cps in (2, 4, 8):
            grid.append(dict(engine=1, vec_bytes=vec, ctas_per_sm=cps))
    for stages, sb, cps in ((4, 16384, 2), (4, 32768, 1), (6, 32768, 1), (3, 65536, 1), (8, 16384, 1), (6, 16384, 2), (4, 8192, 4), (3, 32768, 2), (8, 8192, 2)):
        for store, vec, cw in ((2, 0, 4), (2, 0, 16), (1, 16, 16), (1, 32, 8), (1, 16, 8)):
            grid.append(dict(engine=2, stages=stages, stage_bytes=sb, ctas_per_sm=cps, store_mode=store, vec_bytes=vec, consumer_warps=cw))
    best = {}
    for cfg in grid:
        full = dict(engine=0, stages=4, stage_bytes=16384, ctas_per_sm=2, store_mode=0, vec_bytes=0, consumer_warps=16, l2_hint=0)
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
                torch.cuda.synchronize(0)
                t = timed(stream, fn, args.reps)
            except kvx.KvxError:
                continue
            rec = {"op": name, "ms": round(t, 4), "GBps": round(payload / t / 1e6, 1), **full}
            out.write(json.dumps(rec) + "\n")
            if name not in best or rec["GBps"] > best[name]["GBps"]:
                best[name] = rec
    # pull sweep (kernel on GPU 1)
    for cfg in grid:
        full = dict(engine=0, stages=4, stage_bytes=16384, ctas_per_sm=2, store_mode=0, vec_bytes=0, consumer_warps=16, l2_hint=0)
        full.update(cfg)
        for name, (c, fn) in pull_ops.items():
            try:
                for k, v in full.items():
                    c.set_option(k, v)
                torch.cuda.set_device(1)
                for k in range(2):
                    fn(k)
                torch.cuda.synchronize(1)
                t = timed(stream1, fn, args.reps)
            except kvx.KvxError as e:
                continue
            finally:
                torch.cuda.set_device(0)
            rec = {"op": name, "ms": round(t, 4), "GBps": round(payload / t / 1e6, 1), **full}
            out.write(json.dumps(rec) + "\n")
            if name not in best or rec["GBps"] > best[name]["GBps"]:
                best[name] = rec
    out.close()
    ctx.check()
    print("=== best per op ===")
    for rec in best.values():
        print(json.dumps(rec))


if __name__ == "__main__":
    main()
