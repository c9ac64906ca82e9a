"""Synthetic KV-handoff request stream over a manifest topology: one OS process per worker rank, one GPU
each (the manifests give every worker `gpu: "1"`, examples/deploy/vllm/disagg.yaml:23-25,43-45).

What `run_benchmark()` of the reference runner hands to `benchmarks.utils.benchmark`
(run-benchmarks.sh:56-68) — an endpoint, a model, an output directory — drives here the path the
disaggregated deployment exercises per request: the decode worker allocates fresh blocks, receives the
prefill worker's block ids (kv_transfer_params), pulls the blocks over NVLink with the layout transform
/ TP reshard in flight, and releases the prefill blocks.  Result names follow the metrics the
reference's dashboard and images export (grafana-dynamo-dashboard-configmap.yaml:214 TTFT;
vllm:nixl_xfer_time_seconds, nixl_bytes_transferred, nixl_num_descriptors).
"""
from __future__ import annotations

import json
import os
import socket
import statistics
import time
from dataclasses import asdict, dataclass, field
from typing import Dict, List, Optional

import numpy as np

from .manifest import Topology
from .shapes import KV_SHAPES, KVShape
from .tp import heads_per_rank, plan_edges

SEED = 0xD15A66
RESIDENCY = 4  # a request's blocks are freed this many transfers later (stands in for OSL-256 decode residency)


@dataclass
class StreamConfig:
    model: str
    isl: List[int]  # one entry per request of a pair's stream
    concurrency: int = 1
    verify: bool = True
    decode_layout: Optional[int] = None  # None = manifest's
    batch: int = 1  # > 1: the decode worker pulls this many queued requests per scheduler step in ONE launch
    #                  (kvx_handoff_batch; same-size TP only, route "pair")
    route: str = "pair"  # "pair": decode replica i reads prefill replica i % nP; "spread": request q of decode
    #                       replica i reads prefill replica (i + q) % nP (with concurrency >= nP this is an incast)
    options: Dict[str, int] = field(default_factory=dict)  # kvx context options


def draw_isls(requests: int, choices: List[int], seed: int) -> List[int]:
    """Log-uniform draw over the given ISL set (config #5), deterministic per pair."""
    rng = np.random.Generator(np.random.PCG64(seed))
    return [int(choices[i]) for i in rng.integers(0, len(choices), size=requests)]


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@dataclass
class RankPlan:
    rank: int
    role: str
    replica: int
    tp_rank: int
    tp_size: int
    layout: int
    device: int


def rank_plans(topo: Topology, n_devices: int) -> List[RankPlan]:
    """prefill ranks first (replica-major, tp-minor), then decode ranks; devices round-robin."""
    plans: List[RankPlan] = []
    for role in ("prefill", "decode"):
        for w in topo.role(role):
            for rep in range(w.replicas):
                for t in range(w.tp):
                    r = len(plans)
                    plans.append(RankPlan(r, role, rep, t, w.tp, w.layout, r % max(1, n_devices)))
    return plans


def pool_blocks(shape: KVShape, isl: List[int], concurrency: int) -> int:
    nmax = max(shape.blocks_for(x) for x in isl)
    return max(4096, (RESIDENCY + concurrency + 1) * nmax) + 1


def _pre_allocate(num_blocks: int, counts: List[int], seed: int, hold: int) -> List[np.ndarray]:
    """Block ids per request with realistic churn: request q's blocks return to the free list once
    request q+hold has been allocated."""
    from .workers import BlockAllocator

    alloc = BlockAllocator(num_blocks, seed)
    out = []
    for q, n in enumerate(counts):
        out.append(alloc.allocate(f"r{q}", n).copy())
        if q >= hold:
            alloc.free(f"r{q - hold}")
    return out


def _worker_main(rank: int, world: int, port: int, topo: Topology, cfg: StreamConfig, n_devices: int, out_dir: str):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    import torch
    import torch.distributed as dist

    from . import synth
    from .api import HND, NHD
    from .workers import DecodeWorker, PrefillWorker

    plans = rank_plans(topo, n_devices)
    me = plans[rank]
    torch.cuda.set_device(me.device)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    shape = KV_SHAPES[cfg.model]
    p_ranks = [p for p in plans if p.role == "prefill"]
    d_ranks = [p for p in plans if p.role == "decode"]
    p_tp, d_tp = p_ranks[0].tp_size, d_ranks[0].tp_size
    n_p_rep = 1 + max(p.replica for p in p_ranks)
    counts = [shape.blocks_for(x) for x in cfg.isl]
    nb = pool_blocks(shape, cfg.isl, cfg.concurrency)
    hold = RESIDENCY + cfg.concurrency
    edges = plan_edges(p_tp, d_tp, shape.kv_heads)
    layout = me.layout if (me.role == "prefill" or cfg.decode_layout is None) else cfg.decode_layout
    Worker = PrefillWorker if me.role == "prefill" else DecodeWorker
    w = Worker(me.device, shape, nb, layout, tp_rank=me.tp_rank, tp_size=me.tp_size, seed=SEED + rank, **cfg.options)
    if me.role == "prefill":
        synth.fill_random(w.kv, SEED + 1000 * me.replica + me.tp_rank)
    else:
        synth.fill_canary(w.kv)
    torch.cuda.synchronize()
    metas = [None] * world
    dist.all_gather_object(metas, w.metadata())

    # which prefill replica feeds request q of decode replica i
    def src_replica(d_rep: int, q: int = 0) -> int:
        return (d_rep + q) % n_p_rep if cfg.route == "spread" else d_rep % n_p_rep

    def p_rank_of(rep: int, t: int) -> int:
        return next(p.rank for p in p_ranks if p.replica == rep and p.tp_rank == t)

    result = None
    if me.role == "prefill":
        # block ids of every request of every stream this replica serves (same on all its TP ranks)
        served = [d for d in d_ranks if d.tp_rank == 0 and
                  any(src_replica(d.replica, q) == me.replica for q in range(len(counts)))]
        ids = {d.replica: _pre_allocate(nb, counts, SEED + 7 + 100 * me.replica + d.replica, hold) for d in served}
        box = [ids]
    else:
        box = [None]
    all_ids = [None] * world
    dist.all_gather_object(all_ids, box[0])

    if me.role == "decode":
        srcs = [e for e in edges if e.d_rank == me.tp_rank]
        p_reps = sorted({src_replica(me.replica, q) for q in range(len(counts))})
        for pr in p_reps:
            for e in srcs:
                w.add_remote(pr * 1000 + e.p_rank, metas[p_rank_of(pr, e.p_rank)])
        remote_ids = [all_ids[p_rank_of(src_replica(me.replica, q), 0)][me.replica][q] for q in range(len(counts))]
        local_ids = _pre_allocate(nb, counts, SEED + 13 + 100 * me.replica, hold)  # same on all TP ranks of a replica
        c = max(1, cfg.concurrency)
        streams = [[torch.cuda.Stream() for _ in srcs] for _ in range(c)]
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in cfg.isl]
        chunk_local = shape.block_tokens * w.kv.heads * shape.head_dim * shape.elem_bytes
        per_req, t_sub, t_done = [], [0.0] * len(counts), [0.0] * len(counts)
        dist.barrier()
        torch.cuda.synchronize()
        t_begin = time.perf_counter()
        batch = cfg.batch if (cfg.batch > 1 and len(srcs) == 1 and p_tp == d_tp and cfg.route == "pair") else 1
        group_of = {}
        if batch > 1:
            # one scheduler step = up to `batch` queued requests, received in one launch; closed loop over steps
            lane0 = streams[0][0]
            key = src_replica(me.replica, 0) * 1000 + srcs[0].p_rank
            for g0 in range(0, len(counts), batch):
                qs = list(range(g0, min(len(counts), g0 + batch)))
                t0 = time.perf_counter()
                ev[g0][0].record(lane0)
                w.pull_many(key, [(remote_ids[q], local_ids[q], (q * d_tp + me.tp_rank) % w.num_flags, q + 1, None) for q in qs],
                            stream=lane0)
                ev[g0][1].record(lane0)
                ev[g0][1].synchronize()
                t1 = time.perf_counter()
                for q in qs:
                    t_sub[q], t_done[q], group_of[q] = t0, t1, (g0, qs)
        for q, n in enumerate(counts):
            if batch > 1:
                break
            if q >= c:  # closed loop: at most c requests in flight
                ev[q - c][1].synchronize()
                t_done[q - c] = time.perf_counter()
            lane = streams[q % c]
            t_sub[q] = time.perf_counter()
            ev[q][0].record(lane[0])
            for j, e in enumerate(srcs):
                if j:
                    lane[j].wait_event(ev[q][0])
                # release slot: one flag per (request, consumer TP rank) in the prefill rank's memory
                w.pull(src_replica(me.replica, q) * 1000 + e.p_rank, remote_ids[q], local_ids[q], notify_slot=(q * d_tp + me.tp_rank) % w.num_flags,  # flag slot in the prefill rank's memory
                       value=q + 1, edge=e if (p_tp != d_tp) else None, host_tables=True, stream=lane[j])
            for j in range(1, len(srcs)):
                done_j = torch.cuda.Event()
                done_j.record(lane[j])
                lane[0].wait_event(done_j)
            ev[q][1].record(lane[0])
        for q in range(max(0, len(counts) - c), len(counts)):
            if batch > 1:
                break
            ev[q][1].synchronize()
            t_done[q] = time.perf_counter()
        torch.cuda.synchronize()
        t_end = time.perf_counter()
        w.ctx.check()
        for q, n in enumerate(counts):
            b = n * shape.regions * chunk_local
            if q in group_of:  # batched: the step's requests complete together; rate = the step's bytes over its time
                g0, qs = group_of[q]
                xfer = ev[g0][0].elapsed_time(ev[g0][1]) / 1e3
                rate = sum(counts[x] for x in qs) * shape.regions * chunk_local / xfer / 1e9
            else:
                xfer = ev[q][0].elapsed_time(ev[q][1]) / 1e3
                rate = b / xfer / 1e9
            per_req.append({"request": q, "isl": cfg.isl[q], "num_blocks": n, "bytes_transferred": b,
                            "num_descriptors": n * shape.regions * len(srcs), "xfer_time_seconds": xfer,
                            "throughput_GBps": rate, "batched_with": len(group_of[q][1]) if q in group_of else 1,
                            "time_to_first_token_added_ms": 1e3 * (t_done[q] - t_sub[q])})
        ok = True
        if cfg.verify:
            ok = _verify_tail(torch, w, shape, srcs, p_tp, d_tp, src_replica(me.replica, len(counts) - 1), remote_ids, local_ids, counts, nb,
                              me, HND, NHD)
        total_b = sum(r["bytes_transferred"] for r in per_req)
        result = {"rank": rank, "role": "decode", "replica": me.replica, "tp_rank": me.tp_rank, "device": me.device,
                  "sources": [e.p_rank for e in srcs], "wall_seconds": t_end - t_begin, "bytes_transferred": total_b,
                  "ingress_GBps": total_b / (t_end - t_begin) / 1e9, "verified": bool(ok), "requests": per_req}
    else:
        dist.barrier()
        # the producer side only learns of completion through release flags written by the consumers
        consumers = d_tp // p_tp if d_tp > p_tp else 1
        deadline = time.time() + 120
        pending = set()
        for dr in sorted(ids):  # the last request of every stream that routes to this replica
            mine = [q for q in range(len(counts)) if src_replica(dr, q) == me.replica]
            if mine:
                for t in range(d_tp):
                    if any(e.p_rank == me.tp_rank and e.d_rank == t for e in edges):
                        pending.add((mine[-1], t))
        while pending and time.time() < deadline:
            for key in list(pending):
                q_last, t = key
                if w.poll((q_last * d_tp + t) % w.num_flags, q_last + 1):
                    pending.discard(key)
            time.sleep(0.002)
        result = {"rank": rank, "role": "prefill", "replica": me.replica, "tp_rank": me.tp_rank, "device": me.device,
                  "released": not pending, "consumers_per_request": consumers}
    gathered = [None] * world
    dist.all_gather_object(gathered, result)
    dist.barrier()
    w.close()
    if rank == 0:
        with open(os.path.join(out_dir, "ranks.json"), "w") as f:
            json.dump(gathered, f)
    dist.destroy_process_group()


def _verify_tail(torch, w, shape, srcs, p_tp, d_tp, p_rep, remote_ids, local_ids, counts, nb, me, HND, NHD) -> bool:
    """Regenerate each source prefill rank's pool contents on this GPU (same seed, same generator) and
    compare the last request's destination blocks with the expected transform, with torch ops only."""
    q = len(counts) - 1
    n = counts[q]
    N, Dh = shape.block_tokens, shape.head_dim
    hd = w.kv.heads
    words_d = N * hd * Dh
    dv = w.kv.tensor().view(torch.int16).view(shape.regions, nb, words_d)
    li = torch.from_numpy(local_ids[q].astype(np.int64)).cuda()
    got = dv[:, li]
    got = got.view(shape.regions, n, N, hd, Dh) if w.layout == NHD else got.view(shape.regions, n, hd, N, Dh).transpose(2, 3)
    ok = True
    hp = heads_per_rank(shape.kv_heads, p_tp)
    for e in srcs:
        g = torch.Generator(device=f"cuda:{me.device}")
        g.manual_seed(SEED + 1000 * p_rep + e.p_rank)
        words_p = N * hp * Dh
        src = torch.empty(shape.regions * nb * words_p, dtype=torch.int16, device="cuda")
        step = 1 << 27
        for off in range(0, src.numel(), step):  # same chunking as synth.fill_random (1<<28 bytes)
            src[off:off + step].random_(-32768, 32768, generator=g)
        sv = src.view(shape.regions, nb, words_p)
        ri = torch.from_numpy(remote_ids[q][len(remote_ids[q]) - n:].astype(np.int64)).cuda()
        blk = sv[:, ri].view(shape.regions, n, hp, N, Dh).transpose(2, 3)  # prefill pools are HND here → [.., N, hp, D]
        if p_tp != d_tp:
            want = blk[:, :, :, e.src_head_off:e.src_head_off + e.heads]
            have = got[:, :, :, e.dst_head_off:e.dst_head_off + e.heads]
        else:
            want, have = blk, got
        ok = ok and bool(torch.equal(want, have))
        del src
    return ok


def run_stream(topo: Topology, cfg: StreamConfig, out_dir: str, n_devices: Optional[int] = None) -> Dict:
    """Spawn one process per worker rank, run the stream, collect per-request records and a summary."""
    import torch
    import torch.multiprocessing as mp

    if not torch.cuda.is_available():
        raise RuntimeError("the kvx harness needs CUDA devices: there is no CPU fallback")
    if not topo.disaggregated:
        raise ValueError("the manifest has no prefill/decode split: nothing is handed off in aggregated mode")
    if any(w.layout != 1 for w in topo.role("prefill")) and cfg.verify:
        cfg.verify = False  # the torch cross-check assumes HND prefill pools
    n_devices = n_devices or torch.cuda.device_count()
    plans = rank_plans(topo, n_devices)
    world = len(plans)
    os.makedirs(out_dir, exist_ok=True)
    port = _free_port()
    mp.spawn(_worker_main, args=(world, port, topo, cfg, n_devices, out_dir), nprocs=world, join=True)
    with open(os.path.join(out_dir, "ranks.json")) as f:
        ranks = json.load(f)
    dec = [r for r in ranks if r and r["role"] == "decode"]
    pre = [r for r in ranks if r and r["role"] == "prefill"]
    reqs = [q for r in dec for q in r["requests"]]
    ttft = sorted(q["time_to_first_token_added_ms"] for q in reqs)
    xfer = sorted(q["xfer_time_seconds"] for q in reqs)

    def pct(v, p):
        return v[min(len(v) - 1, int(p * len(v)))] if v else None

    wall = max(r["wall_seconds"] for r in dec)
    total = sum(r["bytes_transferred"] for r in dec)
    shape = KV_SHAPES[cfg.model]
    summary = {
        "topology": topo.describe(), "model": asdict(shape), "isl": cfg.isl, "concurrency": cfg.concurrency,
        "gpus_visible": n_devices, "ranks": world, "oversubscribed": world > n_devices,
        "requests": len(reqs), "bytes_transferred": total, "wall_seconds": wall,
        "aggregate_throughput_GBps": total / wall / 1e9,
        "per_decode_rank_ingress_GBps": {f"replica{r['replica']}/tp{r['tp_rank']}": r["ingress_GBps"] for r in dec},
        "per_edge_throughput_GBps_median": statistics.median(q["throughput_GBps"] for q in reqs) / max(1, len(dec[0]["sources"])),
        "per_request_throughput_GBps_median": statistics.median(q["throughput_GBps"] for q in reqs),
        "pct_of_nvlink_900": 100 * statistics.median(q["throughput_GBps"] for q in reqs) / 900.0,
        "xfer_time_seconds": {"p50": pct(xfer, 0.5), "p90": pct(xfer, 0.9), "max": xfer[-1]},
        "time_to_first_token_added_ms": {"p50": pct(ttft, 0.5), "p90": pct(ttft, 0.9), "max": ttft[-1]},
        "num_descriptors": sum(q["num_descriptors"] for q in reqs),
        "verified_bit_exact": all(r["verified"] for r in dec) if cfg.verify else None,
        "prefill_blocks_released": all(r["released"] for r in pre),
        "seed": SEED,
    }
    with open(os.path.join(out_dir, "results.json"), "w") as f:
        json.dump({"summary": summary, "decode_ranks": dec, "prefill_ranks": pre}, f, indent=1)
    return summary
