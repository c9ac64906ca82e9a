"""Synthetic KV-handoff request stream over a manifest topology: one OS process per worker rank, one GPU
each (the manifests give every worker `gpu: "1"`, examples/deploy/vllm/disagg.yaml:23-25,43-45).

What `run_benchmark()` of the reference runner hands to `benchmarks.utils.benchmark`
(run-benchmarks.sh:56-68) — an endpoint, a model, an output directory — drives here the path the
disaggregated deployment exercises per request: the decode worker allocates fresh blocks, receives the
prefill worker's block ids (kv_transfer_params), pulls the blocks over NVLink with the layout transform
/ TP reshard in flight, and releases the prefill blocks.  Result names follow the metrics the
reference's dashboard and images export (grafana-dynamo-dashboard-configmap.yaml:214 TTFT;
vllm:nixl_xfer_time_seconds, nixl_bytes_transferred, nixl_num_descriptors).

Every request is verified: when it has landed, its destination blocks are compared on the device with the
expected transform of the source blocks (source contents regenerated from the prefill rank's seed on the decode
GPU — not read back through the mapping that was just used — unless that does not fit in memory); at the end
of the stream every block of the decode pool is checked once more: blocks a request still owns against their
source, every other block against the canary.  A run without verification reports `verified_bit_exact: null`
and the runner treats that as "not verified", never as success.

`run_rank` is the per-process body; `run_stream` spawns the processes itself, `bench.py` calls `run_rank`
from the ranks torchrun started.
"""
from __future__ import annotations

import json
import os
import socket
import sys
import statistics
import time
from dataclasses import asdict, dataclass, field
from typing import Dict, List, Optional, Tuple

import numpy as np

from .manifest import Topology
from .shapes import KV_SHAPES, KVShape
from .tp import heads_per_rank, plan_edges

SEED = 0xD15A66
RESIDENCY = 4  # a request's blocks are freed this many transfers later (stands in for OSL-256 decode residency)


@dataclass
class StreamConfig:
    model: str
    isl: List[int]  # closed loop: one entry per request of EVERY decode replica's stream; with arrivals: one global list
    concurrency: int = 1
    verify: bool = True
    decode_layout: Optional[int] = None  # None = manifest's
    batch: int = 1  # > 1: the decode worker pulls this many queued requests per scheduler step in ONE launch
    #                  (kvx_handoff_batch; same-size TP only, route "pair")
    route: str = "pair"  # "pair": decode replica i reads prefill replica i % nP; "spread": request q of decode
    #                       replica i reads prefill replica (i + q) % nP (with concurrency >= nP this is an incast);
    #                       "sched": kvx/router.py books each request on the ingress/egress ports that are free first
    options: Dict[str, int] = field(default_factory=dict)  # kvx context options
    per_replica_isl: Optional[List[List[int]]] = None  # closed loop: a list per decode replica (config #5: a seed per pair)
    ready_ms: Optional[List[float]] = None  # arrival-driven: when each request's KV is ready (prefill done), from stream start
    layerwise: int = 1        # > 1: hand each request over in this many layer groups, each as soon as it is "prefilled"
    prefill_ms: float = 0.0   # layerwise: synthetic prefill time of one request on the prefill GPU (spread over its layers)
    queue: bool = False       # pull through the resident receiver (kvx_queue_*: no launch per request); same-size TP, not layer-wise
    corrupt: Optional[Tuple[int, int]] = None  # test hook: flip one byte of request (decode replica, index) after it landed


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


# ----------------------------------------------------------------------------------------------------------
# Request plan: who reads what from whom, identical on every rank (pure function of the configuration)
# ----------------------------------------------------------------------------------------------------------
@dataclass
class Req:
    d_rep: int
    idx: int                 # position in decode replica d_rep's pull order
    isl: int
    n: int                   # blocks
    p_rep: int
    p_serial: int = 0        # position among the requests prefill replica p_rep serves (its allocation order)
    ready_ms: Optional[float] = None  # None = closed loop (pull as soon as a slot frees)
    q: int = 0               # global index (arrival-driven streams), else == idx


def plan_requests(cfg: StreamConfig, shape: KVShape, n_p: int, n_d: int, payload_of) -> Dict[int, List[Req]]:
    """Every decode replica's ordered list of requests.  `payload_of(n_blocks)` = bytes one request moves."""
    from . import router

    per_d: Dict[int, List[Req]] = {d: [] for d in range(n_d)}
    if cfg.ready_ms is not None:
        if len(cfg.ready_ms) != len(cfg.isl):
            raise ValueError("ready_ms needs one entry per request")
        counts = [shape.blocks_for(x) for x in cfg.isl]
        plan = router.plan(cfg.route, cfg.ready_ms, [payload_of(n) for n in counts], n_p, n_d)
        for a in sorted(plan, key=lambda a: (a.d, a.order)):
            per_d[a.d].append(Req(a.d, a.order, cfg.isl[a.q], counts[a.q], a.p, ready_ms=a.ready_ms, q=a.q))
    else:
        if cfg.route == "sched":
            raise ValueError("route=sched books requests by the time their KV is ready: give ready_ms (arrival=...)")
        for d in range(n_d):
            isls = cfg.per_replica_isl[d] if cfg.per_replica_isl else cfg.isl
            for q, isl in enumerate(isls):
                p = (d + q) % n_p if cfg.route == "spread" else d % n_p
                per_d[d].append(Req(d, q, isl, shape.blocks_for(isl), p, q=q))
    # serial number of each request on its prefill replica: order by (ready time, decode replica, index)
    serial = [0] * n_p
    for r in sorted((r for lst in per_d.values() for r in lst), key=lambda r: (r.ready_ms or 0.0, r.idx, r.d_rep)):
        r.p_serial = serial[r.p_rep]
        serial[r.p_rep] += 1
    return per_d


def _served_by(per_d: Dict[int, List[Req]], p_rep: int) -> List[Req]:
    return sorted((r for lst in per_d.values() for r in lst if r.p_rep == p_rep), key=lambda r: r.p_serial)


def _release_slot(p_serial: int, consumer: int, n_consumers: int, num_flags: int) -> int:
    """Release flag in the prefill rank's memory: one per (request it serves, consuming decode TP rank), in the
    lower half of its flags (the upper half announces layer groups)."""
    return (p_serial * n_consumers + consumer) % (num_flags // 2)


def regen_random(torch, nbytes: int, seed: int, device: int):
    """The bytes synth.fill_random(pool, seed) wrote, regenerated on `device` (same generator, same chunking)."""
    g = torch.Generator(device=f"cuda:{device}")
    g.manual_seed(seed)
    t = torch.empty(nbytes // 2, dtype=torch.int16, device=f"cuda:{device}")
    step = 1 << 27  # int16 elements = 1 << 28 bytes, as in synth.fill_random
    for off in range(0, t.numel(), step):
        t[off:off + step].random_(-32768, 32768, generator=g)
    return t


class _Verifier:
    """Device-side bit-exact checks of a decode rank, on a stream of its own so that they never sit between two
    transfers of the stream being timed."""

    REGION_STEP = 8  # regions compared per torch op (bounds the temporaries at ISL 32k to 0.5 GiB)

    def __init__(self, torch, w, shape, nb, me, sources, p_tp, d_tp, hp, p_layouts):
        self.torch, self.w, self.shape, self.nb, self.me = torch, w, shape, nb, me
        self.p_tp, self.d_tp, self.hp = p_tp, d_tp, hp
        self.stream = torch.cuda.Stream()
        self.N, self.Dh = shape.block_tokens, shape.head_dim
        self.hd = w.kv.heads
        self.dv = w.kv.tensor().view(torch.int16).view(shape.regions, nb, self.N * self.hd * self.Dh)
        self.src: Dict[int, object] = {}     # remote key → [regions, nb, words_p] int16 view of that prefill rank's pool
        self.src_layout: Dict[int, int] = {}
        self.mode: Dict[int, str] = {}
        words_p = self.N * hp * self.Dh
        need = shape.regions * nb * words_p * 2
        for key, (p_rep, edge) in sources.items():
            free, _ = torch.cuda.mem_get_info()
            if need < 0.5 * free:
                t = regen_random(torch, need, SEED + 1000 * p_rep + edge.p_rank, me.device)
                self.mode[key] = "regenerated from the prefill rank's seed"
            else:  # does not fit beside the pool: read the prefill rank's bytes through the peer mapping
                from . import synth

                t = synth.as_torch(w.remotes[key].pool.region_base(0), need, me.device).view(torch.int16)
                self.mode[key] = "read through the peer mapping (regeneration would not fit)"
            self.src[key] = t.view(shape.regions, nb, words_p)
            self.src_layout[key] = p_layouts[key]
        self.ok = torch.ones(1, dtype=torch.int32, device=f"cuda:{me.device}")
        self.stream.wait_stream(torch.cuda.current_stream())  # the regenerated source was written on the current stream
        self.per_req: List[object] = []
        self.canary = np.array(0xC0DE, dtype=np.uint16).view(np.int16).item()

    def _canon(self, blocks, heads, layout):
        """[regions', n, words] → [regions', n, N, heads, D] whatever the row order."""
        from .api import NHD

        r, n = blocks.shape[0], blocks.shape[1]
        if layout == NHD:
            return blocks.view(r, n, self.N, heads, self.Dh)
        return blocks.view(r, n, heads, self.N, self.Dh).transpose(2, 3)

    def _compare(self, remote: Dict[int, np.ndarray], local_ids: np.ndarray, edges: Dict[int, object], mask=None):
        """Device scalar: do the destination blocks `local_ids` (optionally only those selected by the boolean
        `mask`) hold the transform of the request's source blocks?"""
        torch = self.torch
        good = torch.ones((), dtype=torch.bool, device=self.ok.device)
        if len(local_ids) == 0 or (mask is not None and not mask.any()):
            return good
        sel = np.nonzero(mask)[0] if mask is not None else np.arange(len(local_ids))
        li = torch.from_numpy(local_ids[sel].astype(np.int64)).to(self.ok.device)
        for key, edge in edges.items():
            rid = remote[key]
            tail = rid[len(rid) - len(local_ids):]  # prefix hit on the decode side: the LAST len(local) source blocks
            ri = torch.from_numpy(tail[sel].astype(np.int64)).to(self.ok.device)
            for r0 in range(0, self.shape.regions, self.REGION_STEP):
                sl = slice(r0, min(self.shape.regions, r0 + self.REGION_STEP))
                got = self._canon(self.dv[sl][:, li], self.hd, self.w.layout)
                want = self._canon(self.src[key][sl][:, ri], self.hp, self.src_layout[key])
                if self.p_tp != self.d_tp:
                    want = want[:, :, :, edge.src_head_off:edge.src_head_off + edge.heads]
                    got = got[:, :, :, edge.dst_head_off:edge.dst_head_off + edge.heads]
                good = good & (got == want).all()
        return good

    def check_request(self, after, remote: Dict[int, np.ndarray], local_ids, edges, corrupt: bool = False):
        """Queue the check of one landed request behind event `after`; returns the event that marks the check done."""
        torch = self.torch
        if after is not None:   # None: the host has already seen the request complete (resident receiver)
            self.stream.wait_event(after)
        with torch.cuda.stream(self.stream):
            if corrupt and len(local_ids):
                self.dv[self.shape.regions // 2, int(local_ids[len(local_ids) // 2]), 5] ^= 1
            good = self._compare(remote, np.asarray(local_ids), edges)
            self.per_req.append(good.to(torch.int32).reshape(1))
            done = torch.cuda.Event()
            done.record(self.stream)
        return done

    def finish(self, requests: List[Tuple[Dict[int, np.ndarray], np.ndarray, Dict[int, object]]]):
        """End of stream: every block of the decode pool once more.  Walk the requests newest first; the blocks a
        request still owns (no later request reused them) must hold its bytes, all blocks no request ever touched
        must hold the canary (block 0 included)."""
        torch = self.torch
        with torch.cuda.stream(self.stream):
            owned = np.zeros(self.nb, dtype=bool)
            final_ok = torch.ones((), dtype=torch.bool, device=self.ok.device)
            blocks_final = 0
            for remote, local_ids, edges in reversed(requests):
                local_ids = np.asarray(local_ids)
                mine = ~owned[local_ids]
                if mine.any():
                    final_ok = final_ok & self._compare(remote, local_ids, edges, mask=mine)
                    blocks_final += int(mine.sum())
                owned[local_ids] = True
            untouched = torch.from_numpy(np.nonzero(~owned)[0].astype(np.int64)).to(self.ok.device)
            canary_ok = torch.ones((), dtype=torch.bool, device=self.ok.device)
            for r0 in range(0, self.shape.regions, self.REGION_STEP):
                sl = slice(r0, min(self.shape.regions, r0 + self.REGION_STEP))
                canary_ok = canary_ok & (self.dv[sl][:, untouched] == self.canary).all()
            per = torch.cat(self.per_req) if self.per_req else torch.ones(0, dtype=torch.int32, device=self.ok.device)
            self.stream.synchronize()
            per_host = per.cpu().tolist()
        return {"requests_verified": int(sum(per_host)), "requests_checked": len(per_host),
                "first_bad_request": next((i for i, v in enumerate(per_host) if not v), None),
                "final_blocks_checked": blocks_final, "final_blocks_ok": bool(final_ok),
                "canary_blocks_checked": int(untouched.numel()), "canary_blocks_ok": bool(canary_ok),
                "regions_per_block": self.shape.regions, "source_bytes": sorted(set(self.mode.values()))}


def run_rank(rank: int, plans: List[RankPlan], cfg: StreamConfig, group=None, world_size: Optional[int] = None) -> Optional[Dict]:
    """One worker rank's part of the stream.  torch.distributed is initialised (any backend that moves Python
    objects and barriers on CPU tensors), the rank's device is current.  Returns this rank's result record.
    `world_size` > len(plans): the job has more ranks than the topology needs (TP4 -> TP2 uses 6 of 8 GPUs); the extra
    ranks only take part in the collectives, so no sub-group has to be created."""
    import torch
    import torch.distributed as dist

    from . import synth
    from .api import HND, NHD  # noqa: F401
    from .workers import DecodeWorker, PrefillWorker

    world = world_size or len(plans)
    if rank >= len(plans):  # spectator: the same collectives in the same order, nothing else
        dist.all_gather_object([None] * world, None, group=group)
        for _ in range(2 if cfg.verify else 1):
            dist.barrier(group=group)
        dist.all_gather_object([None] * world, None, group=group)
        dist.barrier(group=group)
        return None
    me = plans[rank]
    shape = KV_SHAPES[cfg.model]
    trace = (lambda m: (sys.stderr.write(f"[harness rank {rank} {me.role}{me.replica}/tp{me.tp_rank}] {m}\n"), sys.stderr.flush())) \
        if os.environ.get("KVX_HARNESS_TRACE") else (lambda m: None)
    p_ranks = [p for p in plans if p.role == "prefill"]
    d_ranks = [p for p in plans if p.role == "decode"]
    p_tp, d_tp = p_ranks[0].tp_size, d_ranks[0].tp_size
    n_p_rep = 1 + max(p.replica for p in p_ranks)
    n_d_rep = 1 + max(p.replica for p in d_ranks)
    edges = plan_edges(p_tp, d_tp, shape.kv_heads)
    hp, hd = heads_per_rank(shape.kv_heads, p_tp), heads_per_rank(shape.kv_heads, d_tp)
    chunk_d = shape.block_tokens * hd * shape.head_dim * shape.elem_bytes
    per_d = plan_requests(cfg, shape, n_p_rep, n_d_rep, lambda n: n * shape.regions * chunk_d)
    all_isl = [r.isl for lst in per_d.values() for r in lst] or [16]
    c = max(1, cfg.concurrency)
    nb = pool_blocks(shape, all_isl, c)
    hold = RESIDENCY + c
    layout = me.layout if (me.role == "prefill" or cfg.decode_layout is None) else cfg.decode_layout
    Worker = PrefillWorker if me.role == "prefill" else DecodeWorker
    w = Worker(me.device, shape, nb, layout, tp_rank=me.tp_rank, tp_size=me.tp_size, seed=SEED + rank, **cfg.options)
    if me.role == "prefill":
        synth.fill_random(w.kv, SEED + 1000 * me.replica + me.tp_rank)
    else:
        synth.fill_canary(w.kv)
    torch.cuda.synchronize()
    trace("pool ready, handshake")
    metas = [None] * world
    dist.all_gather_object(metas, w.metadata(), group=group)
    trace("handshake done")

    def p_rank_of(rep: int, t: int) -> int:
        return next(p.rank for p in p_ranks if p.replica == rep and p.tp_rank == t)

    def d_rank_of(rep: int, t: int) -> int:
        return next(p.rank for p in d_ranks if p.replica == rep and p.tp_rank == t)

    # block ids: a pure function of the plan, so every rank computes the lists it needs itself
    def prefill_ids(p_rep: int) -> List[np.ndarray]:
        # prefill contents never change during a stream, so its ids only need to look like a live allocator's
        return _pre_allocate(nb, [r.n for r in _served_by(per_d, p_rep)], SEED + 7 + 100 * p_rep, hold)

    # consumers of prefill TP rank t: the decode TP ranks with an edge from it (fan-out when d_tp > p_tp)
    def consumers_of(t: int) -> List[int]:
        return sorted(e.d_rank for e in edges if e.p_rank == t)

    G = max(1, cfg.layerwise)
    result = None
    if me.role == "decode":
        mine = per_d[me.replica]
        srcs = [e for e in edges if e.d_rank == me.tp_rank]
        p_reps = sorted({r.p_rep for r in mine})
        sources, p_layouts = {}, {}
        for pr in p_reps:
            for e in srcs:
                key = pr * 1000 + e.p_rank
                w.add_remote(key, metas[p_rank_of(pr, e.p_rank)])
                sources[key] = (pr, e)
                p_layouts[key] = plans[p_rank_of(pr, e.p_rank)].layout
        ids_of_p = {pr: prefill_ids(pr) for pr in p_reps}
        remote_ids = [ids_of_p[r.p_rep][r.p_serial] for r in mine]
        local_ids = _pre_allocate(nb, [r.n for r in mine], SEED + 13 + 100 * me.replica, hold)  # same on all TP ranks of a replica
        served_on = {pr: len(_served_by(per_d, pr)) for pr in p_reps}  # flag values grow from pass to pass

        def run_pass(pass_idx: int, ver):
            """One run of the whole stream.  `ver` = the verifier that checks every request as it lands (the
            verification pass), or None (the timed pass: nothing but the transfers touches the GPU)."""
            val_of = lambda r: pass_idx * served_on[r.p_rep] + r.p_serial + 1  # noqa: E731
            trace(f"pass {pass_idx} starts")
            lanes = [[torch.cuda.Stream() for _ in srcs] for _ in range(c)]
            ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in mine]
            ev_ready = [torch.cuda.Event(enable_timing=True) for _ in mine]  # gated streams: last layer group is ready
            gated = G > 1 or cfg.prefill_ms > 0   # the pulls wait on the GPU for the prefill worker's "layers ready" flags
            t_sub, t_done, t_ready = [0.0] * len(mine), [0.0] * len(mine), [0.0] * len(mine)
            group_of: Dict[int, Tuple[int, List[int]]] = {}
            checked_args: List[object] = [None] * len(mine)
            # checks run on the verifier's stream in the order requests LAND; before request i reuses the blocks of
            # requests <= i - hold - 1, every one of their checks must have read them: wait for the latest-queued of them
            chk_events: List[object] = []
            chk_pos: List[Optional[int]] = [None] * len(mine)
            prefix = {"upto": -1, "pos": -1}

            def wait_for_checks(i: int, streams) -> None:
                limit = i - hold - 1
                if ver is None or limit < 0:
                    return
                while prefix["upto"] < limit:
                    prefix["upto"] += 1
                    prefix["pos"] = max(prefix["pos"], chk_pos[prefix["upto"]])
                for s_ in streams:
                    s_.wait_event(chk_events[prefix["pos"]])

            def remote_of(i: int) -> Dict[int, np.ndarray]:
                return {mine[i].p_rep * 1000 + e.p_rank: remote_ids[i] for e in srcs}

            def edges_of(i: int) -> Dict[int, object]:
                return {mine[i].p_rep * 1000 + e.p_rank: e for e in srcs}

            def flag_slot(i: int, e) -> int:
                cons = consumers_of(e.p_rank)
                return _release_slot(mine[i].p_serial, cons.index(me.tp_rank), len(cons), w.remotes[mine[i].p_rep * 1000 + e.p_rank].meta.num_flags)

            use_queue = cfg.queue and cfg.batch <= 1 and len(srcs) == 1 and p_tp == d_tp and not gated
            tickets: Dict[int, int] = {}
            if use_queue:   # the receivers exist before the clock starts (their set-up allocates and is not part of a request)
                for pr in p_reps:
                    w.pull_queue(pr * 1000 + srcs[0].p_rank)

            def issue(i: int, lane) -> None:
                """Queue request i on a lane: one pull per source edge (fan-in when p_tp > d_tp), joined on lane[0]."""
                r = mine[i]
                if use_queue:
                    # the receiver runs on a stream of its own: the checks that must have read the blocks this request
                    # reuses are waited for on the host, and completion is the request's ticket
                    limit = i - hold - 1
                    if ver is not None and limit >= 0:
                        while prefix["upto"] < limit:
                            prefix["upto"] += 1
                            prefix["pos"] = max(prefix["pos"], chk_pos[prefix["upto"]])
                        chk_events[prefix["pos"]].synchronize()
                    tickets[i] = w.pull_queued(r.p_rep * 1000 + srcs[0].p_rank, remote_ids[i], local_ids[i],
                                               notify_slot=flag_slot(i, srcs[0]), value=val_of(r))
                    return
                wait_for_checks(i, lane)
                ev[i][0].record(lane[0])
                for j, e in enumerate(srcs):
                    if j:
                        lane[j].wait_event(ev[i][0])
                    key = r.p_rep * 1000 + e.p_rank
                    edge = e if (p_tp != d_tp) else None
                    if gated:
                        for g in range(G):
                            w.wait(w.layer_ready_slot(r.p_serial, g, G, w.num_flags, j, len(srcs)), val_of(r), 60000, lane[j])
                            if g == G - 1 and j == 0:
                                ev_ready[i].record(lane[0])
                            w.pull_layers(key, remote_ids[i], local_ids[i], g, G,
                                          notify_slot=flag_slot(i, e) if g == G - 1 else None, value=val_of(r),
                                          edge=edge, stream=lane[j])
                    else:
                        w.pull(key, remote_ids[i], local_ids[i], notify_slot=flag_slot(i, e), value=val_of(r),
                               edge=edge, host_tables=True, stream=lane[j])
                for j in range(1, len(srcs)):
                    done_j = torch.cuda.Event()
                    done_j.record(lane[j])
                    lane[0].wait_event(done_j)
                ev[i][1].record(lane[0])

            def landed(i: int) -> None:
                if ver is not None:
                    chk_events.append(ver.check_request(None if use_queue else ev[i][1], remote_of(i), local_ids[i], edges_of(i),
                                                        corrupt=cfg.corrupt is not None and tuple(cfg.corrupt) == (me.replica, i)))
                    chk_pos[i] = len(chk_events) - 1
                checked_args[i] = (remote_of(i), local_ids[i], edges_of(i))

            dist.barrier(group=group)
            torch.cuda.synchronize()
            t_begin = time.perf_counter()
            now_ms = lambda: 1e3 * (time.perf_counter() - t_begin)  # noqa: E731
            batch = cfg.batch if (cfg.batch > 1 and len(srcs) == 1 and p_tp == d_tp and cfg.route == "pair"
                                  and cfg.ready_ms is None and G == 1) else 1
            if batch > 1:
                # one scheduler step = up to `batch` queued requests, received in one launch; closed loop over steps
                lane0 = lanes[0][0]
                for g0 in range(0, len(mine), batch):
                    qs = list(range(g0, min(len(mine), g0 + batch)))
                    key = mine[g0].p_rep * 1000 + srcs[0].p_rank
                    wait_for_checks(g0, [lane0])
                    t0 = time.perf_counter()
                    ev[g0][0].record(lane0)
                    w.pull_many(key, [(remote_ids[i], local_ids[i], flag_slot(i, srcs[0]), val_of(mine[i]), None) for i in qs],
                                stream=lane0)
                    ev[g0][1].record(lane0)
                    ev[g0][1].synchronize()
                    t1 = time.perf_counter()
                    for i in qs:
                        t_sub[i], t_done[i], group_of[i] = t0, t1, (g0, qs)
                        ev[i] = ev[g0]
                        landed(i)
            else:
                # at most c requests in flight, each on a lane of its own; a request becomes eligible when its KV is
                # ready (arrival-driven) or at once (closed loop); FIFO in the planned order
                nxt, inflight, free_lanes = 0, {}, list(range(c))
                while nxt < len(mine) or inflight:
                    progressed = False
                    for i in list(inflight):
                        if (w.pull_queue(mine[i].p_rep * 1000 + srcs[0].p_rank).poll(tickets[i]) if use_queue else ev[i][1].query()):
                            t_done[i] = time.perf_counter()
                            free_lanes.append(inflight.pop(i))
                            landed(i)
                            progressed = True
                    while nxt < len(mine) and free_lanes and (mine[nxt].ready_ms is None or mine[nxt].ready_ms <= now_ms()):
                        lane_id = free_lanes.pop(0)
                        t_sub[nxt] = time.perf_counter()
                        t_ready[nxt] = t_begin + (mine[nxt].ready_ms or 0.0) / 1e3 if mine[nxt].ready_ms is not None else t_sub[nxt]
                        issue(nxt, lanes[lane_id])
                        inflight[nxt] = lane_id
                        nxt += 1
                        progressed = True
                    if not progressed and not inflight and nxt < len(mine):
                        wait = mine[nxt].ready_ms - now_ms()
                        if wait > 0.2:
                            time.sleep((wait - 0.1) / 1e3)
            t_end = time.perf_counter() if use_queue else None   # every ticket has been seen: nothing is in flight
            if use_queue:
                for q_ in w.__dict__.get("_queues", {}).values():
                    q_.park()
            torch.cuda.synchronize()
            t_end = t_end or time.perf_counter()
            w.ctx.check()
            per_req = []
            for i, r in enumerate(mine):
                b = r.n * shape.regions * chunk_d
                if i in group_of:  # batched: the step's requests complete together; rate = the step's bytes over its time
                    g0, qs = group_of[i]
                    xfer = ev[g0][0].elapsed_time(ev[g0][1]) / 1e3
                    rate = sum(mine[x].n for x in qs) * shape.regions * chunk_d / xfer / 1e9
                    ttft = 1e3 * (t_done[i] - t_sub[i])
                elif use_queue:   # no launch to bracket with events: host clock, submit -> ticket seen
                    xfer = t_done[i] - t_sub[i]
                    rate = b / xfer / 1e9 if xfer > 0 else 0.0
                    ttft = 1e3 * (t_done[i] - t_ready[i])
                else:
                    xfer = ev[i][0].elapsed_time(ev[i][1]) / 1e3
                    rate = b / xfer / 1e9 if xfer > 0 else 0.0
                    ttft = 1e3 * (t_done[i] - t_ready[i])
                after_prefill = ev_ready[i].elapsed_time(ev[i][1]) if (gated and i not in group_of) else None
                per_req.append({"request": r.q, "added_after_prefill_ms": after_prefill, "isl": r.isl, "num_blocks": r.n, "bytes_transferred": b, "prefill_replica": r.p_rep,
                                "num_descriptors": r.n * shape.regions * len(srcs), "xfer_time_seconds": xfer,
                                "throughput_GBps": rate, "batched_with": len(group_of[i][1]) if i in group_of else 1,
                                "ready_ms": r.ready_ms, "queued_ms": 1e3 * (t_sub[i] - t_ready[i]) if i not in group_of else 0.0,
                                "time_to_first_token_added_ms": ttft})
            return per_req, t_end - t_begin, [a_ for a_ in checked_args if a_ is not None]

        # Pass 0 is timed and runs nothing but the transfers; its final state is checked whole-pool.  Pass 1 repeats
        # the identical stream (same ids, same order, same concurrency) into a canary-filled pool and compares every
        # request with its source the moment it lands, which costs GPU time and therefore is not the timed pass.
        per_req, wall, landed_args = run_pass(0, None)
        ver = _Verifier(torch, w, shape, nb, me, sources, p_tp, d_tp, hp, p_layouts) if cfg.verify else None
        verdict = None
        if ver is not None:
            timed_final = ver.finish(landed_args)
            synth.fill_canary(w.kv)
            torch.cuda.synchronize()
            _, _, landed_args = run_pass(1, ver)
            verdict = ver.finish(landed_args)
            verdict["timed_pass_final_blocks_ok"] = timed_final["final_blocks_ok"]
            verdict["timed_pass_canary_blocks_ok"] = timed_final["canary_blocks_ok"]
        ok = None
        if verdict is not None:
            ok = (verdict["requests_verified"] == len(mine) and verdict["final_blocks_ok"] and verdict["canary_blocks_ok"]
                  and verdict["timed_pass_final_blocks_ok"] and verdict["timed_pass_canary_blocks_ok"])
        total_b = sum(r["bytes_transferred"] for r in per_req)
        result = {"rank": rank, "role": "decode", "replica": me.replica, "tp_rank": me.tp_rank, "device": me.device,
                  "sources": [e.p_rank for e in srcs], "wall_seconds": wall, "bytes_transferred": total_b,
                  "ingress_GBps": total_b / wall / 1e9, "verified": ok, "verification": verdict,
                  "requests": per_req}
    else:
        served = _served_by(per_d, me.replica)
        cons = consumers_of(me.tp_rank)
        # which decode ranks read this prefill rank: handshake with them when the stream is layer-wise (the
        # "layers ready" flags live in THEIR memory)
        gated = G > 1 or cfg.prefill_ms > 0
        if gated:
            for r_ in {(r.d_rep, t) for r in served for t in cons}:
                w.add_remote(r_[0] * 1000 + r_[1], metas[d_rank_of(*r_)])
        from .control import LeaseTable

        leases = LeaseTable(lease_seconds=600.0)
        passes = 2 if cfg.verify else 1   # the decode side repeats the stream once more to check every request
        released = [0] * passes
        all_released = True
        for pass_idx in range(passes):
            base = pass_idx * len(served)
            for r in served:
                if cons:
                    leases.grant(f"{pass_idx}:{r.p_serial}", consumers=len(cons))
            trace(f"pass {pass_idx}: barrier")
            dist.barrier(group=group)
            if gated:
                # synthetic prefill: each request's layers become ready group by group, prefill_ms per request, one
                # request after the other on the prefill GPU (a device-side spin stands in for the layer kernels)
                stream = torch.cuda.current_stream()
                per_group_cycles = int(cfg.prefill_ms / G * 1e-3 * 1.9e9)  # ~SM clock; only the order of magnitude matters
                for r in served:
                    for g in range(G):
                        if per_group_cycles > 0:
                            torch.cuda._sleep(per_group_cycles)
                        for t in cons:  # my position among the prefill ranks decode rank t reads (fan-in when p_tp > d_tp)
                            feeders = sorted(e.p_rank for e in edges if e.d_rank == t)
                            w.publish_layers(r.d_rep * 1000 + t, r.p_serial, g, G, stream, src=feeders.index(me.tp_rank),
                                             n_src=len(feeders), value=base + r.p_serial + 1)
            # the producer side only learns of completion through release flags written by the consumers
            deadline = time.time() + 600
            pending = {(r.p_serial, ci) for r in served for ci in range(len(cons))}
            while pending and time.time() < deadline:
                for key in sorted(pending)[:64]:
                    serial, ci = key
                    if w.poll(_release_slot(serial, ci, len(cons), w.num_flags), base + serial + 1):
                        pending.discard(key)
                        if leases.release(f"{pass_idx}:{serial}"):
                            released[pass_idx] += 1
                time.sleep(0.001)
            all_released = all_released and not pending
            trace(f"pass {pass_idx}: {released[pass_idx]} of {len(served)} released")
        result = {"rank": rank, "role": "prefill", "replica": me.replica, "tp_rank": me.tp_rank, "device": me.device,
                  "released": all_released and len(leases) == 0, "requests_served": len(served), "requests_released": released[0],
                  "requests_released_in_verification_pass": released[-1] if passes > 1 else None,
                  "consumers_per_request": len(cons)}
    trace("stream done, gathering results")
    gathered = [None] * world
    dist.all_gather_object(gathered, result, group=group)
    dist.barrier(group=group)
    w.close()
    result["all"] = gathered if rank == plans[0].rank else None
    return result


def summarize(ranks: List[Dict], topo_desc: Dict, cfg: StreamConfig, world: int, n_devices: int) -> Dict:
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
    by_isl: Dict[int, List[Dict]] = {}
    for q in reqs:
        by_isl.setdefault(q["isl"], []).append(q)
    verified = None
    if cfg.verify:
        verified = all(r["verified"] is True for r in dec)
    ver = [r["verification"] for r in dec if r.get("verification")]
    summary = {
        "topology": topo_desc, "model": asdict(shape), "isl": sorted(by_isl), "concurrency": cfg.concurrency,
        "route": cfg.route, "layerwise": cfg.layerwise, "resident_receiver": bool(cfg.queue), "arrival_driven": cfg.ready_ms is not None,
        "gpus_visible": n_devices, "ranks": world, "oversubscribed": world > n_devices,
        "requests": len(reqs), "bytes_transferred": total, "wall_seconds": wall,
        "aggregate_throughput_GBps": total / wall / 1e9,
        "per_decode_rank_ingress_GBps": {f"replica{r['replica']}/tp{r['tp_rank']}": r["ingress_GBps"] for r in dec},
        "per_edge_throughput_GBps_median": statistics.median(q["throughput_GBps"] for q in reqs) / max(1, len(dec[0]["sources"])),
        "per_request_throughput_GBps_median": statistics.median(q["throughput_GBps"] for q in reqs),
        "pct_of_nvlink_900": 100 * statistics.median(q["throughput_GBps"] for q in reqs) / 900.0,
        "per_isl": {str(k): {"requests": len(v), "throughput_GBps_median": statistics.median(q["throughput_GBps"] for q in v),
                             "time_to_first_token_added_ms_p50": pct(sorted(q["time_to_first_token_added_ms"] for q in v), 0.5),
                             "time_to_first_token_added_ms_p90": pct(sorted(q["time_to_first_token_added_ms"] for q in v), 0.9)}
                    for k, v in sorted(by_isl.items())},
        "xfer_time_seconds": {"p50": pct(xfer, 0.5), "p90": pct(xfer, 0.9), "max": xfer[-1]},
        "time_to_first_token_added_ms": {"p50": pct(ttft, 0.5), "p90": pct(ttft, 0.9), "max": ttft[-1],
                                         "mean": sum(ttft) / len(ttft)},
        "num_descriptors": sum(q["num_descriptors"] for q in reqs),
        "verified_bit_exact": verified, "verification": ("timed pass: the whole decode pool at the end; verification pass (same stream again into a canary-filled pool): "
                                                            "every request when it lands + the whole pool at the end") if cfg.verify else "off (verify=0)",
        "requests_verified": sum(v["requests_verified"] for v in ver) if ver else 0,
        "blocks_checked": sum(v["final_blocks_checked"] for v in ver) if ver else 0,
        "canary_blocks_checked": sum(v["canary_blocks_checked"] for v in ver) if ver else 0,
        "verification_source_bytes": sorted({s for v in ver for s in v["source_bytes"]}),
        "verification_failures": [{"replica": r["replica"], "tp_rank": r["tp_rank"],
                                   **{k: r["verification"][k] for k in ("requests_verified", "requests_checked", "first_bad_request", "final_blocks_ok",
                                                                        "canary_blocks_ok", "timed_pass_final_blocks_ok", "timed_pass_canary_blocks_ok")}}
                                  for r in dec if r.get("verification") and r["verified"] is not True],
        "prefill_blocks_released": all(r["released"] for r in pre),
        "releases_observed": sum(r["requests_released"] for r in pre), "releases_expected": sum(r["requests_served"] for r in pre if r["consumers_per_request"]),
        "seed": SEED,
    }
    return summary


def _worker_main(rank: int, world: int, port: int, topo: Topology, cfg: StreamConfig, n_devices: int, out_dir: str):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    import torch
    import torch.distributed as dist

    plans = rank_plans(topo, n_devices)
    torch.cuda.set_device(plans[rank].device)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    result = run_rank(rank, plans, cfg)
    if rank == 0:
        with open(os.path.join(out_dir, "ranks.json"), "w") as f:
            json.dump(result["all"], f)
    dist.destroy_process_group()


def run_stream(topo: Topology, cfg: StreamConfig, out_dir: str, n_devices: Optional[int] = None) -> Dict:
    """Spawn one process per worker rank, run the stream, collect per-request records and a summary."""
    import torch
    import torch.multiprocessing as mp

    if not torch.cuda.is_available():
        raise RuntimeError("the kvx harness needs CUDA devices: there is no CPU fallback")
    if not topo.disaggregated:
        raise ValueError("the manifest has no prefill/decode split: nothing is handed off in aggregated mode")
    n_devices = n_devices or torch.cuda.device_count()
    plans = rank_plans(topo, n_devices)
    world = len(plans)
    os.makedirs(out_dir, exist_ok=True)
    port = _free_port()
    mp.spawn(_worker_main, args=(world, port, topo, cfg, n_devices, out_dir), nprocs=world, join=True)
    with open(os.path.join(out_dir, "ranks.json")) as f:
        ranks = json.load(f)
    summary = summarize(ranks, topo.describe(), cfg, world, n_devices)
    dec = [r for r in ranks if r and r["role"] == "decode"]
    pre = [r for r in ranks if r and r["role"] == "prefill"]
    with open(os.path.join(out_dir, "results.json"), "w") as f:
        json.dump({"summary": summary, "decode_ranks": dec, "prefill_ranks": pre}, f, indent=1)
    return summary
