"""Which prefill worker feeds which decode worker, and when: the handoff side of the Frontend / router the
reference's manifests put in front of the workers (`componentType: frontend`,
examples/deploy/vllm/disagg.yaml:12-17) and scale with `replicas` (:22,42).  The manifests never say how the
router pairs workers; what the links can carry does.  On an NVSwitch node every GPU has ONE ingress and ONE
egress NVLink port at ~780 GB/s achievable: two handoffs that share a port each get half of it, so both finish
late (4 prefill workers into 1 decode worker: 203 GB/s per request, profiles/r01s2_results incast_4p_to_1d),
while the same two handoffs one after the other finish at T and 2T.

`Router` is an online list scheduler over those 2·(nP + mD) ports.  A request is admitted with the time its KV
will be ready (prefill done, block table known) and its payload bytes; the router picks the (prefill, decode)
pair whose two ports are both free earliest, books both ports for the transfer's estimated duration
t = fixed + bytes / rate (the fit DESIGN.md §6 gives: 20 us + bytes / 794 GB/s), and tells the decode worker
the order in which to pull.  Each decode worker then pulls ONE request at a time in that order, so no port
ever carries two full-rate streams.  `complete()` feeds measured completion times back, which keeps the
books honest when a transfer runs faster or slower than the estimate.

Static policies for comparison (what kvx/harness.py did before): "pair" (request k → decode k % mD, prefill
= decode index) and "spread" (prefill rotates per request), both issued as soon as the KV is ready.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence

LINK_GBPS = 780.0   # achievable per port and direction (pull), profiles/r01s2_bench_n2.json
FIXED_MS = 0.020    # launch + first/last NVLink round trip of one pull


@dataclass(frozen=True)
class Assignment:
    q: int            # request index
    p: int            # prefill replica
    d: int            # decode replica
    ready_ms: float   # KV ready on the prefill worker
    start_ms: float   # when the router expects the pull to start
    est_ms: float     # estimated duration
    order: int        # position in decode replica d's pull order

    @property
    def end_ms(self) -> float:
        return self.start_ms + self.est_ms


def estimate_ms(nbytes: int, link_gbps: float = LINK_GBPS, fixed_ms: float = FIXED_MS) -> float:
    return fixed_ms + nbytes / (link_gbps * 1e6)


class Router:
    """Earliest-start list scheduling of handoffs over per-GPU ingress / egress ports."""

    def __init__(self, n_prefill: int, n_decode: int, *, link_gbps: float = LINK_GBPS, fixed_ms: float = FIXED_MS):
        if n_prefill < 1 or n_decode < 1:
            raise ValueError("a router needs at least one prefill and one decode worker")
        self.n_p, self.n_d = n_prefill, n_decode
        self.link_gbps, self.fixed_ms = link_gbps, fixed_ms
        self.egress_free = [0.0] * n_prefill   # when each prefill worker's egress port is free again (ms)
        self.ingress_free = [0.0] * n_decode   # same for each decode worker's ingress port
        self.egress_bytes = [0] * n_prefill    # bytes booked so far (tie-break: spread the load)
        self.ingress_bytes = [0] * n_decode
        self._order = [0] * n_decode
        self.booked: Dict[int, Assignment] = {}

    def assign(self, q: int, ready_ms: float, nbytes: int, *, prefill: Optional[int] = None,
               decode: Optional[int] = None) -> Assignment:
        """Book the request.  `prefill` / `decode` pin one side (e.g. prefill already ran somewhere)."""
        if q in self.booked:
            raise KeyError(f"request {q} already assigned")
        est = estimate_ms(nbytes, self.link_gbps, self.fixed_ms)
        best = None
        for p in ([prefill] if prefill is not None else range(self.n_p)):
            for d in ([decode] if decode is not None else range(self.n_d)):
                start = max(ready_ms, self.egress_free[p], self.ingress_free[d])
                key = (start, self.ingress_bytes[d] + self.egress_bytes[p], d, p)
                if best is None or key < best[0]:
                    best = (key, p, d, start)
        _, p, d, start = best
        a = Assignment(q, p, d, ready_ms, start, est, self._order[d])
        self._order[d] += 1
        self.egress_free[p] = self.ingress_free[d] = start + est
        self.egress_bytes[p] += nbytes
        self.ingress_bytes[d] += nbytes
        self.booked[q] = a
        return a

    def complete(self, q: int, done_ms: float) -> None:
        """Measured completion: if the transfer ended at another time than booked and nothing was booked behind
        it on a port yet, move that port's free time to the truth."""
        a = self.booked[q]
        if self.egress_free[a.p] == a.end_ms:
            self.egress_free[a.p] = done_ms
        if self.ingress_free[a.d] == a.end_ms:
            self.ingress_free[a.d] = done_ms


def plan(policy: str, ready_ms: Sequence[float], nbytes: Sequence[int], n_prefill: int, n_decode: int, **router_kw
         ) -> List[Assignment]:
    """Assign a whole (synthetic) request list, in admission order.  "sched" = Router; "pair" / "spread" = the
    static formulas, with start = ready (they pull as soon as the KV is there)."""
    out: List[Assignment] = []
    if policy == "sched":
        r = Router(n_prefill, n_decode, **router_kw)
        for q, (t, b) in enumerate(zip(ready_ms, nbytes)):
            out.append(r.assign(q, t, b))
        return out
    if policy not in ("pair", "spread"):
        raise ValueError(f"unknown routing policy '{policy}'")
    order = [0] * n_decode
    for q, (t, b) in enumerate(zip(ready_ms, nbytes)):
        d = q % n_decode
        p = d % n_prefill if policy == "pair" else (d + q // n_decode) % n_prefill
        out.append(Assignment(q, p, d, t, t, estimate_ms(b, **router_kw), order[d]))
        order[d] += 1
    return out


def port_overlaps(assignments: Sequence[Assignment]) -> int:
    """How many pairs of booked transfers overlap in time on one port (0 for every Router plan)."""
    bad = 0
    for key in ("p", "d"):
        by_port: Dict[int, List[Assignment]] = {}
        for a in assignments:
            by_port.setdefault(getattr(a, key), []).append(a)
        for lst in by_port.values():
            lst.sort(key=lambda a: a.start_ms)
            for x, y in zip(lst, lst[1:]):
                if y.start_ms < x.end_ms - 1e-9:
                    bad += 1
    return bad


def skewed_arrivals(requests: int, seed: int, mean_gap_ms: float = 1.0, burst: int = 4) -> List[float]:
    """KV-ready times of a bursty stream: bursts of up to `burst` requests that become ready together, the gaps
    between bursts exponential with the given mean (numpy PCG64, deterministic)."""
    import numpy as np

    rng = np.random.Generator(np.random.PCG64(seed))
    t, out = 0.0, []
    while len(out) < requests:
        k = int(rng.integers(1, burst + 1))
        out += [t] * min(k, requests - len(out))
        t += float(rng.exponential(mean_gap_ms))
    return out
