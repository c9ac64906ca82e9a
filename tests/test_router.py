"""kvx/router.py: which prefill worker feeds which decode worker (the Frontend/router role of
examples/deploy/vllm/disagg.yaml:12-17 for this path), and the request plan the harness derives from it."""
import numpy as np
import pytest

from kvx import router
from kvx.harness import StreamConfig, draw_isls, plan_requests
from kvx.shapes import KV_SHAPES

SHAPE = KV_SHAPES["meta-llama/Llama-3.1-8B"]
BYTES = lambda n: n * SHAPE.regions * SHAPE.chunk_bytes()  # noqa: E731


def _stream(requests=64, seed=5):
    isl = draw_isls(requests, [1024, 2048, 4096, 8192, 16384, 32768], seed)
    ready = router.skewed_arrivals(requests, seed + 1, mean_gap_ms=1.0, burst=4)
    return isl, ready, [BYTES(SHAPE.blocks_for(x)) for x in isl]


def test_router_never_puts_two_streams_on_one_port():
    isl, ready, nbytes = _stream()
    plan = router.plan("sched", ready, nbytes, 4, 4)
    assert len(plan) == len(isl) and sorted(a.q for a in plan) == list(range(len(isl)))
    assert router.port_overlaps(plan) == 0
    assert all(a.start_ms >= a.ready_ms for a in plan)
    # the static policies do overlap on this stream (that is what the router is for)
    assert router.port_overlaps(router.plan("pair", ready, nbytes, 4, 4)) > 0
    # 4 prefill workers into 1 decode worker: the incast the round-1 harness measured at 203 GB/s per request
    incast = router.plan("sched", [0.0] * 8, [BYTES(512)] * 8, 4, 1)
    assert router.port_overlaps(incast) == 0 and sorted(a.order for a in incast) == list(range(8))
    ends = sorted(a.end_ms for a in incast)
    assert ends[0] == pytest.approx(router.estimate_ms(BYTES(512))) and ends[-1] == pytest.approx(8 * router.estimate_ms(BYTES(512)))


def test_router_balances_bytes_and_is_deterministic():
    isl, ready, nbytes = _stream(128, seed=9)
    a, b = router.plan("sched", ready, nbytes, 4, 4), router.plan("sched", ready, nbytes, 4, 4)
    assert a == b
    per_d = np.zeros(4)
    for x in a:
        per_d[x.d] += nbytes[x.q]
    static = np.zeros(4)
    for x in router.plan("pair", ready, nbytes, 4, 4):
        static[x.d] += nbytes[x.q]
    assert per_d.max() / per_d.mean() <= static.max() / static.mean() + 1e-9
    # modelled mean wait (start - ready) with one-at-a-time ports: booking by free port beats round robin
    def mean_wait(policy):
        r = router.Router(4, 4)
        waits = []
        for q, (t, nb) in enumerate(zip(ready, nbytes)):
            if policy == "sched":
                x = r.assign(q, t, nb)
            else:
                x = r.assign(q, t, nb, prefill=q % 4, decode=q % 4)
            waits.append(x.end_ms - t)
        return sum(waits) / len(waits)
    assert mean_wait("sched") <= mean_wait("pair")


def test_router_feedback_and_pins():
    r = router.Router(2, 2)
    a = r.assign(0, 0.0, BYTES(512))
    r.complete(0, a.end_ms * 0.5)  # finished early: the ports are free earlier than booked
    b = r.assign(1, 0.0, BYTES(512), prefill=a.p, decode=a.d)
    assert b.start_ms == pytest.approx(a.end_ms * 0.5)
    with pytest.raises(KeyError):
        r.assign(0, 0.0, 1)
    with pytest.raises(ValueError):
        router.plan("nope", [0.0], [1], 1, 1)


def test_plan_requests_is_consistent_for_every_policy():
    isl, ready, _ = _stream(24)
    for route in ("pair", "spread", "sched"):
        cfg = StreamConfig(model=SHAPE.name, isl=isl, ready_ms=ready, route=route)
        per_d = plan_requests(cfg, SHAPE, 4, 4, BYTES)
        reqs = [r for lst in per_d.values() for r in lst]
        assert sorted(r.q for r in reqs) == list(range(24))
        for d, lst in per_d.items():
            assert [r.idx for r in lst] == list(range(len(lst))) and all(r.d_rep == d for r in lst)
        for p in range(4):  # serial numbers on a prefill replica are 0..k-1, each once
            ser = sorted(r.p_serial for r in reqs if r.p_rep == p)
            assert ser == list(range(len(ser)))
    # closed loop: a stream per decode replica; spread rotates the prefill replica per request
    cfg = StreamConfig(model=SHAPE.name, isl=[1024] * 6, route="spread")
    per_d = plan_requests(cfg, SHAPE, 4, 2, BYTES)
    assert [r.p_rep for r in per_d[1]] == [1, 2, 3, 0, 1, 2]
    cfg = StreamConfig(model=SHAPE.name, isl=[1024], per_replica_isl=[[1024, 2048], [4096]])
    per_d = plan_requests(cfg, SHAPE, 2, 2, BYTES)
    assert [r.isl for r in per_d[0]] == [1024, 2048] and [r.isl for r in per_d[1]] == [4096]
    with pytest.raises(ValueError):
        plan_requests(StreamConfig(model=SHAPE.name, isl=[1024], route="sched"), SHAPE, 1, 1, BYTES)
