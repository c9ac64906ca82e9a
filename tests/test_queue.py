"""kvx_queue_*: the resident receiver (a persistent kernel fed from a queue in pinned host memory) must move
exactly the bytes the per-request entry points move — compared here with the oracle over the ENTIRE destination
pool — and must never make the process hang: it parks when idle and comes back on the next submit."""
import time

import numpy as np
import pytest
import torch

import kvx
from gpu_util import DevPool
from oracle import kvx_oracle as o

pytestmark = pytest.mark.gpu

GEOMS = [dict(), dict(max_ctas=5), dict(stages=6, stage_bytes=32768, max_ctas=43), dict(stages=2, stage_bytes=16384, consumer_warps=4)]


def _geom_id(g):
    return "-".join(f"{k}{v}" for k, v in g.items()) or "default"


@pytest.mark.parametrize("geom", GEOMS, ids=_geom_id)
@pytest.mark.parametrize("layouts", [(o.HND, o.NHD), (o.NHD, o.HND), (o.HND, o.HND)], ids=lambda l: f"{l[0]}to{l[1]}")
def test_queue_matches_oracle(geom, layouts):
    geom = dict(geom)
    max_ctas = geom.pop("max_ctas", 0)
    c = kvx.Context(0, **geom)
    L, H, D, nb = 3, 8, 128, 400
    src = o.make_pool(L, H, D, nb, layouts[0], seed=11)
    dst = o.make_pool(L, H, D, nb, layouts[1], fill="canary")
    ds, dd = DevPool(c, src), DevPool(c, dst)
    flags, _ = c.flags()
    words = torch.zeros(16, dtype=torch.int64).pin_memory()
    perm_s, perm_d = o.make_table(nb, 380, 1), o.make_table(nb, 380, 2)
    counts = [64, 1, 0, 150, 7, 33, 100, 25]   # back to back: the CTAs run from one request into the next
    with c.queue(ds.kp, dd.kp, max_ctas=max_ctas) as q:
        at, tickets = 0, []
        for k, n in enumerate(counts):
            ts, td = perm_s[at:at + n], perm_d[at:at + n]
            tickets.append(q.submit(ts, td, n, flag=flags + 8 * (20 + k), flag_val=100 + k, h_done=words[k:k + 1]))
            at += n
        for t in tickets:
            q.wait(t)
        assert q.launches >= 1 and all(q.poll(t) for t in tickets)
        assert words[:len(counts)].tolist() == [100 + k for k in range(len(counts))]
        for k in range(len(counts)):
            assert c.flag_poll(flags + 8 * (20 + k), 100 + k)
        q.park()
        at = 0
        for n in counts:
            o.handoff(src, perm_s[at:at + n], dst, perm_d[at:at + n])
            at += n
        assert np.array_equal(dd.bytes(), dst.buf)
        assert np.array_equal(ds.bytes(), src.buf)
    c.check()
    c.close()


def test_queue_parks_when_idle_and_comes_back():
    c = kvx.Context(0)
    nb = 64
    src = o.make_pool(1, 8, 64, nb, o.HND, seed=2)
    dst = o.make_pool(1, 8, 64, nb, o.NHD, fill="canary")
    ds, dd = DevPool(c, src), DevPool(c, dst)
    q = c.queue(ds.kp, dd.kp, idle_us=300)
    t1 = q.submit(np.array([3, 4], np.int32), np.array([9, 8], np.int32))
    q.wait(t1)
    # a resident kernel would block this for ever; a parked one does not
    t0 = time.perf_counter()
    torch.cuda.synchronize()
    assert time.perf_counter() - t0 < 1.0
    time.sleep(0.01)
    t2 = q.submit(np.array([5], np.int32), np.array([7], np.int32))   # parked: this submit launches the kernel again
    q.wait(t2)
    assert q.launches == 2
    torch.cuda.synchronize()
    o.handoff(src, [3, 4], dst, [9, 8])
    o.handoff(src, [5], dst, [7])
    assert np.array_equal(dd.bytes(), dst.buf)
    # submit storms around the park: every request must be served exactly once whatever the interleaving
    rng = np.random.default_rng(0)
    for k in range(200):
        a, b = int(rng.integers(1, nb)), int(rng.integers(1, nb))
        t = q.submit(np.array([a], np.int32), np.array([b], np.int32))
        o.handoff(src, [a], dst, [b])
        if k % 3 == 0:
            q.wait(t)
            time.sleep(float(rng.uniform(0.0002, 0.0005)))   # around the 300 us idle limit
    q.wait(t)
    q.park()
    assert np.array_equal(dd.bytes(), dst.buf)
    assert q.launches >= 3
    c.close()   # stops the queue itself: no hang


def test_queue_ring_and_table_ring_wrap():
    """More requests than ring slots (256) and more ids than the table ring holds (1 Mi entries): submit waits for
    the slot / span to come free instead of overwriting a request in flight."""
    c = kvx.Context(0)
    nb = 70_001
    src = o.make_pool(1, 1, 8, nb, o.NHD, seed=4, block_tokens=1)
    dst = o.make_pool(1, 1, 8, nb, o.NHD, fill="canary", block_tokens=1)
    ds, dd = DevPool(c, src), DevPool(c, dst)
    want = dst.buf.copy()
    sv, wv = src.buf.reshape(2, nb, 16), want.reshape(2, nb, 16)
    with c.queue(ds.kp, dd.kp) as q:
        for k in range(600):  # tiny requests: slot ring wraps twice
            a, b = 1 + (7 * k) % (nb - 1), 1 + (11 * k) % (nb - 1)
            t = q.submit(np.array([a], np.int32), np.array([b], np.int32))
            wv[:, b] = sv[:, a]
        for k in range(20):   # 20 x 140k ids: the table ring (1 Mi) wraps
            ts, td = o.make_table(nb, 70_000, 300 + k), o.make_table(nb, 70_000, 400 + k)
            t = q.submit(ts, td)
            wv[:, td] = sv[:, ts]
        q.wait(t, 20000)
        q.park()
        assert np.array_equal(dd.bytes(), want.reshape(-1))
    c.close()


def test_queue_scheduler_step_in_one_call():
    """kvx_queue_submit_batch: a step's requests with one call, each with its own flag and completion word."""
    c = kvx.Context(0)
    nb = 300
    src = o.make_pool(2, 8, 128, nb, o.HND, seed=8)
    dst = o.make_pool(2, 8, 128, nb, o.NHD, fill="canary")
    ds, dd = DevPool(c, src), DevPool(c, dst)
    flags, _ = c.flags()
    words = torch.zeros(12, dtype=torch.int64).pin_memory()
    perm_s, perm_d = o.make_table(nb, 240, 1), o.make_table(nb, 240, 2)
    reqs, at = [], 0
    for k, n in enumerate([8, 40, 0, 3, 64, 16, 1, 30, 8, 8, 50, 12]):
        reqs.append((perm_s[at:at + n] if n else None, perm_d[at:at + n] if n else None, flags + 8 * (40 + k), 7 + k, words[k:k + 1]))
        at += n
    with c.queue(ds.kp, dd.kp) as q:
        last = q.submit_many(q.pack(reqs))
        for t in range(last - 11, last + 1):   # tickets of one call are consecutive
            q.wait(t)
        assert words.tolist() == [7 + k for k in range(12)]
        q.park()
    for s_ids, d_ids, *_ in reqs:
        if d_ids is not None:
            o.handoff(src, s_ids, dst, d_ids)
    assert np.array_equal(dd.bytes(), dst.buf)
    for k in range(12):
        assert c.flag_poll(flags + 8 * (40 + k), 7 + k)
    c.close()


def test_queue_tp_slice_and_errors():
    c = kvx.Context(0)
    nb = 50
    src = o.make_pool(2, 2, 128, nb, o.HND, seed=5)            # a TP4 shard: 2 heads
    dst = o.make_pool(2, 4, 128, nb, o.HND, fill="canary")     # a TP2 shard: 4 heads; this edge fills slot 1
    ds, dd = DevPool(c, src), DevPool(c, dst)
    xf = kvx.Xform(0, 2, 2, kvx.HND)
    ts, td = o.make_table(nb, 30, 1), o.make_table(nb, 30, 2)
    with c.queue(ds.kp, dd.kp, xf) as q:
        q.wait(q.submit(ts, td))
        bad = td.copy(); bad[3] = nb
        with pytest.raises(kvx.KvxError) as e:
            q.submit(ts, bad)
        assert e.value.code == kvx.api.ERANGE
        with pytest.raises(kvx.KvxError) as e:   # the completion word must be pinned: the kernel writes it itself
            q.submit(ts, td, h_done=np.zeros(1, np.int64))
        assert e.value.code == kvx.api.EINVAL
        q.park()
    o.handoff(src, ts, dst, td, o.Xform(0, 2, 2))
    assert np.array_equal(dd.bytes(), dst.buf)
    c.check()
    c.close()
