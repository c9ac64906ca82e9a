"""CPU tests of the oracle itself: golden vectors, numpy vs C restatement, domain properties."""
import json
import os

import numpy as np
import pytest
from hypothesis import given, settings, strategies as st

from oracle import kvx_oracle as o
from oracle import kvx_oracle_c as oc

pytestmark = pytest.mark.filterwarnings("ignore")


def _pool_from_regions(arr: np.ndarray, layout: int, heads: int, head_dim: int, N: int = 16, E: int = 2) -> o.Pool:
    """arr: [regions, num_blocks, ...chunk...] uint16 → split-arrangement Pool over its bytes."""
    regions, nb = arr.shape[0], arr.shape[1]
    buf = np.ascontiguousarray(arr).view(np.uint8).reshape(-1).copy()
    chunk = buf.size // (regions * nb)
    return o.Pool(buf, [r * nb * chunk for r in range(regions)], nb, chunk, N, heads, head_dim, E, layout)


def test_golden_layout_hnd_to_nhd(golden_dir):
    """Oracle handoff HND→NHD == raw landing + vLLM kv_postprocess_layout_on_receive."""
    g = np.load(os.path.join(golden_dir, "layout_hnd_to_nhd.npz"))
    L, H, N, D, nb_p, nb_d, n = [int(x) for x in g["shape"]]
    src = _pool_from_regions(g["p_pool"], o.HND, H, D, N)
    dst = _pool_from_regions(np.full((2 * L, nb_d, N, H, D), o.CANARY16, dtype=np.uint16), o.NHD, H, D, N)
    o.handoff(src, g["p_table"], dst, g["d_table"])
    want = np.ascontiguousarray(g["d_expect"]).view(np.uint8).reshape(-1)
    assert np.array_equal(dst.buf, want)
    # and the C restatement
    dst2 = _pool_from_regions(np.full((2 * L, nb_d, N, H, D), o.CANARY16, dtype=np.uint16), o.NHD, H, D, N)
    assert oc.handoff(src, g["p_table"], dst2, g["d_table"], threads=3) == 0
    assert np.array_equal(dst2.buf, want)


def test_golden_blocksize_ratio(golden_dir):
    """Oracle regroup (4-token P blocks → 16-token D blocks) == raw landing + vLLM's
    kv_postprocess_blksize_on_receive (HND kept) / kv_postprocess_blksize_and_layout_on_receive (→ NHD)."""
    g = np.load(os.path.join(golden_dir, "blocksize_ratio.npz"))
    L, H, D, Np, ratio, nb_p, nb_d, n_p = [int(x) for x in g["shape"]]
    src = _pool_from_regions(g["p_pool"], o.HND, H, D, Np)
    for name, lay in (("hnd", o.HND), ("nhd", o.NHD)):
        want = np.ascontiguousarray(g["d_expect_" + name]).view(np.uint8).reshape(-1)
        for impl in ("numpy", "c"):
            dst = _pool_from_regions(np.full((2 * L, nb_d, Np * ratio * H * D), o.CANARY16, dtype=np.uint16), lay, H, D, Np * ratio)
            if impl == "numpy":
                o.handoff(src, g["p_table"], dst, g["d_table"])
            else:
                assert oc.handoff(src, g["p_table"], dst, g["d_table"], threads=2) == 0
            assert np.array_equal(dst.buf, want), (name, impl)


@pytest.mark.parametrize("Ns,Nd", [(16, 64), (64, 16), (16, 32), (32, 16)])
@pytest.mark.parametrize("layouts", [(o.HND, o.HND), (o.HND, o.NHD), (o.NHD, o.NHD), (o.NHD, o.HND)])
def test_regroup_numpy_vs_c_and_round_trip(Ns, Nd, layouts):
    ls, ld = layouts
    N = min(Ns, Nd)
    n = 2 * (max(Ns, Nd) // N) + 1  # a partially filled last big block
    ns, nd = -(-n // (Ns // N)), -(-n // (Nd // N))
    src = o.make_pool(1, 4, 32, ns + 2, ls, block_tokens=Ns, seed=8)
    d1 = o.make_pool(1, 4, 32, nd + 3, ld, block_tokens=Nd, fill="canary")
    d2 = o.make_pool(1, 4, 32, nd + 3, ld, block_tokens=Nd, fill="canary")
    ts, td = o.make_table(ns + 2, ns, 1), o.make_table(nd + 3, nd, 2)
    o.handoff_regroup(src, ts, d1, td, n_small=n)
    s, d = oc._paged(src, ts, 0, keep := []), oc._paged(d2, td, 0, keep)
    import ctypes
    assert oc.lib().kvxo_xfer(ctypes.byref(s), ctypes.byref(d), 2, n, N, 4, 64, 2) == 0
    assert np.array_equal(d1.buf, d2.buf)
    # and back again into a fresh source-shaped pool: the moved granules are restored bit for bit
    back = o.make_pool(1, 4, 32, ns + 2, ls, block_tokens=Ns, fill="canary")
    o.handoff_regroup(d1, td, back, ts, n_small=n)
    for r in range(2):
        a = o._rows(src.region_view(r)[ts], ls, Ns, 4, 64).reshape(-1, 4, 64)[: n * N]
        b = o._rows(back.region_view(r)[ts], ls, Ns, 4, 64).reshape(-1, 4, 64)[: n * N]
        assert np.array_equal(a, b)


def test_golden_tp_mapping(golden_dir):
    rows = json.load(open(os.path.join(golden_dir, "tp_mapping.json")))["rows"]
    assert len(rows) >= 30
    for row in rows:
        m = o.tp_mapping(row["d_tp"], row["p_tp"], row["d_rank"], row["total_kv_heads"])
        assert list(m.source_ranks) == row["source_ranks"], row
        assert {str(k): v for k, v in m.rank_to_slot.items()} == row["rank_to_slot"], row
        assert m.rank_offset_factor == row["rank_offset_factor"], row


def test_survey_known_answers():
    """SURVEY.md §8c table."""
    m = o.tp_mapping(2, 4, 1, 8)
    assert m.tp_ratio == -2 and m.source_ranks == (2, 3) and m.rank_to_slot == {2: 0, 3: 1}
    m = o.tp_mapping(4, 2, 3, 8)
    assert m.tp_ratio == 2 and m.source_ranks == (1,) and m.rank_offset_factor == 1
    m = o.tp_mapping(1, 4, 0, 8)
    assert m.tp_ratio == -4 and m.rank_to_slot == {0: 0, 1: 1, 2: 2, 3: 3}


CASES = [
    # (layers, heads, head_dim, src_layout, dst_layout, arrangement_src, arrangement_dst)
    (2, 8, 64, o.HND, o.NHD, "split", "split"),
    (2, 8, 64, o.NHD, o.HND, "split", "blocks_first"),
    (1, 4, 128, o.NHD, o.NHD, "blocks_first", "split"),
    (3, 2, 32, o.HND, o.HND, "split", "split"),
]


@pytest.mark.parametrize("case", CASES)
def test_numpy_vs_c_and_canary(case):
    L, H, D, ls, ld, as_, ad = case
    src = o.make_pool(L, H, D, 23, ls, arrangement=as_, seed=1)
    dst = o.make_pool(L, H, D, 29, ld, arrangement=ad, fill="canary")
    dst_c = o.make_pool(L, H, D, 29, ld, arrangement=ad, fill="canary")
    ts, td = o.make_table(23, 9, 5), o.make_table(29, 9, 6)
    o.handoff(src, ts, dst, td)
    assert oc.handoff(src, ts, dst_c, td, threads=4) == 0
    assert np.array_equal(dst.buf, dst_c.buf)
    # untouched blocks keep the canary; touched blocks hold the source rows
    canary = o.make_pool(L, H, D, 29, ld, arrangement=ad, fill="canary")
    touched = np.zeros(29, bool)
    touched[td] = True
    for r in range(dst.regions):
        v, c = dst.region_view(r), canary.region_view(r)
        assert np.array_equal(v[~touched], c[~touched])
        a = o._rows(src.region_view(r)[ts], ls, 16, H, D * 2)
        b = o._rows(v[td], ld, 16, H, D * 2)
        assert np.array_equal(a, b)


def test_gather_scatter_compose_equals_handoff():
    src = o.make_pool(2, 8, 64, 17, o.HND, seed=3)
    d1 = o.make_pool(2, 8, 64, 19, o.NHD, fill="canary")
    d2 = o.make_pool(2, 8, 64, 19, o.NHD, fill="canary")
    ts, td = o.make_table(17, 7, 1), o.make_table(19, 7, 2)
    for stage_layout in (o.NHD, o.HND):
        img = o.gather(src, ts, o.Xform(stage_layout=stage_layout))
        img_c = oc.gather(src, ts, o.Xform(stage_layout=stage_layout), threads=2)
        assert np.array_equal(img, img_c)
        o.scatter(o.copy(img), d1, td, o.Xform(stage_layout=stage_layout))
    o.handoff(src, ts, d2, td)
    assert np.array_equal(d1.buf, d2.buf)


def test_round_trip_is_identity():
    """encode → decode: P(HND) → D(NHD) → P'(HND) restores the request's blocks bit for bit."""
    src = o.make_pool(2, 8, 128, 13, o.HND, seed=9)
    mid = o.make_pool(2, 8, 128, 13, o.NHD, fill="canary")
    back = o.make_pool(2, 8, 128, 13, o.HND, fill="canary")
    t = o.make_table(13, 12, 4)
    o.handoff(src, t, mid, t)
    o.handoff(mid, t, back, t)
    for r in range(src.regions):
        assert np.array_equal(src.region_view(r)[t], back.region_view(r)[t])


def test_prefix_hits():
    src = o.make_pool(1, 2, 32, 9, o.NHD, seed=2)
    dst = o.make_pool(1, 2, 32, 9, o.NHD, fill="canary")
    before = dst.buf.copy()
    o.handoff(src, [1, 2, 3], dst, [])  # full prefix hit: nothing moves
    assert np.array_equal(dst.buf, before)
    o.handoff(src, [1, 2, 3, 4], dst, [7, 8])  # partial: last two source blocks
    for r in range(2):
        assert np.array_equal(dst.region_view(r)[[7, 8]], src.region_view(r)[[3, 4]])


def test_bad_block_ids_raise():
    src = o.make_pool(1, 2, 32, 5, o.NHD, seed=2)
    with pytest.raises(IndexError):
        o.gather(src, [1, 5])
    with pytest.raises(IndexError):
        o.gather(src, [-1])


@pytest.mark.parametrize("p_tp,d_tp", [(4, 2), (2, 4), (4, 1), (1, 2), (2, 2)])
@pytest.mark.parametrize("layouts", [(o.HND, o.HND), (o.HND, o.NHD), (o.NHD, o.NHD), (o.NHD, o.HND)])
def test_tp_reshard_against_global_truth(p_tp, d_tp, layouts):
    """Build a global KV [regions, blocks, N, H_total, R]; shard heads over P ranks; after all edges
    every D rank's blocks equal its own head shard of the global truth (SURVEY.md §8c identity)."""
    lp, ld = layouts
    L, Ht, D, N, nb, n = 2, 8, 32, 16, 11, 6
    R = D * 2
    rng = np.random.Generator(np.random.PCG64(77))
    truth = rng.integers(0, 256, size=(2 * L, nb, N, Ht, R), dtype=np.uint8)
    hp, hd = Ht // p_tp, Ht // d_tp
    t_src = o.make_table(nb, n, 1)
    p_pools = []
    for r in range(p_tp):
        pool = o.make_pool(L, hp, D, nb, lp, fill="canary", tp_rank=r, tp_size=p_tp)
        for reg in range(2 * L):
            pool.region_view(reg)[:] = o._pack(truth[reg][:, :, r * hp : (r + 1) * hp, :], lp)
        p_pools.append(pool)
    for j in range(d_tp):
        dpool = o.make_pool(L, hd, D, nb + 2, ld, fill="canary", tp_rank=j, tp_size=d_tp)
        dpool_c = o.make_pool(L, hd, D, nb + 2, ld, fill="canary", tp_rank=j, tp_size=d_tp)
        t_dst = o.make_table(nb + 2, n, 100 + j)
        m = o.tp_mapping(d_tp, p_tp, j, Ht)
        for e in m.edges:
            o.handoff(p_pools[e.p_rank], t_src, dpool, t_dst, e.xform)
            assert oc.handoff(p_pools[e.p_rank], t_src, dpool_c, t_dst, e.xform, threads=2) == 0
        for reg in range(2 * L):
            got = o._rows(dpool.region_view(reg)[t_dst], ld, N, hd, R)
            assert np.array_equal(got, truth[reg][t_src][:, :, j * hd : (j + 1) * hd, :])
        assert np.array_equal(dpool.buf, dpool_c.buf)


@settings(max_examples=40, deadline=None)
@given(
    L=st.integers(1, 3), H=st.sampled_from([1, 2, 4, 8]), D=st.sampled_from([8, 16, 64]),
    N=st.sampled_from([1, 4, 16]), n=st.integers(0, 7), ls=st.sampled_from([0, 1]), ld=st.sampled_from([0, 1]),
    so=st.integers(0, 3), do=st.integers(0, 3), hc=st.integers(1, 4), seed=st.integers(0, 2**16),
)
def test_property_numpy_equals_c(L, H, D, N, n, ls, ld, so, do, hc, seed):
    Hs, Hd = H + so, H + do  # source / destination chunks with room for the offsets
    hc = min(hc, H)
    src = o.make_pool(L, Hs, D, 9, ls, block_tokens=N, seed=seed)
    d1 = o.make_pool(L, Hd, D, 10, ld, block_tokens=N, fill="canary")
    d2 = o.make_pool(L, Hd, D, 10, ld, block_tokens=N, fill="canary")
    ts, td = o.make_table(9, n, seed), o.make_table(10, n, seed + 1)
    xf = o.Xform(so, do, hc)
    o.handoff(src, ts, d1, td, xf)
    assert oc.handoff(src, ts, d2, td, xf, threads=2) == 0
    assert np.array_equal(d1.buf, d2.buf)


def test_torch_cpu_twin_agrees_with_numpy_oracle():
    """A second, independent restatement: the receive path the reference's images run today —
    land the raw blocks, then index_select -> permute -> index_copy_ per region on torch.bfloat16 tensors
    ([vllm] kv_connector/utils.py:255-277) — against the NumPy oracle on random bit patterns (NaN / Inf /
    denormal bf16 included), TP head slot included.  Equality is on the integer views."""
    import torch

    L, H, N, D, nb_p, nb_d, n = 2, 8, 16, 64, 23, 29, 17
    src = o.make_pool(L, H, D, nb_p, o.HND, seed=91)
    dst = o.make_pool(L, H, D, nb_d, o.NHD, fill="canary")
    ts, td = o.make_table(nb_p, n, 1), o.make_table(nb_d, n, 2)
    o.handoff(src, ts, dst, td)
    p = torch.from_numpy(src.buf.copy()).view(torch.bfloat16).view(2 * L, nb_p, H, N, D)            # HND pools
    d = torch.from_numpy(o.make_pool(L, H, D, nb_d, o.NHD, fill="canary").buf).view(torch.bfloat16).view(2 * L, nb_d, N, H, D)
    si, di = torch.from_numpy(ts.astype(np.int64)), torch.from_numpy(td.astype(np.int64))
    for r in range(2 * L):
        blocks = p[r].index_select(0, si)                       # [n, H, N, D]
        d[r].index_copy_(0, di, blocks.permute(0, 2, 1, 3).contiguous())
    assert np.array_equal(d.contiguous().view(torch.int16).numpy().view(np.uint8).reshape(-1), dst.buf)
    # TP4 -> TP2 on the decode side: two 2-head sources land in head slots 0 and 2 of a 4-head NHD chunk
    srcs = [o.make_pool(L, 2, D, nb_p, o.HND, seed=5 + r) for r in range(2)]
    dst2 = o.make_pool(L, 4, D, nb_d, o.NHD, fill="canary")
    d2 = torch.from_numpy(dst2.buf.copy()).view(torch.bfloat16).view(2 * L, nb_d, N, 4, D)
    for s, pool in enumerate(srcs):
        o.handoff(pool, ts, dst2, td, o.Xform(0, 2 * s, 2))
        ps = torch.from_numpy(pool.buf.copy()).view(torch.bfloat16).view(2 * L, nb_p, 2, N, D)
        for r in range(2 * L):
            got = d2[r].index_select(0, di)
            got[:, :, 2 * s:2 * s + 2] = ps[r].index_select(0, si).permute(0, 2, 1, 3)
            d2[r].index_copy_(0, di, got)
    assert np.array_equal(d2.contiguous().view(torch.int16).numpy().view(np.uint8).reshape(-1), dst2.buf)


@pytest.mark.skipif(not os.environ.get("KVX_TEST_VLLM"), reason="set KVX_TEST_VLLM=1 to import the on-box vLLM (about 30 s)")
def test_live_vllm_helper_agrees_with_oracle():
    """The helper the golden fixtures were generated from, called live (only where vLLM is installed)."""
    torch = pytest.importorskip("torch")
    utils = pytest.importorskip("vllm.distributed.kv_transfer.kv_connector.utils")
    L, H, N, D, nb_p, nb_d, n = 1, 4, 16, 8, 9, 11, 6
    src = o.make_pool(L, H, D, nb_p, o.HND, seed=7)
    dst = o.make_pool(L, H, D, nb_d, o.NHD, fill="canary")
    ts, td = o.make_table(nb_p, n, 1), o.make_table(nb_d, n, 2)
    o.handoff(src, ts, dst, td)
    raw = o.make_pool(L, H, D, nb_d, o.NHD, fill="canary")
    for r in range(2 * L):
        cache = torch.from_numpy(raw.region_view(r).view(np.int16).reshape(nb_d, N, H, D))
        cache.view(nb_d, -1)[torch.from_numpy(td.astype(np.int64))] = torch.from_numpy(
            src.region_view(r).view(np.int16)[ts.astype(np.int64)].copy())
        utils.kv_postprocess_layout_on_receive(cache, torch.from_numpy(td.astype(np.int64)))
    assert np.array_equal(raw.buf, dst.buf)


def test_baseline_config1_exact_size_numpy_vs_c():
    """BASELINE.json configs[0]: Llama-3.2-1B KV shape (16 layers, 8 KV heads, d=64, bf16, 16-token blocks), one
    2k-token prompt = 128 blocks x 32 regions x 16 KiB = 64 MiB, prefill HND -> decode NHD and identity; the two
    CPU restatements (NumPy, C with 1 and 4 threads) agree on every byte of the decode pool."""
    L, H, D, n = 16, 8, 64, 128
    nb = 4 * n + 1
    src = o.make_pool(L, H, D, nb, o.HND, seed=0xD15A66 + 1000)
    ts, td = o.make_table(nb, n, 0xD15A66 + 1007), o.make_table(nb, n, 0xD15A66 + 1013)
    assert n * src.regions * src.chunk_bytes == 64 << 20
    for layout in (o.NHD, o.HND):
        want = o.make_pool(L, H, D, nb, layout, fill="canary")
        o.handoff(src, ts, want, td)
        for threads in (1, 4):
            got = o.make_pool(L, H, D, nb, layout, fill="canary")
            assert oc.handoff(src, ts, got, td, threads=threads) == 0
            assert np.array_equal(got.buf, want.buf)
        untouched = np.ones(nb, dtype=bool)
        untouched[td] = False
        assert (want.region_view(0)[untouched].view(np.uint16) == o.CANARY16).all()


# ---- addressing pinned to the on-box connector's own descriptor code (tests/golden/addressing.json) ----------------
def _addressing(golden_dir):
    import json

    with open(os.path.join(golden_dir, "addressing.json")) as f:
        return json.load(f)


def _geom_pool(L, H, D, nb, N, E, arrangement):
    """A pool record with the fixture's geometry and no contents (only offsets matter here)."""
    p = o.make_pool(L, H, D, 2, o.HND, block_tokens=N, elem_bytes=E, arrangement=arrangement, fill="canary")
    chunk = N * H * D * E
    if arrangement == "split":
        p.region_off = [r * nb * chunk for r in range(2 * L)]
    else:
        p.region_off = [l * nb * 2 * chunk + kv * chunk for l in range(L) for kv in range(2)]
    p.num_blocks = nb
    return p


def test_golden_descriptor_addressing(golden_dir):
    """Which bytes a request touches in a pool: descriptor ids and addresses as [vllm] _compute_desc_ids and
    _build_fa_local produce them, split K/V pools and blocks-first joint pools (K and V as two virtual regions)."""
    for row in _addressing(golden_dir)["local"]:
        pool = _geom_pool(row["layers"], row["heads"], row["head_dim"], row["num_blocks"], row["block_tokens"], row["elem_bytes"],
                          row["arrangement"])
        ids = row["block_ids"]
        assert row["desc_ids"] == [r * row["num_blocks"] + b for r in range(pool.regions) for b in ids]
        assert [list(t) for t in o.touched(pool, ids)] == row["touched"], row["arrangement"]


def test_golden_heterogeneous_tp_byte_offsets(golden_dir):
    """Heterogeneous TP in bytes: what a decode rank reads inside every prefill block ([vllm] _build_fa_remote) and where
    each source rank's heads land inside its own blocks (_build_local_splits_from_plan) — against the oracle's head
    slices driven by the product's edge planner (kvx/tp.py)."""
    from kvx.tp import plan_edges

    rows = _addressing(golden_dir)["hetero_tp"]
    assert len(rows) == 28
    for row in rows:
        H, N, D, E, L = row["total_kv_heads"], row["block_tokens"], row["head_dim"], row["elem_bytes"], row["layers"]
        arr = "blocks_first" if row["blocks_first"] else "split"
        hp, hd = H // row["p_tp"], H // row["d_tp"]
        P = _geom_pool(L, hp, D, row["nb_p"], N, E, arr)
        Dp = _geom_pool(L, hd, D, row["nb_d"], N, E, arr)
        edges = [e for e in plan_edges(row["p_tp"], row["d_tp"], H) if e.d_rank == row["d_rank"]]
        assert [e.p_rank for e in edges] == row["source_ranks"]
        for k, e in enumerate(edges):
            # every prefill rank of the request has the same geometry: one remote list serves them all
            assert [list(t) for t in o.touched(P, range(row["nb_p"]), e.src_head_off, e.heads)] == row["remote"], row
            assert [list(t) for t in o.touched(Dp, range(row["nb_d"]), e.dst_head_off, e.heads)] == row["local_per_source"][k], row


def test_golden_prefix_hit_alignment(golden_dir):
    """Prefix-cache hit on the decode side: the LAST len(local) remote blocks are read ([vllm] _apply_prefix_caching) —
    the product's align_tables and the oracle's handoff must pick the same ones."""
    from kvx.workers import align_tables

    for row in _addressing(golden_dir)["prefix"]:
        s, d = align_tables(np.array(row["remote"], np.int32), np.array(row["local"], np.int32))
        assert s.tolist() == row["remote_read"] and d.tolist() == row["local_written"]
        src = o.make_pool(1, 2, 8, 42, o.NHD, seed=1)
        dst = o.make_pool(1, 2, 8, 42, o.NHD, fill="canary")
        want = o.make_pool(1, 2, 8, 42, o.NHD, fill="canary")
        o.handoff(src, row["remote"], dst, row["local"])
        if row["local"]:
            o.handoff(src, row["remote_read"], want, row["local_written"])
        assert np.array_equal(dst.buf, want.buf)
