"""Parity tests proper: libkvx.so (through the C ABI) against the oracle, bit for bit."""
import os

import numpy as np
import pytest
import torch

import kvx
from gpu_util import ENGINE_CONFIGS, DevPool, cfg_id, dev_table, xf_c
from oracle import kvx_oracle as o

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    c = kvx.Context(0)
    yield c
    c.close()


@pytest.fixture(params=ENGINE_CONFIGS, ids=cfg_id)
def cfg_ctx(request):
    c = kvx.Context(0, **request.param)
    yield c
    c.check()
    c.close()


SHAPES = [
    # L, H, D, src layout, dst layout, src arrangement, dst arrangement, n
    (2, 8, 64, o.NHD, o.NHD, "split", "split", 37),       # Llama-3.2-1B chunk, identity
    (2, 8, 128, o.HND, o.NHD, "split", "split", 70),      # vLLM P (HND) → NHD decode
    (2, 8, 128, o.NHD, o.HND, "split", "blocks_first", 33),
    (1, 8, 128, o.HND, o.HND, "blocks_first", "split", 5),
    (3, 2, 128, o.HND, o.NHD, "split", "split", 1),        # TP4 shard chunk (8 KiB)
    (1, 4, 32, o.NHD, o.HND, "split", "split", 300),       # many small chunks, > one table batch
]


@pytest.mark.parametrize("shape", SHAPES, ids=lambda s: f"L{s[0]}H{s[1]}D{s[2]}-{s[3]}{s[4]}-{s[5][0]}{s[6][0]}-n{s[7]}")
def test_fused_handoff_matches_oracle(cfg_ctx, shape):
    L, H, D, ls, ld, as_, ad, n = shape
    nbp, nbd = n + 9, n + 14
    src = o.make_pool(L, H, D, nbp, ls, arrangement=as_, seed=11)
    dst = o.make_pool(L, H, D, nbd, ld, arrangement=ad, fill="canary")
    ts, td = o.make_table(nbp, n, 3), o.make_table(nbd, n, 4)
    dsrc, ddst = DevPool(cfg_ctx, src), DevPool(cfg_ctx, dst)
    cfg_ctx.handoff_fused(dsrc.kp, dev_table(ts), ddst.kp, dev_table(td), n)
    torch.cuda.synchronize()
    o.handoff(src, ts, dst, td)
    assert np.array_equal(ddst.bytes(), dst.buf)          # touched == expected, untouched == canary
    assert np.array_equal(dsrc.bytes(), src.buf)          # source unchanged


@pytest.mark.parametrize("stage_layout", [o.NHD, o.HND])
@pytest.mark.parametrize("shape", SHAPES[:4], ids=lambda s: f"L{s[0]}H{s[1]}D{s[2]}-{s[3]}{s[4]}")
def test_gather_copy_scatter_matches_oracle(cfg_ctx, shape, stage_layout):
    L, H, D, ls, ld, as_, ad, n = shape
    nbp, nbd = n + 5, n + 6
    src = o.make_pool(L, H, D, nbp, ls, arrangement=as_, seed=21)
    dst = o.make_pool(L, H, D, nbd, ld, arrangement=ad, fill="canary")
    ts, td = o.make_table(nbp, n, 8), o.make_table(nbd, n, 9)
    dsrc, ddst = DevPool(cfg_ctx, src), DevPool(cfg_ctx, dst)
    img_bytes = src.regions * n * src.chunk_bytes
    stage_p = torch.full((img_bytes,), 0xAB, dtype=torch.uint8, device="cuda")
    stage_d = torch.full((img_bytes + 64,), 0xCD, dtype=torch.uint8, device="cuda")
    xf = kvx.Xform(0, 0, 0, stage_layout)
    cfg_ctx.gather(dsrc.kp, dev_table(ts), n, stage_p, xf)
    flags, _ = cfg_ctx.flags()
    cfg_ctx.flag_reset(flags + 8 * 5, 0)
    cfg_ctx.copy(stage_p, stage_d, img_bytes, flag=flags + 8 * 5, flag_val=42)
    cfg_ctx.scatter(stage_d, ddst.kp, dev_table(td), n, xf)
    torch.cuda.synchronize()
    want_img = o.gather(src, ts, o.Xform(stage_layout=stage_layout))
    assert np.array_equal(stage_p.cpu().numpy(), want_img.reshape(-1))
    got_d = stage_d.cpu().numpy()
    assert np.array_equal(got_d[:img_bytes], want_img.reshape(-1))
    assert (got_d[img_bytes:] == 0xCD).all()              # copy wrote exactly `bytes`
    assert cfg_ctx.flag_poll(flags + 8 * 5, 42)
    o.scatter(want_img, dst, td, o.Xform(stage_layout=stage_layout))
    assert np.array_equal(ddst.bytes(), dst.buf)


@pytest.mark.parametrize("layouts", [(o.HND, o.HND), (o.HND, o.NHD), (o.NHD, o.NHD), (o.NHD, o.HND)],
                         ids=["HH", "HN", "NN", "NH"])
@pytest.mark.parametrize("tp", [(4, 2), (2, 4), (1, 2), (4, 1)], ids=lambda t: f"P{t[0]}D{t[1]}")
def test_tp_reshard_edges(cfg_ctx, tp, layouts):
    """Config #4 in small: every P rank → D rank edge with its head slice, fan-in into one D pool."""
    p_tp, d_tp = tp
    lp, ld = layouts
    L, Ht, D, n = 2, 8, 128, 19
    hp, hd = Ht // p_tp, Ht // d_tp
    ts = o.make_table(n + 3, n, 31)
    p_pools = [o.make_pool(L, hp, D, n + 3, lp, seed=100 + r, tp_rank=r, tp_size=p_tp) for r in range(p_tp)]
    dev_p = [DevPool(cfg_ctx, p) for p in p_pools]
    for j in range(d_tp):
        dpool = o.make_pool(L, hd, D, n + 7, ld, fill="canary", tp_rank=j, tp_size=d_tp)
        ddev = DevPool(cfg_ctx, dpool)
        td = o.make_table(n + 7, n, 200 + j)
        for e in o.tp_mapping(d_tp, p_tp, j, Ht).edges:
            o.handoff(p_pools[e.p_rank], ts, dpool, td, e.xform)
            cfg_ctx.handoff_fused(dev_p[e.p_rank].kp, dev_table(ts), ddev.kp, dev_table(td), n, xf_c(e.xform, ld))
        torch.cuda.synchronize()
        assert np.array_equal(ddev.bytes(), dpool.buf)


def test_tp_reshard_staged(ctx):
    """gather(head slice) → copy → scatter(head slot) for TP4→TP2 equals the oracle."""
    L, D, n = 2, 128, 11
    p_pools = [o.make_pool(L, 2, D, n + 2, o.HND, seed=5 + r) for r in range(2)]
    dpool = o.make_pool(L, 4, D, n + 4, o.NHD, fill="canary")
    ddev = DevPool(ctx, dpool)
    ts, td = o.make_table(n + 2, n, 1), o.make_table(n + 4, n, 2)
    for r, pp in enumerate(p_pools):
        xf = o.Xform(0, 2 * r, 2, o.NHD)
        dp = DevPool(ctx, pp)
        img = torch.empty(pp.regions * n * pp.chunk_bytes, dtype=torch.uint8, device="cuda")
        ctx.gather(dp.kp, dev_table(ts), n, img, xf_c(xf, o.NHD))
        ctx.scatter(img, ddev.kp, dev_table(td), n, xf_c(xf, o.NHD))
        torch.cuda.synchronize()
        want = o.gather(pp, ts, xf)
        assert np.array_equal(img.cpu().numpy(), want.reshape(-1))
        o.scatter(want, dpool, td, xf)
    assert np.array_equal(ddev.bytes(), dpool.buf)


REGROUP = [
    # Ns, Nd, src layout, dst layout, H_src, H_dst, (src_off, dst_off, heads), granules
    (16, 64, o.HND, o.HND, 8, 8, None, 23),
    (16, 32, o.HND, o.NHD, 8, 8, None, 16),
    (16, 64, o.NHD, o.NHD, 8, 8, None, 9),
    (16, 32, o.NHD, o.HND, 8, 8, None, 31),
    (64, 16, o.HND, o.NHD, 8, 8, None, 20),     # big prefill blocks into small decode blocks
    (32, 16, o.NHD, o.HND, 4, 4, None, 7),
    (16, 64, o.HND, o.NHD, 2, 4, (0, 2, 2), 12),  # regroup + TP slot in one pass
    (16, 32, o.NHD, o.HND, 4, 2, (2, 0, 2), 10),
]


@pytest.mark.parametrize("case", REGROUP, ids=lambda c: f"N{c[0]}to{c[1]}-{c[2]}{c[3]}-H{c[4]}to{c[5]}-n{c[7]}")
def test_block_size_regroup_matches_oracle(cfg_ctx, case):
    Ns, Nd, ls, ld, Hs, Hd, sl, n = case
    N = min(Ns, Nd)
    ks, kd = Ns // N, Nd // N
    ns, nd = -(-n // ks), -(-n // kd)
    src = o.make_pool(2, Hs, 128, ns + 5, ls, block_tokens=Ns, seed=61)
    dst = o.make_pool(2, Hd, 128, nd + 6, ld, block_tokens=Nd, fill="canary")
    ts, td = o.make_table(ns + 5, ns, 1), o.make_table(nd + 6, nd, 2)
    xf_o = o.Xform(*sl) if sl else None
    xf_k = kvx.Xform(sl[0], sl[1], sl[2], ld) if sl else None
    dsrc, ddst = DevPool(cfg_ctx, src), DevPool(cfg_ctx, dst)
    cfg_ctx.handoff_fused(dsrc.kp, dev_table(ts), ddst.kp, dev_table(td), n, xf_k)
    torch.cuda.synchronize()
    o.handoff_regroup(src, ts, dst, td, xf_o, n_small=n)
    assert np.array_equal(ddst.bytes(), dst.buf)
    # the host-table entry uploads ceil(n/k) ids per side
    ddst2 = DevPool(cfg_ctx, o.make_pool(2, Hd, 128, nd + 6, ld, block_tokens=Nd, fill="canary"))
    cfg_ctx.handoff_host(dsrc.kp, ts, ddst2.kp, td, n, xf_k)
    torch.cuda.synchronize()
    assert np.array_equal(ddst2.bytes(), dst.buf)


def test_golden_blocksize_fixture_on_gpu(cfg_ctx, golden_dir):
    """vLLM's kv_postprocess_blksize[_and_layout]_on_receive outputs, reproduced in flight."""
    from test_oracle import _pool_from_regions

    g = np.load(os.path.join(golden_dir, "blocksize_ratio.npz"))
    L, H, D, Np, ratio, nb_p, nb_d, n_p = [int(x) for x in g["shape"]]
    src = _pool_from_regions(g["p_pool"], o.HND, H, D, Np)
    for name, lay in (("hnd", o.HND), ("nhd", o.NHD)):
        dst = _pool_from_regions(np.full((2 * L, nb_d, Np * ratio * H * D), o.CANARY16, dtype=np.uint16), lay, H, D, Np * ratio)
        dsrc, ddst = DevPool(cfg_ctx, src), DevPool(cfg_ctx, dst)
        cfg_ctx.handoff_fused(dsrc.kp, dev_table(g["p_table"]), ddst.kp, dev_table(g["d_table"]), n_p)
        torch.cuda.synchronize()
        assert np.array_equal(ddst.bytes(), np.ascontiguousarray(g["d_expect_" + name]).view(np.uint8).reshape(-1))


ODD = [
    # L, H, D, E, src layout, dst layout, n   — fp8 KV (E=1) and row sizes that are not powers of two
    (2, 8, 128, 1, o.HND, o.NHD, 21),
    (2, 8, 64, 1, o.NHD, o.HND, 9),
    (1, 4, 96, 2, o.HND, o.NHD, 14),
    (1, 5, 80, 2, o.NHD, o.HND, 6),
    (1, 40, 128, 2, o.HND, o.NHD, 3),   # a 160 KiB chunk: head unit 4 KiB, 10 pieces per chunk
]


@pytest.mark.parametrize("case", ODD, ids=lambda c: f"H{c[1]}D{c[2]}E{c[3]}-{c[4]}{c[5]}")
def test_other_element_and_row_sizes(cfg_ctx, case):
    L, H, D, E, ls, ld, n = case
    src = o.make_pool(L, H, D, n + 3, ls, elem_bytes=E, seed=5)
    dst = o.make_pool(L, H, D, n + 4, ld, elem_bytes=E, fill="canary")
    ts, td = o.make_table(n + 3, n, 1), o.make_table(n + 4, n, 2)
    dsrc, ddst = DevPool(cfg_ctx, src), DevPool(cfg_ctx, dst)
    cfg_ctx.handoff_fused(dsrc.kp, dev_table(ts), ddst.kp, dev_table(td), n)
    torch.cuda.synchronize()
    o.handoff(src, ts, dst, td)
    assert np.array_equal(ddst.bytes(), dst.buf)


def test_golden_fixture_on_gpu(cfg_ctx, golden_dir):
    g = np.load(os.path.join(golden_dir, "layout_hnd_to_nhd.npz"))
    L, H, N, D, nb_p, nb_d, n = [int(x) for x in g["shape"]]
    from test_oracle import _pool_from_regions

    src = _pool_from_regions(g["p_pool"], o.HND, H, D, N)
    dst = _pool_from_regions(np.full((2 * L, nb_d, N, H, D), o.CANARY16, dtype=np.uint16), o.NHD, H, D, N)
    dsrc, ddst = DevPool(cfg_ctx, src), DevPool(cfg_ctx, dst)
    cfg_ctx.handoff_fused(dsrc.kp, dev_table(g["p_table"]), ddst.kp, dev_table(g["d_table"]), n)
    torch.cuda.synchronize()
    assert np.array_equal(ddst.bytes(), np.ascontiguousarray(g["d_expect"]).view(np.uint8).reshape(-1))


def test_golden_addressing_fixture_on_gpu(ctx, golden_dir):
    """The kernel against the connector's own byte ranges (tests/golden/addressing.json, generated from [vllm]
    _build_fa_remote / _build_local_splits_from_plan): for every heterogeneous-TP row, each edge must copy exactly
    remote range -> local range for the request's blocks and nothing else.  The oracle is not involved."""
    import json

    from kvx.tp import plan_edges

    rows = json.load(open(os.path.join(golden_dir, "addressing.json")))["hetero_tp"]
    rng = np.random.Generator(np.random.PCG64(99))
    for row in rows:
        H, N, D, E, L = row["total_kv_heads"], row["block_tokens"], row["head_dim"], row["elem_bytes"], row["layers"]
        arr = "blocks_first" if row["blocks_first"] else "split"
        hp, hd = H // row["p_tp"], H // row["d_tp"]
        edges = [e for e in plan_edges(row["p_tp"], row["d_tp"], H) if e.d_rank == row["d_rank"]]
        dst = o.make_pool(L, hd, D, row["nb_d"], o.HND, block_tokens=N, elem_bytes=E, arrangement=arr, fill="canary")
        want = dst.buf.copy()
        ddst = DevPool(ctx, dst)
        n = 4
        for k, e in enumerate(edges):
            src = o.make_pool(L, hp, D, row["nb_p"], o.HND, block_tokens=N, elem_bytes=E, arrangement=arr, seed=100 + e.p_rank)
            dsrc = DevPool(ctx, src)
            ts = (1 + rng.permutation(row["nb_p"] - 1)[:n]).astype(np.int32)
            td = (1 + rng.permutation(row["nb_d"] - 1)[:n]).astype(np.int32)   # every edge of the request writes the same blocks
            if k:
                td = td_first
            td_first = td
            ctx.handoff_fused(dsrc.kp, dev_table(ts), ddst.kp, dev_table(td), n, kvx.Xform(e.src_head_off, e.dst_head_off, e.heads, kvx.HND))
            torch.cuda.synchronize()
            for r in range(2 * L):
                for i in range(n):
                    so, sl = row["remote"][r * row["nb_p"] + int(ts[i])]
                    do_, dl = row["local_per_source"][k][r * row["nb_d"] + int(td[i])]
                    assert sl == dl
                    want[do_:do_ + dl] = src.buf[so:so + sl]
            dsrc.kp.release()
        assert np.array_equal(ddst.bytes(), want), row
        ddst.kp.release()
    ctx.check()


def test_empty_and_single(ctx):
    src = o.make_pool(1, 8, 64, 4, o.NHD, seed=1)
    dst = o.make_pool(1, 8, 64, 4, o.HND, fill="canary")
    dsrc, ddst = DevPool(ctx, src), DevPool(ctx, dst)
    flags, _ = ctx.flags()
    ctx.flag_reset(flags, 0)
    ctx.handoff_fused(dsrc.kp, None, ddst.kp, None, 0, flag=flags, flag_val=7)  # full prefix hit: notify only
    torch.cuda.synchronize()
    assert ctx.flag_poll(flags, 7)
    assert np.array_equal(ddst.bytes(), dst.buf)
    ctx.handoff_fused(dsrc.kp, dev_table([3]), ddst.kp, dev_table([1]), 1, flag=flags, flag_val=8)
    torch.cuda.synchronize()
    assert ctx.flag_poll(flags, 8) and not ctx.flag_poll(flags, 9)
    o.handoff(src, [3], dst, [1])
    assert np.array_equal(ddst.bytes(), dst.buf)


def test_host_table_entry(ctx):
    """kvx_handoff_host: block ids arrive in host memory; completion comes back to the host."""
    src = o.make_pool(2, 8, 128, 70, o.HND, seed=4)
    dst = o.make_pool(2, 8, 128, 80, o.NHD, fill="canary")
    ts, td = o.make_table(70, 64, 5), o.make_table(80, 64, 6)
    dsrc, ddst = DevPool(ctx, src), DevPool(ctx, dst)
    done = torch.zeros(1, dtype=torch.int64).pin_memory()
    ctx.handoff_host(dsrc.kp, ts, ddst.kp, td, 64, flag_val=99, h_done=done)
    torch.cuda.synchronize()
    assert int(done[0]) == 99
    o.handoff(src, ts, dst, td)
    assert np.array_equal(ddst.bytes(), dst.buf)
    bad = td.copy()
    bad[3] = 80
    with pytest.raises(kvx.KvxError) as e:
        ctx.handoff_host(dsrc.kp, ts, ddst.kp, bad, 64)
    assert e.value.code == kvx.api.ERANGE


@pytest.mark.parametrize("engine", ["ldg", "tma"])
def test_bad_device_table_is_skipped_and_reported(engine):
    c = kvx.Context(0, engine=engine)
    src = o.make_pool(1, 8, 64, 6, o.NHD, seed=1)
    dst = o.make_pool(1, 8, 64, 6, o.NHD, fill="canary")
    dsrc, ddst = DevPool(c, src), DevPool(c, dst)
    c.handoff_fused(dsrc.kp, dev_table([1, 2, 3]), ddst.kp, dev_table([1, 6, 3]), 3)  # id 6 is out of range
    torch.cuda.synchronize()
    with pytest.raises(kvx.KvxError) as e:
        c.check()
    assert e.value.code == kvx.api.ERANGE
    o.handoff(src, [1, 3], dst, [1, 3])  # the bad item moved nothing
    assert np.array_equal(ddst.bytes(), dst.buf)
    c.check()
    c.close()


@pytest.mark.parametrize("nbytes", [16, 4096 + 16, 32768, 3 * 32768 + 48, (1 << 20) + 32, 9 << 20])
def test_copy_sizes(cfg_ctx, nbytes):
    a = torch.randint(0, 256, (nbytes + 32,), dtype=torch.uint8, device="cuda")
    b = torch.full((nbytes + 32,), 0x5A, dtype=torch.uint8, device="cuda")
    cfg_ctx.copy(a[16:], b[16:], nbytes)
    torch.cuda.synchronize()
    assert torch.equal(a[16 : 16 + nbytes], b[16 : 16 + nbytes])
    assert (b[:16] == 0x5A).all() and (b[16 + nbytes :] == 0x5A).all()


def test_argument_errors(ctx):
    src = o.make_pool(1, 8, 64, 4, o.NHD, seed=1)
    dsrc = DevPool(ctx, src)
    with pytest.raises(kvx.KvxError):  # misaligned staging
        ctx.gather(dsrc.kp, dev_table([1]), 1, dsrc.t.data_ptr() + 8)
    with pytest.raises(kvx.KvxError):  # head slice out of range
        ctx.gather(dsrc.kp, dev_table([1]), 1, dsrc.t, kvx.Xform(6, 0, 4, 0))
    with pytest.raises(kvx.KvxError):  # copy size not a multiple of 16
        ctx.copy(dsrc.t, dsrc.t, 24)
    other = o.make_pool(1, 8, 128, 4, o.NHD, seed=1)
    with pytest.raises(kvx.KvxError):  # shapes differ
        ctx.handoff_fused(dsrc.kp, dev_table([1]), DevPool(ctx, other).kp, dev_table([1]), 1)


@pytest.mark.parametrize("cfg", [dict(engine="tma"), dict(engine="ldg")], ids=cfg_id)
def test_full_size_properties_qwen3_32k(cfg):
    """BASELINE config #2 at its largest size (Qwen3-0.6B KV shape, ISL 32k, 3.5 GiB payload), checked
    by size-independent properties on the device: HND→NHD→HND round trip restores the request's
    blocks bit for bit, a word checksum is preserved by each hop, untouched blocks keep the canary."""
    _round_trip_properties(cfg, "Qwen/Qwen3-0.6B", 32768)


def test_full_size_properties_llama8b_32k():
    """BASELINE configs #3 / #5 at their largest request (Llama-3.1-8B KV shape, ISL 32k, 4 GiB payload), the
    same properties on one GPU (the multi-GPU runs verify the NVLink path bit for bit through the harness)."""
    _round_trip_properties({}, "meta-llama/Llama-3.1-8B", 32768)


def _round_trip_properties(cfg, model, isl):
    from kvx import synth

    c = kvx.Context(0, **cfg)
    shape = kvx.KV_SHAPES[model]
    n = shape.blocks_for(isl)
    nb = n + 129
    P = synth.alloc_pool(c, shape, nb, kvx.HND)
    Dp = synth.alloc_pool(c, shape, nb, kvx.NHD)
    P2 = synth.alloc_pool(c, shape, nb, kvx.HND)
    synth.fill_random(P, 1234)
    synth.fill_canary(Dp)
    synth.fill_canary(P2)
    t1, t2, t3 = synth.make_table(nb, n, 1), synth.make_table(nb, n, 2), synth.make_table(nb, n, 3)
    c.handoff_fused(P.pool, dev_table(t1), Dp.pool, dev_table(t2), n)
    c.handoff_fused(Dp.pool, dev_table(t2), P2.pool, dev_table(t3), n)
    torch.cuda.synchronize()
    c.check()
    chunk_words = shape.chunk_bytes() // 2

    def view(dp):
        return dp.tensor().view(torch.int16).view(shape.regions, nb, chunk_words)

    i1, i2, i3 = (torch.from_numpy(t.astype(np.int64)).cuda() for t in (t1, t2, t3))
    a, m, b = view(P)[:, i1], view(Dp)[:, i2], view(P2)[:, i3]
    assert torch.equal(a, b)
    # per-chunk checksum survives the permutation
    assert torch.equal(a.to(torch.int64).sum(-1), m.to(torch.int64).sum(-1))
    # the middle hop really is the NHD image of the HND source
    N, H, Dh = shape.block_tokens, shape.kv_heads, shape.head_dim
    assert torch.equal(a.view(shape.regions, n, H, N, Dh).transpose(2, 3), m.view(shape.regions, n, N, H, Dh))
    canary = np.array(synth.CANARY16, dtype=np.uint16).view(np.int16).item()
    mask = torch.ones(nb, dtype=torch.bool, device="cuda")
    mask[i2] = False
    assert (view(Dp)[:, mask] == canary).all()
    for dp in (P, Dp, P2):
        dp.free()
    c.close()


def test_layerwise_handoff_equals_whole(ctx):
    """Layer-wise transfer (overlap with the tail of prefill): one registered pool per layer group, one
    handoff per group as its KV becomes ready, equals the whole-request handoff."""
    L, H, D, n = 6, 8, 128, 25
    src = o.make_pool(L, H, D, n + 4, o.HND, seed=17)
    whole = o.make_pool(L, H, D, n + 5, o.NHD, fill="canary")
    ts, td = o.make_table(n + 4, n, 1), o.make_table(n + 5, n, 2)
    o.handoff(src, ts, whole, td)
    t_src = torch.from_numpy(src.buf.copy()).cuda()
    t_dst = torch.from_numpy(o.make_pool(L, H, D, n + 5, o.NHD, fill="canary").buf).cuda()
    streams = [torch.cuda.Stream() for _ in range(3)]
    dts, dtd = dev_table(ts), dev_table(td)
    torch.cuda.synchronize()
    pools = []
    for g in range(3):  # three groups of two layers (= four regions) each
        regs = range(4 * g, 4 * g + 4)
        ps = ctx.register_pool([t_src.data_ptr() + src.region_off[r] for r in regs], num_blocks=n + 4, heads=H, head_dim=D, layout=kvx.HND)
        pd = ctx.register_pool([t_dst.data_ptr() + whole.region_off[r] for r in regs], num_blocks=n + 5, heads=H, head_dim=D, layout=kvx.NHD)
        pools += [ps, pd]
        ctx.handoff_fused(ps, dts, pd, dtd, n, stream=streams[g])
    torch.cuda.synchronize()
    assert np.array_equal(t_dst.cpu().numpy(), whole.buf)
    for p in pools:
        p.release()


def test_cuda_graph_capture_and_replay(ctx):
    """The device-table entry makes no allocation and no synchronisation: it can be captured in a CUDA
    graph and replayed with new table CONTENTS (same buffers), as a decode loop would."""
    L, H, D, n = 2, 8, 128, 40
    src = o.make_pool(L, H, D, 3 * n, o.HND, seed=23)
    dst = o.make_pool(L, H, D, 3 * n, o.NHD, fill="canary")
    dsrc, ddst = DevPool(ctx, src), DevPool(ctx, dst)
    ts, td = dev_table(o.make_table(3 * n, n, 1)), dev_table(o.make_table(3 * n, n, 2))
    flags, _ = ctx.flags()
    ctx.flag_reset(flags + 8 * 9, 0)
    ctx.handoff_fused(dsrc.kp, ts, ddst.kp, td, n, flag=flags + 8 * 9, flag_val=1)  # warm-up outside capture
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    with torch.cuda.graph(g, stream=s):
        ctx.handoff_fused(dsrc.kp, ts, ddst.kp, td, n, flag=flags + 8 * 9, flag_val=2, stream=s)
    o.handoff(src, ts.cpu().numpy(), dst, td.cpu().numpy())
    for rep in range(2):  # replay with fresh tables written into the same device buffers
        t1, t2 = o.make_table(3 * n, n, 10 + rep), o.make_table(3 * n, n, 20 + rep)
        ts.copy_(torch.from_numpy(t1)); td.copy_(torch.from_numpy(t2))
        g.replay()
        torch.cuda.synchronize()
        o.handoff(src, t1, dst, t2)
        assert np.array_equal(ddst.bytes(), dst.buf)
        assert ctx.flag_poll(flags + 8 * 9, 2)
    ctx.check()


@pytest.mark.parametrize("pdl", [0, 1], ids=["serial", "pdl"])
@pytest.mark.parametrize("engine", ["tma", "ldg"])
def test_dependent_chain_back_to_back(engine, pdl):
    """Eight handoffs queued back to back, each reading the blocks the previous one wrote, with
    programmatic dependent launch on and off: a kernel that became resident early must still see
    everything its predecessor stored (and the predecessor's completion flag must come first)."""
    c = kvx.Context(0, engine=engine, pdl=pdl)
    L, H, D, n, hops = 2, 8, 128, 24, 8
    nb = n + 7
    layouts = [o.HND, o.NHD]
    pools = [o.make_pool(L, H, D, nb, layouts[k % 2], seed=77, fill="canary" if k else "random") for k in range(hops + 1)]
    tabs = [o.make_table(nb, n, 300 + k) for k in range(hops + 1)]
    devs = [DevPool(c, p) for p in pools]
    dtabs = [dev_table(t) for t in tabs]
    flags, _ = c.flags()
    c.flag_reset(flags + 8 * 3, 0)
    torch.cuda.synchronize()
    for k in range(hops):
        c.handoff_fused(devs[k].kp, dtabs[k], devs[k + 1].kp, dtabs[k + 1], n, flag=flags + 8 * 3, flag_val=k + 1)
    torch.cuda.synchronize()
    assert c.flag_poll(flags + 8 * 3, hops)
    for k in range(hops):
        o.handoff(pools[k], tabs[k], pools[k + 1], tabs[k + 1])
        assert np.array_equal(devs[k + 1].bytes(), pools[k + 1].buf), f"hop {k}"
    c.check()
    c.close()


@pytest.mark.parametrize("n", [1, 63, 64, 65, 255, 256, 257, 1023, 1024, 1025, 3999, 4000, 4001])
@pytest.mark.parametrize("param_tables", [8000, 0], ids=["ptab", "upload"])
def test_host_entry_table_transport(param_tables, n):
    """kvx_handoff_host moves id lists inside the kernel's parameter block (size classes of 128, 512,
    2048 and 8000 ids for both sides together) and uploads longer ones; the completion word lands in
    pinned memory (written by the kernel) or pageable memory (copied back).  Same bytes either way."""
    c = kvx.Context(0, param_tables=param_tables)
    nb = n + 11
    src = o.make_pool(1, 8, 64, nb, o.HND, seed=n)
    dst = o.make_pool(1, 8, 64, nb + 2, o.NHD, fill="canary")
    ts, td = o.make_table(nb, n, 5), o.make_table(nb + 2, n, 6)
    dsrc, ddst = DevPool(c, src), DevPool(c, dst)
    pinned = torch.zeros(2, dtype=torch.int64).pin_memory()
    pageable = np.zeros(2, dtype=np.int64)
    flags, _ = c.flags()
    c.flag_reset(flags + 8 * 7, 0)
    c.handoff_host(dsrc.kp, ts, ddst.kp, td, n, flag=flags + 8 * 7, flag_val=41, h_done=pinned[1:])
    torch.cuda.synchronize()
    assert int(pinned[1]) == 41 and int(pinned[0]) == 0 and c.flag_poll(flags + 8 * 7, 41)
    o.handoff(src, ts, dst, td)
    assert np.array_equal(ddst.bytes(), dst.buf)
    # second request into the same pools: pageable completion word, no device flag
    ts2, td2 = o.make_table(nb, n, 15), o.make_table(nb + 2, n, 16)
    c.handoff_host(dsrc.kp, ts2, ddst.kp, td2, n, flag=None, flag_val=42, h_done=pageable[1:])
    torch.cuda.synchronize()
    assert pageable[1] == 42 and pageable[0] == 0
    o.handoff(src, ts2, dst, td2)
    assert np.array_equal(ddst.bytes(), dst.buf)
    # empty request: notify only, both kinds of completion word
    c.handoff_host(dsrc.kp, None, ddst.kp, None, 0, flag=flags + 8 * 7, flag_val=43, h_done=pinned[:1])
    torch.cuda.synchronize()
    assert int(pinned[0]) == 43 and c.flag_poll(flags + 8 * 7, 43)
    assert np.array_equal(ddst.bytes(), dst.buf)
    bad = td.copy()
    bad[0] = nb + 2
    with pytest.raises(kvx.KvxError) as e:
        c.handoff_host(dsrc.kp, ts, ddst.kp, bad, n)
    assert e.value.code == kvx.api.ERANGE
    c.check()
    c.close()


@pytest.mark.parametrize("param_tables", [8000, 0], ids=["ptab", "upload"])
def test_host_lists_are_consumed_on_return(param_tables):
    """kvx.h: "the lists are consumed before the call returns" — also when they sit in PINNED memory, from which
    a plain cudaMemcpyAsync would read them after the call has returned.  Several requests are queued behind a
    long-running copy and each one's pinned lists are overwritten the moment its call returns."""
    c = kvx.Context(0, param_tables=param_tables)
    nb, n = 300, 64
    src = o.make_pool(2, 8, 64, nb, o.HND, seed=3)
    dst = o.make_pool(2, 8, 64, nb, o.NHD, fill="canary")
    dsrc, ddst = DevPool(c, src), DevPool(c, dst)
    perm_s, perm_d = o.make_table(nb, 4 * n, 21), o.make_table(nb, 4 * n, 22)
    pin_s, pin_d = torch.zeros(n, dtype=torch.int32).pin_memory(), torch.zeros(n, dtype=torch.int32).pin_memory()
    big_a = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    big_b = torch.empty_like(big_a)
    for _ in range(4):  # keep the stream busy so that the handoffs below are still queued when their calls return
        big_b.copy_(big_a)
    for k in range(4):
        pin_s.copy_(torch.from_numpy(perm_s[k * n:(k + 1) * n]))
        pin_d.copy_(torch.from_numpy(perm_d[k * n:(k + 1) * n]))
        c.handoff_host(dsrc.kp, pin_s, ddst.kp, pin_d, n)
        pin_s.fill_(1); pin_d.fill_(1)   # valid ids, wrong blocks: a late read would move block 1 onto block 1
        o.handoff(src, perm_s[k * n:(k + 1) * n], dst, perm_d[k * n:(k + 1) * n])
    torch.cuda.synchronize()
    assert np.array_equal(ddst.bytes(), dst.buf)
    c.check()
    c.close()


def test_table_ring_wraps_under_backlog():
    """Lists too long for the parameter block travel through the context's table ring (4 Mi entries).  Spans are
    reused only after their reader has finished: 12 requests of 2 x 400k ids wrap the ring twice while earlier ones
    are still queued, and every one must move exactly its own blocks."""
    c = kvx.Context(0)
    nb, n = 400_001, 400_000
    src = o.make_pool(1, 1, 8, nb, o.NHD, seed=9, block_tokens=1)     # 16-byte chunks: a big table, few bytes
    dst = o.make_pool(1, 1, 8, nb, o.NHD, fill="canary", block_tokens=1)
    dsrc, ddst = DevPool(c, src), DevPool(c, dst)
    big_a = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    big_b = torch.empty_like(big_a)
    for _ in range(4):
        big_b.copy_(big_a)
    want = dst.buf.copy()
    sv, wv = src.buf.view(np.uint8).reshape(2, nb, 16), want.view(np.uint8).reshape(2, nb, 16)  # [K|V][block][chunk]
    for k in range(12):
        ts, td = o.make_table(nb, n, 100 + k), o.make_table(nb, n, 200 + k)
        c.handoff_host(dsrc.kp, ts, ddst.kp, td, n)
        wv[:, td] = sv[:, ts]
    torch.cuda.synchronize()
    assert np.array_equal(ddst.bytes(), want.view(np.uint8).reshape(-1))
    c.check()
    c.close()


def _cross_layer_nhd_pool(L, H, D, nb, fill, seed=0, N=16, E=2):
    """vLLM's cross-layer block pool in NHD order: one allocation (num_blocks, num_layers, 2, N, H, D)
    ([vllm] v1/attention/backends/flash_attn.py:158-160).  As a kvx pool: region (l, kv) starts at
    (2l + kv) * chunk and every block id advances by the whole cross-layer block."""
    chunk = N * H * D * E
    if fill == "random":
        words = np.random.Generator(np.random.PCG64(seed)).integers(0, 65536, size=nb * 2 * L * chunk // 2, dtype=np.uint16)
    else:
        words = np.full(nb * 2 * L * chunk // 2, o.CANARY16, dtype=np.uint16)
    return o.Pool(words.view(np.uint8).copy(), [r * chunk for r in range(2 * L)], nb, 2 * L * chunk, N, H, D, E, o.NHD)


def test_cross_layer_block_pools(cfg_ctx):
    """Cross-layer block pools on either side, against split per-layer pools on the other, and the
    result checked twice: against the oracle and against numpy indexing of the 6-d array itself."""
    L, H, D, n, nb = 3, 8, 128, 21, 40
    src = _cross_layer_nhd_pool(L, H, D, nb, "random", seed=9)
    dst = o.make_pool(L, H, D, nb + 3, o.HND, fill="canary")
    ts, td = o.make_table(nb, n, 1), o.make_table(nb + 3, n, 2)
    dsrc, ddst = DevPool(cfg_ctx, src), DevPool(cfg_ctx, dst)
    cfg_ctx.handoff_fused(dsrc.kp, dev_table(ts), ddst.kp, dev_table(td), n)
    torch.cuda.synchronize()
    got = ddst.bytes()
    o.handoff(src, ts, dst, td)
    assert np.array_equal(got, dst.buf)
    a6 = src.buf.view(np.uint16).reshape(nb, L, 2, 16, H, D)                    # the engine's own view
    g5 = got.view(np.uint16).reshape(2 * L, nb + 3, H, 16, D)                   # split HND regions
    for l in range(L):
        for kv in range(2):
            assert np.array_equal(g5[2 * l + kv, td], a6[ts, l, kv].transpose(0, 2, 1, 3))
    # and back into a cross-layer pool
    back = _cross_layer_nhd_pool(L, H, D, nb, "canary")
    dback = DevPool(cfg_ctx, back)
    cfg_ctx.handoff_fused(ddst.kp, dev_table(td), dback.kp, dev_table(ts), n)
    torch.cuda.synchronize()
    b6 = dback.bytes().view(np.uint16).reshape(nb, L, 2, 16, H, D)
    assert np.array_equal(b6[ts], a6[ts])
    rest = np.ones(nb, dtype=bool)
    rest[ts] = False
    assert (b6[rest] == o.CANARY16).all()


@pytest.mark.parametrize("N", [16, 64])
def test_mla_latent_pools(cfg_ctx, N):
    """MLA-style caches: one latent vector per token and layer (no separate V, a single 'head' of
    512 + 64 elements) — one region per layer, 1152-byte rows; block sizes 16 and 64."""
    L, Dl, n, nb = 5, 576, 13, 20
    chunk = N * Dl * 2
    words = np.random.Generator(np.random.PCG64(3)).integers(0, 65536, size=L * nb * chunk // 2, dtype=np.uint16)
    src = o.Pool(words.view(np.uint8).copy(), [r * nb * chunk for r in range(L)], nb, chunk, N, 1, Dl, 2, o.NHD)
    dst = o.Pool(np.full(L * (nb + 1) * chunk // 2, o.CANARY16, dtype=np.uint16).view(np.uint8).copy(),
                 [r * (nb + 1) * chunk for r in range(L)], nb + 1, chunk, N, 1, Dl, 2, o.HND)
    ts, td = o.make_table(nb, n, 4), o.make_table(nb + 1, n, 5)
    dsrc, ddst = DevPool(cfg_ctx, src), DevPool(cfg_ctx, dst)
    cfg_ctx.handoff_fused(dsrc.kp, dev_table(ts), ddst.kp, dev_table(td), n)
    torch.cuda.synchronize()
    o.handoff(src, ts, dst, td)
    assert np.array_equal(ddst.bytes(), dst.buf)
    s3 = src.buf.view(np.uint16).reshape(L, nb, N * Dl)
    d3 = ddst.bytes().view(np.uint16).reshape(L, nb + 1, N * Dl)
    assert np.array_equal(d3[:, td], s3[:, ts])          # with one head NHD and HND are the same bytes


@pytest.mark.parametrize("mode", ["one_launch", "upload_fallback", "oversize_fallback", "pageable_fallback"])
def test_batched_requests(mode):
    """kvx_handoff_batch: several requests (one of them empty) between the same pools, each with its own
    release flag and pinned completion word, in one launch — and the same through every fallback."""
    opts = dict(param_tables=0) if mode == "upload_fallback" else {}
    c = kvx.Context(0, **opts)
    counts = [7, 0, 33, 1, 120] if mode != "oversize_fallback" else [7, 0, 2100, 1, 2100]
    nb = sum(counts) + 9
    src = o.make_pool(2, 8, 64, nb, o.HND, seed=3)
    dst = o.make_pool(2, 8, 64, nb + 5, o.NHD, fill="canary")
    ps = o.make_table(nb, sum(counts), 1)
    pd = o.make_table(nb + 5, sum(counts), 2)
    dsrc, ddst = DevPool(c, src), DevPool(c, dst)
    flags, _ = c.flags()
    pinned = torch.zeros(len(counts), dtype=torch.int64).pin_memory()
    pageable = np.zeros(len(counts), dtype=np.int64)
    reqs, at = [], 0
    for k, n in enumerate(counts):
        c.flag_reset(flags + 8 * (20 + k), 0)
        s_ids, d_ids = (ps[at:at + n], pd[at:at + n]) if n else (None, None)
        done = pageable[k:k + 1] if (mode == "pageable_fallback" and k == 2) else pinned[k:k + 1]
        reqs.append((s_ids, d_ids, flags + 8 * (20 + k) if k != 3 else None, 100 + k, done if k != 0 else None))
        at += n
    torch.cuda.synchronize()
    l0 = c.launch_count
    c.handoff_batch(dsrc.kp, ddst.kp, reqs)
    torch.cuda.synchronize()
    launches = c.launch_count - l0
    assert launches == (1 if mode == "one_launch" else sum(1 for n in counts if n))
    o.handoff(src, ps, dst, pd)
    assert np.array_equal(ddst.bytes(), dst.buf)
    for k in range(len(counts)):
        if k != 3:
            assert c.flag_poll(flags + 8 * (20 + k), 100 + k)
        want = 0 if k == 0 else 100 + k
        got = pageable[k] if (mode == "pageable_fallback" and k == 2) else int(pinned[k])
        assert got == want, (k, got)
    bad = [(np.array([1], np.int32), np.array([nb + 5], np.int32), None, 1, None)] * 2
    with pytest.raises(kvx.KvxError) as e:
        c.handoff_batch(dsrc.kp, ddst.kp, bad)
    assert e.value.code == kvx.api.ERANGE
    c.check()
    c.close()


def test_flag_wait_times_out_and_context_recovers(ctx):
    """A transfer that never completes (the peer died) must surface as KVX_ETIMEOUT at the next check,
    never as a hang; the context keeps working afterwards."""
    flags, _ = ctx.flags()
    slot = flags + 8 * 30
    ctx.flag_reset(slot, 0)
    torch.cuda.synchronize()
    ctx.flag_wait(slot, 1, timeout_ms=50)      # nobody will ever write 1
    torch.cuda.synchronize()
    with pytest.raises(kvx.KvxError) as e:
        ctx.check()
    assert e.value.code == kvx.api.ETIMEOUT
    ctx.check()                                 # the error is reported once
    src = o.make_pool(1, 8, 64, 6, o.NHD, seed=2)
    dst = o.make_pool(1, 8, 64, 6, o.HND, fill="canary")
    dsrc, ddst = DevPool(ctx, src), DevPool(ctx, dst)
    ctx.handoff_fused(dsrc.kp, dev_table([2, 4]), ddst.kp, dev_table([5, 1]), 2, flag=slot, flag_val=1)
    ctx.flag_wait(slot, 1, timeout_ms=2000)     # now it arrives (same stream: already there)
    torch.cuda.synchronize()
    ctx.check()
    o.handoff(src, [2, 4], dst, [5, 1])
    assert np.array_equal(ddst.bytes(), dst.buf)


def test_cross_layer_hnd_block_pools(cfg_ctx):
    """vLLM's cross-layer pool in HND order is (num_blocks, H, L, 2, N, D) ([vllm] flash_attn.py:162-165): the
    heads of one (layer, K|V) chunk lie 2L*N*D*E bytes apart.  kvx_pool_desc.head_stride describes it; checked
    against numpy indexing of the engine's own 6-d array, both as source and (with a TP head slot) as destination."""
    L, H, D, N, n, nb = 2, 4, 128, 16, 13, 22
    E = 2
    row, hs, blk = D * E, 2 * L * N * D * E, H * 2 * L * N * D * E
    rng = np.random.Generator(np.random.PCG64(12))
    a6 = rng.integers(0, 65536, size=(nb, H, L, 2, N, D), dtype=np.uint16)
    t_src = torch.from_numpy(a6.view(np.int16)).cuda()
    bases = [t_src.data_ptr() + (2 * l + kv) * N * row for l in range(L) for kv in range(2)]
    src = cfg_ctx.register_pool(bases, num_blocks=nb, heads=H, head_dim=D, layout=kvx.HND, block_stride=blk, head_stride=hs)
    dst = o.make_pool(L, H, D, nb + 2, o.NHD, fill="canary")
    ddst = DevPool(cfg_ctx, dst)
    ts, td = o.make_table(nb, n, 1), o.make_table(nb + 2, n, 2)
    cfg_ctx.handoff_fused(src, dev_table(ts), ddst.kp, dev_table(td), n)
    torch.cuda.synchronize()
    got = ddst.bytes().view(np.uint16).reshape(2 * L, nb + 2, N, H, D)
    for l in range(L):
        for kv in range(2):
            assert np.array_equal(got[2 * l + kv, td], a6[ts, :, l, kv].transpose(0, 2, 1, 3))
    rest = np.ones(nb + 2, dtype=bool)
    rest[td] = False
    assert (got[:, rest] == o.CANARY16).all()
    # destination side: two 2-head split HND sources into head slots 0 and 2 of a 4-head cross-layer HND pool
    b6 = np.full((nb, H, L, 2, N, D), o.CANARY16, dtype=np.uint16)
    t_dst = torch.from_numpy(b6.view(np.int16)).cuda()
    dbases = [t_dst.data_ptr() + (2 * l + kv) * N * row for l in range(L) for kv in range(2)]
    dpool = cfg_ctx.register_pool(dbases, num_blocks=nb, heads=H, head_dim=D, layout=kvx.HND, block_stride=blk, head_stride=hs)
    want = b6.copy()
    for s in range(2):
        half = o.make_pool(L, 2, D, nb + 2, o.HND, seed=40 + s)
        dhalf = DevPool(cfg_ctx, half)
        cfg_ctx.handoff_fused(dhalf.kp, dev_table(td), dpool, dev_table(ts), n, kvx.Xform(0, 2 * s, 2, kvx.HND))
        h5 = half.buf.view(np.uint16).reshape(2 * L, nb + 2, 2, N, D)
        for l in range(L):
            for kv in range(2):
                want[ts, 2 * s:2 * s + 2, l, kv] = h5[2 * l + kv, td]
    torch.cuda.synchronize()
    assert np.array_equal(t_dst.cpu().numpy().view(np.uint16), want)
    cfg_ctx.check()
    src.release(); dpool.release()
    with pytest.raises(kvx.KvxError):   # a head stride on an NHD pool makes no sense
        cfg_ctx.register_pool(bases, num_blocks=nb, heads=H, head_dim=D, layout=kvx.NHD, block_stride=blk, head_stride=hs)


def test_many_streams_many_flags(ctx):
    """160 handoffs from 8 streams into disjoint blocks of one pool, each with its own completion flag:
    the per-launch counter ring, dependent launches and flags must not interfere across streams."""
    L, H, D = 1, 8, 64
    S, PER, nmax = 8, 20, 6
    nb = S * PER * nmax + 1
    src = o.make_pool(L, H, D, nb, o.HND, seed=8)
    dst = o.make_pool(L, H, D, nb, o.NHD, fill="canary")
    dsrc, ddst = DevPool(ctx, src), DevPool(ctx, dst)
    rng = np.random.Generator(np.random.PCG64(4))
    ids_s = 1 + rng.permutation(nb - 1)
    ids_d = 1 + rng.permutation(nb - 1)
    flags, nflags = ctx.flags()
    streams = [torch.cuda.Stream() for _ in range(S)]
    jobs, at = [], 0
    for k in range(S * PER):
        n = int(rng.integers(1, nmax + 1))
        jobs.append((ids_s[at:at + n].astype(np.int32), ids_d[at:at + n].astype(np.int32)))
        ctx.flag_reset(flags + 8 * (100 + k), 0)
        at += n
    tabs = [(dev_table(a), dev_table(b)) for a, b in jobs]
    torch.cuda.synchronize()
    for k, (a, b) in enumerate(jobs):
        st = streams[k % S]
        if k % 3 == 0:   # a third through the host-table entry
            ctx.handoff_host(dsrc.kp, a, ddst.kp, b, len(a), None, flags + 8 * (100 + k), k + 1, None, st)
        else:
            ctx.handoff_fused(dsrc.kp, tabs[k][0], ddst.kp, tabs[k][1], len(a), None, flags + 8 * (100 + k), k + 1, st)
    torch.cuda.synchronize()
    ctx.check()
    for k, (a, b) in enumerate(jobs):
        assert ctx.flag_poll(flags + 8 * (100 + k), k + 1), k
        o.handoff(src, a, dst, b)
    assert np.array_equal(ddst.bytes(), dst.buf)


@pytest.mark.parametrize("dst_layout", [o.NHD, o.HND], ids=["to_nhd", "identity"])
def test_baseline_config1_exact_size(ctx, dst_layout):
    """BASELINE.json configs[0] at its exact size (Llama-3.2-1B KV shape, one 2k-token prompt, 64 MiB): the
    kernel against the oracle, fused and gather -> copy -> scatter."""
    L, H, D, n = 16, 8, 64, 128
    nb = 4 * n + 1
    src = o.make_pool(L, H, D, nb, o.HND, seed=0xD15A66 + 1000)
    want = o.make_pool(L, H, D, nb, dst_layout, fill="canary")
    ts, td = o.make_table(nb, n, 0xD15A66 + 1007), o.make_table(nb, n, 0xD15A66 + 1013)
    o.handoff(src, ts, want, td)
    dsrc = DevPool(ctx, src)
    d1 = DevPool(ctx, o.make_pool(L, H, D, nb, dst_layout, fill="canary"))
    d2 = DevPool(ctx, o.make_pool(L, H, D, nb, dst_layout, fill="canary"))
    ctx.handoff_host(dsrc.kp, ts, d1.kp, td, n)
    img = torch.empty(n * src.regions * src.chunk_bytes, dtype=torch.uint8, device="cuda")
    img2 = torch.empty_like(img)
    xf = kvx.Xform(0, 0, 0, dst_layout)
    ctx.gather(dsrc.kp, dev_table(ts), n, img, xf)
    ctx.copy(img, img2, img.numel())
    ctx.scatter(img2, d2.kp, dev_table(td), n, xf)
    torch.cuda.synchronize()
    ctx.check()
    assert np.array_equal(d1.bytes(), want.buf)
    assert np.array_equal(d2.bytes(), want.buf)


def test_full_size_properties_tp4_to_tp2_llama70b():
    """BASELINE config #4 at full size on one GPU (Llama-3.1-70B KV shape, ISL 8192: two prefill TP4 shards of
    2 heads -> one decode TP2 shard of 4 heads, 640 MiB per edge), checked by size-independent properties:
    every head slot of the decode chunk equals its source shard's chunk (HND -> NHD), and sending the slots back
    (TP2 -> TP4 direction, head slice on the source side) restores both shards bit for bit."""
    from kvx import synth

    c = kvx.Context(0)
    shape = kvx.KV_SHAPES["meta-llama/Llama-3.1-70B"]
    n = shape.blocks_for(8192)
    nb = n + 77
    P = [synth.alloc_pool(c, shape, nb, kvx.HND, heads=2, tp_rank=r, tp_size=4) for r in range(2)]
    Dp = synth.alloc_pool(c, shape, nb, kvx.NHD, heads=4, tp_rank=0, tp_size=2)
    back = [synth.alloc_pool(c, shape, nb, kvx.HND, heads=2, tp_rank=r, tp_size=4) for r in range(2)]
    for r in range(2):
        synth.fill_random(P[r], 700 + r)
        synth.fill_canary(back[r])
    synth.fill_canary(Dp)
    tp, td, tb = (synth.make_table(nb, n, s) for s in (1, 2, 3))
    dtp, dtd, dtb = dev_table(tp), dev_table(td), dev_table(tb)
    for r in range(2):
        c.handoff_fused(P[r].pool, dtp, Dp.pool, dtd, n, kvx.Xform(0, 2 * r, 2, kvx.NHD))      # slot r of the decode chunk
    for r in range(2):
        c.handoff_fused(Dp.pool, dtd, back[r].pool, dtb, n, kvx.Xform(2 * r, 0, 2, kvx.HND))   # head slice back out
    torch.cuda.synchronize()
    c.check()
    N, Dh, R = shape.block_tokens, shape.head_dim, shape.regions
    ip, idd, ib = (torch.from_numpy(t.astype(np.int64)).cuda() for t in (tp, td, tb))
    dview = Dp.tensor().view(torch.int16).view(R, nb, N, 4, Dh)[:, idd]
    for r in range(2):
        a = P[r].tensor().view(torch.int16).view(R, nb, 2, N, Dh)[:, ip]
        b = back[r].tensor().view(torch.int16).view(R, nb, 2, N, Dh)[:, ib]
        assert torch.equal(a, b)
        assert torch.equal(a.transpose(2, 3), dview[:, :, :, 2 * r:2 * r + 2])
    canary = np.array(synth.CANARY16, dtype=np.uint16).view(np.int16).item()
    mask = torch.ones(nb, dtype=torch.bool, device="cuda")
    mask[idd] = False
    assert (Dp.tensor().view(torch.int16).view(R, nb, -1)[:, mask] == canary).all()
    for p in P + back + [Dp]:
        p.free()
    c.close()
