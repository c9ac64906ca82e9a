"""Randomised geometry: any (shape, layouts, head slice, block-size ratio, arrangement, table) the plan
accepts must move exactly the oracle's bytes, on both engines."""
import numpy as np
import pytest
import torch
from hypothesis import HealthCheck, given, settings, strategies as st

import kvx
from gpu_util import DevPool, dev_table
from oracle import kvx_oracle as o

pytestmark = pytest.mark.gpu

_ctxs = {}


def _ctx(engine):
    if engine not in _ctxs:
        opts = {"ldg": dict(engine="ldg"), "tma": dict(engine="tma", store_mode=0), "tma_bulk": dict(engine="tma", store_mode=2),
                # host-table entry: ids in the kernel parameter block, a handful of CTAs, no dependent launch
                "tma_host": dict(engine="tma", max_ctas=3, pdl=0)}[engine]
        _ctxs[engine] = kvx.Context(0, **opts)
    return _ctxs[engine]


geometry = st.fixed_dictionaries(dict(
    L=st.integers(1, 2), D=st.sampled_from([16, 32, 64, 80, 128]), E=st.sampled_from([1, 2]),
    N=st.sampled_from([4, 16]), ks=st.sampled_from([1, 1, 2, 4]), kd=st.sampled_from([1, 1, 2, 4]),
    Hc=st.integers(1, 4), so=st.integers(0, 3), do=st.integers(0, 3), pad_s=st.integers(0, 2), pad_d=st.integers(0, 2),
    ls=st.sampled_from([0, 1]), ld=st.sampled_from([0, 1]), arr_s=st.sampled_from(["split", "blocks_first"]),
    arr_d=st.sampled_from(["split", "blocks_first"]), n=st.integers(1, 40), seed=st.integers(0, 2**20)))


@pytest.mark.parametrize("engine", ["tma", "tma_bulk", "ldg", "tma_host"])
@settings(max_examples=60, deadline=None, suppress_health_check=[HealthCheck.function_scoped_fixture, HealthCheck.too_slow])
@given(g=geometry)
def test_random_geometry_matches_oracle(engine, g):
    if (g["D"] * g["E"]) % 16:
        return  # rows must be a multiple of 16 bytes (kvx_pool_register refuses others)
    if g["ks"] > 1 and g["kd"] > 1:
        g["kd"] = 1  # one side holds the granule
    ctx = _ctx(engine)
    Ns, Nd = g["N"] * g["ks"], g["N"] * g["kd"]
    Hs, Hd = g["so"] + g["Hc"] + g["pad_s"], g["do"] + g["Hc"] + g["pad_d"]
    n = g["n"]
    ns, nd = -(-n // g["ks"]), -(-n // g["kd"])
    src = o.make_pool(g["L"], Hs, g["D"], ns + 3, g["ls"], block_tokens=Ns, elem_bytes=g["E"], arrangement=g["arr_s"], seed=g["seed"])
    dst = o.make_pool(g["L"], Hd, g["D"], nd + 4, g["ld"], block_tokens=Nd, elem_bytes=g["E"], arrangement=g["arr_d"], fill="canary")
    ts, td = o.make_table(ns + 3, ns, g["seed"] + 1), o.make_table(nd + 4, nd, g["seed"] + 2)
    dsrc, ddst = DevPool(ctx, src), DevPool(ctx, dst)
    if engine == "tma_host":
        ctx.handoff_host(dsrc.kp, ts, ddst.kp, td, n, kvx.Xform(g["so"], g["do"], g["Hc"], g["ld"]))
    else:
        ctx.handoff_fused(dsrc.kp, dev_table(ts), ddst.kp, dev_table(td), n, kvx.Xform(g["so"], g["do"], g["Hc"], g["ld"]))
    torch.cuda.synchronize()
    ctx.check()
    o.handoff_regroup(src, ts, dst, td, o.Xform(g["so"], g["do"], g["Hc"]), n_small=n)
    assert np.array_equal(ddst.bytes(), dst.buf)
    dsrc.kp.release(); ddst.kp.release()
