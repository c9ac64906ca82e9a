"""kvx_handoff_static — the entry point for callers that must not allocate device memory (an engine
whose HBM is fully committed to its KV pool, or a request-path thread that cannot afford the
device-wide synchronisation of cudaMalloc).  Called through the C ABI and compared bit for bit with
the existing kvx_handoff_host on the same inputs, twice in a row: the second call reuses the
library's fixed scratch array, which still holds the first call's tables, and must not be affected
by them.  Collected last so that a failure here cannot hide any other test."""
import numpy as np
import pytest
import torch

import kvx
from kvx import api
from oracle import kvx_oracle as o

pytestmark = pytest.mark.gpu


def _device_copy(pool):
    return torch.from_numpy(pool.buf.copy()).cuda()


def _desc(t, pool):
    return api.pool_desc([t.data_ptr() + off for off in pool.region_off], num_blocks=pool.num_blocks, heads=pool.heads,
                         head_dim=pool.head_dim, layout=pool.layout, block_tokens=pool.block_tokens,
                         elem_bytes=pool.elem_bytes, block_stride=pool.block_stride)


def _registered(ctx, t, pool):
    return ctx.register_pool([t.data_ptr() + off for off in pool.region_off], num_blocks=pool.num_blocks,
                             heads=pool.heads, head_dim=pool.head_dim, layout=pool.layout,
                             block_tokens=pool.block_tokens, elem_bytes=pool.elem_bytes,
                             block_stride=pool.block_stride)


def test_static_entry_matches_host_entry_twice():
    L, H, D, nb = 4, 8, 128, 200
    src = o.make_pool(L, H, D, nb, o.HND, seed=42)
    dst = o.make_pool(L, H, D, nb, o.NHD, fill="canary")
    t_src = _device_copy(src)
    t_ref, t_new = _device_copy(dst), _device_copy(dst)   # existing entry point / new entry point
    flag = torch.zeros(2, dtype=torch.int64, device="cuda")

    # The existing entry runs with a deeper pipeline (128 KiB of shared memory) than the static entry
    # (64 KiB) on the same kernel: the static entry's one-time set-up must not lower the kernel's
    # shared-memory limit under this context, or its second launch below would fail.
    ctx = kvx.Context(0, stages=4, stage_bytes=32768, consumer_warps=16)
    p_src, p_ref = _registered(ctx, t_src, src), _registered(ctx, t_ref, dst)
    (d_src, keep1), (d_new, keep2) = _desc(t_src, src), _desc(t_new, dst)

    # call 1: 150 blocks; call 2: 37 other blocks with a TP-style head slice (stale tables must not leak)
    calls = [
        (o.make_table(nb, 150, 1), o.make_table(nb, 150, 2), None),
        (o.make_table(nb, 37, 3), o.make_table(nb, 37, 4), kvx.Xform(2, 4, 3, kvx.NHD)),
    ]
    # The first call on a device does the one-time set-up (it may load the kernel image); from then on a call
    # must leave free device memory exactly as it found it — that is what "allocates nothing" means to an engine
    # that has budgeted every byte.  n = 0 is the start-up call the header recommends.
    assert api.handoff_static(0, d_src, None, d_new, None, 0) == 0
    torch.cuda.synchronize()
    for i, (ts, td, xf) in enumerate(calls):
        ctx.handoff_host(p_src, ts, p_ref, td, len(ts), xf, flag_val=i + 1)
        torch.cuda.synchronize()
        # block-id lists in PINNED memory, overwritten as soon as the call returns: the call must have taken them
        pin_s, pin_d = torch.from_numpy(ts.copy()).pin_memory(), torch.from_numpy(td.copy()).pin_memory()
        free_before, _ = torch.cuda.mem_get_info()
        rc = api.handoff_static(0, d_src, pin_s, d_new, pin_d, len(ts), xf, flag=flag.data_ptr() + 8, flag_val=10 + i)
        pin_s.zero_(); pin_d.zero_()
        assert rc == 0, rc
        torch.cuda.synchronize()
        free_after, _ = torch.cuda.mem_get_info()
        assert free_after == free_before, f"call {i + 1} changed free device memory by {free_before - free_after} bytes"
        assert torch.equal(t_new, t_ref), f"call {i + 1}: static entry differs from kvx_handoff_host"
        assert int(flag[1]) == 10 + i
    # and both equal the oracle
    for ts, td, xf in calls:
        o.handoff(src, ts, dst, td, None if xf is None else o.Xform(xf.src_head_off, xf.dst_head_off, xf.heads))
    assert np.array_equal(t_new.cpu().numpy(), dst.buf)
    ctx.check()
    ctx.close()


def test_static_entry_rejects_oversize_and_bad_ids():
    src = o.make_pool(1, 8, 64, 8, o.NHD, seed=1)
    dst = o.make_pool(1, 8, 64, 8, o.NHD, fill="canary")
    t_src, t_dst = _device_copy(src), _device_copy(dst)
    (d_src, k1), (d_dst, k2) = _desc(t_src, src), _desc(t_dst, dst)
    big = np.ones(api.STATIC_MAX_BLOCKS + 1, dtype=np.int32)
    assert api.handoff_static(0, d_src, big, d_dst, big, len(big)) == api.ERANGE
    bad = np.array([1, 8], dtype=np.int32)  # id 8 is out of range
    assert api.handoff_static(0, d_src, bad, d_dst, bad, 2) == api.ERANGE
    torch.cuda.synchronize()
    assert np.array_equal(t_dst.cpu().numpy(), dst.buf)  # nothing was launched
    one = np.array([3], dtype=np.int32)
    assert api.handoff_static(0, d_src, one, d_dst, np.array([5], dtype=np.int32), 1) == 0
    torch.cuda.synchronize()
    o.handoff(src, [3], dst, [5])
    assert np.array_equal(t_dst.cpu().numpy(), dst.buf)
