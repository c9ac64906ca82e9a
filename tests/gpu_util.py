"""Helpers for the GPU parity tests: move oracle pools to the device and back."""
import numpy as np
import torch

import kvx
from oracle import kvx_oracle as o


class DevPool:
    def __init__(self, ctx: kvx.Context, pool: o.Pool):
        self.host = pool
        self.t = torch.from_numpy(pool.buf.copy()).to(f"cuda:{ctx.device}")
        base = self.t.data_ptr()
        self.kp = ctx.register_pool([base + off for off in pool.region_off], num_blocks=pool.num_blocks,
                                    heads=pool.heads, head_dim=pool.head_dim, layout=pool.layout,
                                    block_tokens=pool.block_tokens, elem_bytes=pool.elem_bytes,
                                    block_stride=pool.block_stride, tp_rank=pool.tp_rank, tp_size=pool.tp_size)

    def bytes(self) -> np.ndarray:
        return self.t.cpu().numpy()


def dev_table(table, device=0):
    return torch.from_numpy(np.ascontiguousarray(table, dtype=np.int32)).to(f"cuda:{device}")


def xf_c(x: o.Xform, dflt_layout: int) -> kvx.Xform:
    return kvx.Xform(x.src_head_off, x.dst_head_off, x.heads, dflt_layout if x.stage_layout is None else x.stage_layout)


ENGINE_CONFIGS = [
    dict(engine="ldg", vec_bytes=16),
    dict(engine="ldg", vec_bytes=32),
    dict(engine="tma", store_mode=0, stages=4, stage_bytes=32768, consumer_warps=8),
    dict(engine="tma", store_mode=1, vec_bytes=16, stages=2, stage_bytes=16384, consumer_warps=4),
    dict(engine="tma", store_mode=1, vec_bytes=32, stages=3, stage_bytes=32768, consumer_warps=8),
    dict(engine="tma", store_mode=2, stages=4, stage_bytes=8192, consumer_warps=2),
    dict(engine="tma", store_mode=2, stages=6, stage_bytes=32768, consumer_warps=8, ctas_per_sm=2, l2_hint=1),
    dict(engine="tma", max_ctas=5, pdl=0),          # bounded SM footprint, plain launches
    dict(engine="tma", stages=6, stage_bytes=32768, ctas_per_sm=1, max_ctas=43),   # the geometry NVLink transfers default to
    dict(engine="tma", store_mode=3),                                     # tensor-map stores where the case allows, else thread stores
    dict(engine="tma", store_mode=3, stages=6, stage_bytes=32768, ctas_per_sm=3),
]


def cfg_id(c):
    return "-".join(f"{k}{v}" for k, v in c.items())
