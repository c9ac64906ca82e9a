"""Synthetic workload of SURVEY.md §8d: random-bit KV pools on the device, scattered block tables.

Device memory comes from kvx_mem_alloc (plain cudaMalloc, exportable to a peer process); torch is
used only to move bytes between numpy and that memory.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Optional

import numpy as np

from .api import Context, Pool
from .shapes import KVShape

CANARY16 = 0xC0DE
SEED_BASE = 0xD15A66


class _Raw:
    """A raw device range exposed to torch through __cuda_array_interface__."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 3}


def as_torch(ptr: int, nbytes: int, device: int):
    import torch

    return torch.as_tensor(_Raw(ptr, nbytes), device=f"cuda:{device}")


@dataclass
class DevicePool:
    """One rank's KV pool in device memory: a single allocation cut into regions."""

    ctx: Context
    pool: Pool
    ptr: int
    nbytes: int
    region_off: List[int]
    block_stride: int
    shape: KVShape
    heads: int
    layout: int
    num_blocks: int

    def tensor(self):
        return as_torch(self.ptr, self.nbytes, self.ctx.device)

    def to_numpy(self) -> np.ndarray:
        return self.tensor().cpu().numpy()

    def free(self) -> None:
        self.pool.release()
        self.ctx.mem_free(self.ptr)


def alloc_pool(ctx: Context, shape: KVShape, num_blocks: int, layout: int, *, heads: Optional[int] = None,
               arrangement: str = "split", tp_rank: int = 0, tp_size: int = 1) -> DevicePool:
    heads = heads if heads is not None else max(1, shape.kv_heads // tp_size)
    chunk = shape.block_tokens * heads * shape.head_dim * shape.elem_bytes
    regions = shape.regions
    nbytes = regions * num_blocks * chunk
    ptr = ctx.mem_alloc(nbytes)
    if arrangement == "split":
        region_off, stride = [r * num_blocks * chunk for r in range(regions)], chunk
    elif arrangement == "blocks_first":
        region_off = []
        for l in range(shape.layers):
            region_off += [l * num_blocks * 2 * chunk, l * num_blocks * 2 * chunk + chunk]
        stride = 2 * chunk
    else:
        raise ValueError(arrangement)
    pool = ctx.register_pool([ptr + o for o in region_off], num_blocks=num_blocks, heads=heads,
                             head_dim=shape.head_dim, layout=layout, block_tokens=shape.block_tokens,
                             elem_bytes=shape.elem_bytes, block_stride=stride, tp_rank=tp_rank, tp_size=tp_size)
    return DevicePool(ctx, pool, ptr, nbytes, region_off, stride, shape, heads, layout, num_blocks)


def fill_random(dp: DevicePool, seed: int) -> None:
    """Uniform random 16-bit patterns generated on the device (NaN/Inf/denormal bf16 included)."""
    import torch

    g = torch.Generator(device=f"cuda:{dp.ctx.device}")
    g.manual_seed(seed)
    t = dp.tensor()
    step = 1 << 28
    for off in range(0, dp.nbytes, step):
        part = t[off : off + step]
        part.view(torch.int16).random_(-32768, 32768, generator=g)


def fill_canary(dp: DevicePool) -> None:
    import torch

    dp.tensor().view(torch.int16).fill_(np.array(CANARY16, dtype=np.uint16).view(np.int16).item())


def upload(dp: DevicePool, host_bytes: np.ndarray) -> None:
    import torch

    dp.tensor().copy_(torch.from_numpy(np.ascontiguousarray(host_bytes).view(np.uint8).reshape(-1)))


def make_table(num_blocks: int, n: int, seed: int, contiguous: bool = False) -> np.ndarray:
    """Unique, scattered, non-monotone block ids; never block 0 (vLLM's null block)."""
    if n > num_blocks - 1:
        raise ValueError("table longer than the pool")
    if contiguous:
        return np.arange(1, n + 1, dtype=np.int32)
    rng = np.random.Generator(np.random.PCG64(seed))
    return (1 + rng.permutation(num_blocks - 1)[:n]).astype(np.int32)
