"""Prefill / decode worker split of the reference's disaggregated manifests.

examples/deploy/vllm/disagg.yaml:38-57 (VllmPrefillWorker, --is-prefill-worker) owns the P-side pool
and pushes a finished request's blocks; :18-37 (VllmDecodeWorker, --is-decode-worker) owns the D-side
pool, allocates fresh blocks and waits for completion.  One worker = one OS process = one GPU
(resources.limits.gpu: "1").  The control records mirror what the images exchange:
`WorkerMetadata` ~ NixlAgentMetadata (vllm v1/nixl/metadata.py:41-53), `TransferParams` ~
kv_transfer_params (v1/nixl/scheduler.py:664-674).
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, List, Optional

import numpy as np

from . import synth
from .api import Context, Pool, Xform
from .shapes import KVShape
from .tp import TPEdge


class BlockAllocator:
    """Free-list block allocator; id 0 is the null block and is never handed out."""

    def __init__(self, num_blocks: int, seed: int = 0):
        rng = np.random.Generator(np.random.PCG64(seed))
        self._free: List[int] = (1 + rng.permutation(num_blocks - 1)).astype(np.int32).tolist()
        self._held: Dict[str, np.ndarray] = {}

    @property
    def free_blocks(self) -> int:
        return len(self._free)

    def allocate(self, request_id: str, n: int) -> np.ndarray:
        if request_id in self._held:
            raise KeyError(f"request {request_id} already holds blocks")
        if n > len(self._free):
            raise MemoryError(f"need {n} blocks, {len(self._free)} free")
        ids = np.array([self._free.pop() for _ in range(n)], dtype=np.int32)
        self._held[request_id] = ids
        return ids

    def free(self, request_id: str) -> None:
        ids = self._held.pop(request_id)
        self._free.extend(int(i) for i in ids)

    def table(self, request_id: str) -> np.ndarray:
        return self._held[request_id]


@dataclass
class WorkerMetadata:
    """Handshake record a decode worker publishes once per engine pair."""

    role: str
    tp_rank: int
    tp_size: int
    layout: int
    num_blocks: int
    block_tokens: int
    heads: int
    pool_blob: bytes
    flags_blob: bytes
    num_flags: int


@dataclass
class TransferParams:
    """Per-request control record: who holds the blocks and which ones."""

    request_id: str
    remote_block_ids: np.ndarray
    remote_rank: int
    tp_size: int


def align_tables(src_ids: np.ndarray, dst_ids: np.ndarray):
    """Prefix-cache hits on the decode side (vllm v1/nixl/worker.py:2166-2185, 2307-2316): with fewer
    destination than source blocks take the LAST len(dst) source blocks; an empty destination list
    means notify only."""
    if len(dst_ids) > len(src_ids):
        raise ValueError("decode side holds more blocks than prefill produced")
    return src_ids[len(src_ids) - len(dst_ids):], dst_ids


class _Worker:
    role = "worker"

    def __init__(self, device: int, shape: KVShape, num_blocks: int, layout: int, *, tp_rank: int = 0,
                 tp_size: int = 1, arrangement: str = "split", seed: int = 0, **ctx_options):
        self.ctx = Context(device, **ctx_options)
        self.shape, self.layout, self.tp_rank, self.tp_size = shape, layout, tp_rank, tp_size
        self.kv = synth.alloc_pool(self.ctx, shape, num_blocks, layout, tp_rank=tp_rank, tp_size=tp_size,
                                   arrangement=arrangement)
        self.blocks = BlockAllocator(num_blocks, seed)
        self.flags_ptr, self.num_flags = self.ctx.flags()

    @property
    def pool(self) -> Pool:
        return self.kv.pool

    def metadata(self) -> WorkerMetadata:
        return WorkerMetadata(self.role, self.tp_rank, self.tp_size, self.layout, self.kv.num_blocks,
                              self.shape.block_tokens, self.kv.heads, self.pool.export(),
                              self.ctx.mem_export(self.flags_ptr, 8 * self.num_flags), self.num_flags)

    def close(self) -> None:
        self.kv.free()
        self.ctx.close()


class DecodeWorker(_Worker):
    """Consumer: allocates fresh block-table slots and waits for the completion flag (H5, H6)."""

    role = "decode"

    def flag_addr(self, slot: int) -> int:
        return self.flags_ptr + 8 * (slot % self.num_flags)

    def expect(self, slot: int, stream=None) -> None:
        self.ctx.flag_reset(self.flag_addr(slot), 0, stream)

    def wait(self, slot: int, value: int, timeout_ms: int = 5000, stream=None) -> None:
        """Stream-ordered wait: work queued on `stream` after this sees the request's KV."""
        self.ctx.flag_wait(self.flag_addr(slot), value, timeout_ms, stream)

    def poll(self, slot: int, value: int) -> bool:
        return self.ctx.flag_poll(self.flag_addr(slot), value)


@dataclass
class _Remote:
    meta: WorkerMetadata
    pool: Pool
    flags_ptr: int


class PrefillWorker(_Worker):
    """Producer: gathers a request's scattered blocks and pushes them into a decode worker's slots."""

    role = "prefill"

    def __init__(self, *a, **kw):
        super().__init__(*a, **kw)
        self.remotes: Dict[int, _Remote] = {}

    def add_remote(self, key: int, meta: WorkerMetadata) -> None:
        """Handshake (= add_remote_agent): map the decode worker's pool and flags into this process."""
        self.remotes[key] = _Remote(meta, self.ctx.import_pool(meta.pool_blob), self.ctx.mem_import(meta.flags_blob))

    def push(self, key: int, src_ids, dst_ids, *, slot: int, value: int, edge: Optional[TPEdge] = None,
             host_tables: bool = True, h_done=None, stream=None) -> None:
        """One P rank → D rank edge of one request, fused: gather → (transform) → peer store → flag."""
        rem = self.remotes[key]
        xf = edge.xform(rem.meta.layout) if edge else None
        flag = rem.flags_ptr + 8 * (slot % rem.meta.num_flags)
        if host_tables:
            src_ids, dst_ids = align_tables(np.asarray(src_ids, np.int32), np.asarray(dst_ids, np.int32))
            self.ctx.handoff_host(self.pool, np.ascontiguousarray(src_ids), rem.pool, np.ascontiguousarray(dst_ids),
                                  len(dst_ids), xf, flag, value, h_done, stream)
        else:  # tables already resident on this GPU (equal length, device int32)
            self.ctx.handoff_fused(self.pool, src_ids, rem.pool, dst_ids, len(dst_ids), xf, flag, value, stream)
