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
    """Common part of both roles: a context, a KV pool, a block allocator, completion flags, and the
    peers it has shaken hands with."""

    role = "worker"

    def __init__(self, device: int, shape: KVShape, num_blocks: int, layout: int, *, tp_rank: int = 0,
                 tp_size: int = 1, arrangement: str = "split", seed: int = 0, **ctx_options):
        self.ctx = Context(device, **ctx_options)
        self.shape, self.layout, self.tp_rank, self.tp_size = shape, layout, tp_rank, tp_size
        self.kv = synth.alloc_pool(self.ctx, shape, num_blocks, layout, tp_rank=tp_rank, tp_size=tp_size,
                                   arrangement=arrangement)
        self.blocks = BlockAllocator(num_blocks, seed)
        self.flags_ptr, self.num_flags = self.ctx.flags()
        self.remotes: Dict[int, _Remote] = {}

    @property
    def pool(self) -> Pool:
        return self.kv.pool

    def metadata(self) -> WorkerMetadata:
        return WorkerMetadata(self.role, self.tp_rank, self.tp_size, self.layout, self.kv.num_blocks,
                              self.shape.block_tokens, self.kv.heads, self.pool.export(),
                              self.ctx.mem_export(self.flags_ptr, 8 * self.num_flags), self.num_flags)

    def add_remote(self, key: int, meta: WorkerMetadata) -> None:
        """Handshake (= add_remote_agent): map the peer's pool and flags into this process."""
        self.remotes[key] = _Remote(meta, self.ctx.import_pool(meta.pool_blob), self.ctx.mem_import(meta.flags_blob))

    # -- completion flags in THIS worker's memory ---------------------------------------------------
    def flag_addr(self, slot: int) -> int:
        return self.flags_ptr + 8 * (slot % self.num_flags)

    def expect(self, slot: int, stream=None) -> None:
        self.ctx.flag_reset(self.flag_addr(slot), 0, stream)

    def wait(self, slot: int, value: int, timeout_ms: int = 5000, stream=None) -> None:
        """Stream-ordered wait: work queued on `stream` after this sees the peer's writes."""
        self.ctx.flag_wait(self.flag_addr(slot), value, timeout_ms, stream)

    def poll(self, slot: int, value: int) -> bool:
        return self.ctx.flag_poll(self.flag_addr(slot), value)

    # -- layer groups (layer-wise handoff) -------------------------------------------------------------
    def layer_groups(self, groups: int):
        """Region ranges (first, count) of `groups` contiguous groups of layers; a layer is its K and V region."""
        layers = self.shape.layers
        if not 1 <= groups <= layers:
            raise ValueError(f"{groups} layer groups for {layers} layers")
        cuts = [round(g * layers / groups) for g in range(groups + 1)]
        return [(2 * cuts[g], 2 * (cuts[g + 1] - cuts[g])) for g in range(groups)]

    def _slices(self, pool: Pool, tag, groups: int):
        cache = self.__dict__.setdefault("_slice_cache", {})
        if (tag, groups) not in cache:
            cache[(tag, groups)] = [pool.slice(first, count) for first, count in self.layer_groups(groups)]
        return cache[(tag, groups)]

    @staticmethod
    def layer_ready_slot(seq: int, group: int, groups: int, num_flags: int, src: int = 0, n_src: int = 1) -> int:
        """Flag slot (upper half of the receiver's flags) that says "layer group `group` of request `seq` is ready"
        on the `src`-th of the `n_src` prefill TP ranks this decode rank reads (fan-in when prefill TP > decode TP)."""
        half = num_flags // 2
        return half + ((seq * groups + group) * n_src + src) % half

    def _move(self, src_pool, src_ids, dst_pool, dst_ids, n, xf, flag, value, host_tables, h_done, stream):
        if host_tables:
            self.ctx.handoff_host(src_pool, np.ascontiguousarray(src_ids, dtype=np.int32), dst_pool,
                                  np.ascontiguousarray(dst_ids, dtype=np.int32), n, xf, flag, value, h_done, stream)
        else:  # tables already resident on this GPU (equal length, device int32)
            self.ctx.handoff_fused(src_pool, src_ids, dst_pool, dst_ids, n, xf, flag, value, stream)

    def close(self) -> None:
        for q in self.__dict__.get("_queues", {}).values():
            q.stop()
        for views in self.__dict__.get("_slice_cache", {}).values():
            for v in views:
                v.release()
        for r in self.remotes.values():
            r.pool.release()
        self.kv.free()
        self.ctx.close()


@dataclass
class _Remote:
    meta: WorkerMetadata
    pool: Pool
    flags_ptr: int

    def flag_addr(self, slot: int) -> int:
        return self.flags_ptr + 8 * (slot % self.meta.num_flags)


class DecodeWorker(_Worker):
    """Consumer (--is-decode-worker): allocates fresh block-table slots and PULLS the request's blocks
    out of the prefill worker's pool — the READ the reference's images issue from the decode side
    (vllm v1/nixl/worker.py:2220-2233).  The kernel runs on the decode GPU: remote blocks → shared
    memory → local slots, layout transform in flight; when it is done the decode side's own stream
    order makes the KV usable, and a release-store into the PREFILL worker's flag tells it the blocks
    can be freed (H6).  Reads over NVLink measured 784 GB/s against 714 GB/s for pushed stores."""

    role = "decode"

    def pull(self, key: int, remote_ids, local_ids, *, notify_slot: Optional[int] = None, value: int = 1,
             edge: Optional[TPEdge] = None, host_tables: bool = True, h_done=None, stream=None) -> None:
        rem = self.remotes[key]
        xf = edge.xform(self.layout) if edge else None
        flag = rem.flag_addr(notify_slot) if notify_slot is not None else None
        if host_tables:
            remote_ids, local_ids = align_tables(np.asarray(remote_ids, np.int32), np.asarray(local_ids, np.int32))
        self._move(rem.pool, remote_ids, self.pool, local_ids, len(local_ids), xf, flag, value, host_tables, h_done, stream)


    def pull_layers(self, key: int, remote_ids, local_ids, group: int, groups: int, *, seq: Optional[int] = None,
                    notify_slot: Optional[int] = None, value: int = 1, edge: Optional[TPEdge] = None, h_done=None,
                    timeout_ms: int = 10000, src: int = 0, n_src: int = 1, stream=None) -> None:
        """Layer-wise handoff, receiving side: pull ONE group of layers of a request as soon as the prefill worker
        has produced it, so that the transfer overlaps the rest of prefill (the router-level behaviour SURVEY.md
        §8f.4 names; the reference's manifests leave it to the images).  With `seq` the pull first waits, ON THE
        STREAM, for PrefillWorker.publish_layers(seq, group) — no host round trip between "layers ready" and the
        pull.  The release flag (`notify_slot`) is normally given with the last group only."""
        rem = self.remotes[key]
        xf = edge.xform(self.layout) if edge else None
        src = self._slices(rem.pool, key, groups)[group]
        dst = self._slices(self.pool, "own", groups)[group]
        if seq is not None:
            self.wait(self.layer_ready_slot(seq, group, groups, self.num_flags, src, n_src), seq + 1, timeout_ms, stream)
        flag = rem.flag_addr(notify_slot) if notify_slot is not None else None
        r, l = align_tables(np.asarray(remote_ids, np.int32), np.asarray(local_ids, np.int32))
        self._move(src, r, dst, l, len(l), xf, flag, value, True, h_done, stream)

    def pull_queue(self, key: int, *, edge: Optional[TPEdge] = None, max_ctas: int = 0, idle_us: int = 0):
        """The resident receiver for prefill worker `key` (kvx_queue_start): requests submitted with pull_queued
        are served by one persistent kernel, back to back, without a launch per request."""
        cache = self.__dict__.setdefault("_queues", {})
        if key not in cache:
            xf = edge.xform(self.layout) if edge else None
            cache[key] = self.ctx.queue(self.remotes[key].pool, self.pool, xf, max_ctas=max_ctas, idle_us=idle_us)
        return cache[key]

    def pull_queued(self, key: int, remote_ids, local_ids, *, notify_slot: Optional[int] = None, value: int = 1,
                    h_done=None) -> int:
        """pull() through the resident receiver; returns the ticket (Queue.wait / Queue.poll)."""
        rem = self.remotes[key]
        r, l = align_tables(np.asarray(remote_ids, np.int32), np.asarray(local_ids, np.int32))
        flag = rem.flag_addr(notify_slot) if notify_slot is not None else None
        return self.pull_queue(key).submit(np.ascontiguousarray(r), np.ascontiguousarray(l), len(l), flag, value, h_done)

    def pull_many(self, key: int, requests, *, edge: Optional[TPEdge] = None, stream=None) -> None:
        """All the requests a scheduler step has to receive from one prefill worker, in one launch
        (= the loop over reqs_to_recv in vllm v1/nixl/worker.py:1933-1968).  `requests` is a sequence of
        (remote_ids, local_ids, notify_slot or None, value, h_done or None); every request's release flag
        and completion word are written when the whole batch has landed."""
        rem = self.remotes[key]
        xf = edge.xform(self.layout) if edge else None
        items = []
        for remote_ids, local_ids, slot, value, h_done in requests:
            r, l = align_tables(np.asarray(remote_ids, np.int32), np.asarray(local_ids, np.int32))
            items.append((r, l, rem.flag_addr(slot) if slot is not None else None, value, h_done))
        self.ctx.handoff_batch(rem.pool, self.pool, items, xf, stream)


class PrefillWorker(_Worker):
    """Producer (--is-prefill-worker): owns the P-side pool, publishes a finished request's block ids
    and frees them once every consumer has released them.  `push` is the write-direction variant of
    the same fused kernel (prefill GPU gathers and stores into the decode worker's slots)."""

    role = "prefill"

    def publish_layers(self, key: int, seq: int, group: int, groups: int, stream=None, *, src: int = 0, n_src: int = 1,
                       value: Optional[int] = None) -> None:
        """Layer-wise handoff, producing side: everything queued on `stream` so far has written layer group
        `group` of request `seq` — tell decode worker `key` (an 8-byte store into ITS flag memory, ordered on the
        stream behind the prefill kernels), which lets its pull_layers(…, seq=seq) go."""
        rem = self.remotes[key]
        slot = self.layer_ready_slot(seq, group, groups, rem.meta.num_flags, src, n_src)
        self.ctx.flag_reset(rem.flag_addr(slot), seq + 1 if value is None else value, stream)

    def push(self, key: int, src_ids, dst_ids, *, slot: int, value: int, edge: Optional[TPEdge] = None,
             host_tables: bool = True, h_done=None, stream=None) -> None:
        rem = self.remotes[key]
        xf = edge.xform(rem.meta.layout) if edge else None
        if host_tables:
            src_ids, dst_ids = align_tables(np.asarray(src_ids, np.int32), np.asarray(dst_ids, np.int32))
        self._move(self.pool, src_ids, rem.pool, dst_ids, len(dst_ids), xf, rem.flag_addr(slot), value, host_tables,
                   h_done, stream)
