"""Thin object layer over the C ABI: Context, Pool, and the transfer verbs.

Mirrors the verbs the reference's images use on this path (vLLM NIXL connector, cited in
include/kvx.h): register_kv_caches → Context.register_pool, get_agent_metadata/add_remote_agent →
Pool.export / Context.import_pool, make_prepped_xfer+transfer → gather/copy/scatter/handoff,
check_xfer_state → flag_poll / flag_wait.
"""
from __future__ import annotations

import ctypes
from dataclasses import dataclass
from typing import Optional, Sequence

from . import _lib
from ._lib import BatchItem, PoolDesc, XformC, c_i32, c_i64, c_sz, c_u32, c_u64, c_vp

NHD, HND = 0, 1
OPT = {"engine": 0, "stages": 1, "stage_bytes": 2, "ctas_per_sm": 3, "store_mode": 4, "vec_bytes": 5,
       "consumer_warps": 6, "l2_hint": 7, "pdl": 8, "param_tables": 9, "max_ctas": 10}
ENGINE = {"auto": 0, "ldg": 1, "tma": 2}
OK, EINVAL, ECUDA, ENOMEM, EUNSUPPORTED, ETIMEOUT, ERANGE, EIPC, ENOTREADY = 0, -1, -2, -3, -4, -5, -6, -7, -8


class KvxError(RuntimeError):
    def __init__(self, code: int, where: str, detail: str = ""):
        self.code = code
        msg = _lib.load().kvx_strerror(code).decode()
        super().__init__(f"{where}: {msg} ({code})" + (f": {detail}" if detail else ""))


@dataclass
class Xform:
    """kvx_xform: head slice moved by one edge and the staging row order."""

    src_head_off: int = 0
    dst_head_off: int = 0
    heads: int = 0
    stage_layout: int = NHD

    def c(self) -> XformC:
        return XformC(self.src_head_off, self.dst_head_off, self.heads, self.stage_layout)


def _ptr(x) -> Optional[int]:
    """Device/host address from an int, a torch tensor, a numpy array or None."""
    if x is None:
        return None
    if isinstance(x, int):
        return x
    if hasattr(x, "data_ptr"):
        return x.data_ptr()
    if hasattr(x, "ctypes"):
        return x.ctypes.data
    raise TypeError(type(x))


def _stream(stream) -> Optional[int]:
    if stream is None:
        return None
    if isinstance(stream, int):
        return stream or None
    return stream.cuda_stream or None  # torch.cuda.Stream


class Pool:
    def __init__(self, ctx: "Context", handle: int, *, regions, block_tokens, heads, head_dim, elem_bytes, layout,
                 num_blocks, imported=False):
        self.ctx, self.handle = ctx, handle
        self.regions, self.block_tokens, self.heads = regions, block_tokens, heads
        self.head_dim, self.elem_bytes, self.layout, self.num_blocks = head_dim, elem_bytes, layout, num_blocks
        self.imported = imported

    @property
    def chunk_bytes(self) -> int:
        out = c_u64()
        self.ctx._check(self.ctx.lib.kvx_pool_chunk_bytes(self.handle, ctypes.byref(out)), "kvx_pool_chunk_bytes")
        return out.value

    def export(self) -> bytes:
        need = c_sz(0)
        self.ctx.lib.kvx_pool_export(self.handle, None, ctypes.byref(need))
        buf = ctypes.create_string_buffer(need.value)
        self.ctx._check(self.ctx.lib.kvx_pool_export(self.handle, buf, ctypes.byref(need)), "kvx_pool_export")
        return buf.raw[: need.value]

    def region_base(self, region: int) -> int:
        out = c_vp()
        self.ctx._check(self.ctx.lib.kvx_pool_region_base(self.handle, region, ctypes.byref(out)), "kvx_pool_region_base")
        return out.value

    def slice(self, first_region: int, count: int) -> "Pool":
        """The K/V regions of a group of layers as a pool of their own (kvx_pool_slice); release before `self`."""
        h = c_vp()
        self.ctx._check(self.ctx.lib.kvx_pool_slice(self.handle, first_region, count, ctypes.byref(h)), "kvx_pool_slice")
        return Pool(self.ctx, h.value, regions=count, block_tokens=self.block_tokens, heads=self.heads,
                    head_dim=self.head_dim, elem_bytes=self.elem_bytes, layout=self.layout, num_blocks=self.num_blocks,
                    imported=self.imported)

    @property
    def exportable_across_processes(self) -> bool:
        out = c_i32()
        self.ctx._check(self.ctx.lib.kvx_pool_exportable(self.handle, ctypes.byref(out)), "kvx_pool_exportable")
        return bool(out.value)

    def export_fds(self):
        """(record, fds) of a pool in VMM memory (kvx_pool_export_fds): send the record and pass the fds with SCM_RIGHTS
        (socket.send_fds), then close them."""
        need, cnt = c_sz(0), c_u32(0)
        self.ctx.lib.kvx_pool_export_fds(self.handle, None, ctypes.byref(need), None, ctypes.byref(cnt))
        buf = ctypes.create_string_buffer(need.value)
        fds = (c_i32 * max(1, cnt.value))()
        self.ctx._check(self.ctx.lib.kvx_pool_export_fds(self.handle, buf, ctypes.byref(need), fds, ctypes.byref(cnt)), "kvx_pool_export_fds")
        return buf.raw[: need.value], list(fds[: cnt.value])

    def release(self) -> None:
        if self.handle:
            self.ctx.lib.kvx_pool_release(self.handle)
            self.handle = None


class Context:
    """One worker's transfer context on one GPU (kvx_ctx)."""

    def __init__(self, device: int = 0, **options):
        self.lib = _lib.load()
        h = c_vp()
        rc = self.lib.kvx_ctx_create(device, ctypes.byref(h))
        if rc != OK:
            raise KvxError(rc, "kvx_ctx_create", "a CUDA device is required; kvx has no CPU fallback")
        self.handle, self.device = h.value, device
        for k, v in options.items():
            self.set_option(k, v)

    # -- plumbing -------------------------------------------------------------------------------
    def _check(self, rc: int, where: str) -> None:
        if rc != OK:
            raise KvxError(rc, where, self.lib.kvx_last_error(self.handle).decode() if rc == ECUDA else "")

    def close(self) -> None:
        if self.handle:
            self.lib.kvx_ctx_destroy(self.handle)
            self.handle = None

    def __enter__(self) -> "Context":
        return self

    def __exit__(self, *exc) -> None:
        self.close()

    def set_option(self, name: str, value) -> None:
        if name == "engine" and isinstance(value, str):
            value = ENGINE[value]
        self._check(self.lib.kvx_ctx_set_option(self.handle, OPT[name], int(value)), f"kvx_ctx_set_option({name})")

    def get_option(self, name: str) -> int:
        out = c_i64()
        self._check(self.lib.kvx_ctx_get_option(self.handle, OPT[name], ctypes.byref(out)), "kvx_ctx_get_option")
        return out.value

    @property
    def launch_count(self) -> int:
        out = c_u64()
        self._check(self.lib.kvx_ctx_launch_count(self.handle, ctypes.byref(out)), "kvx_ctx_launch_count")
        return out.value

    @property
    def sm_count(self) -> int:
        out = c_i32()
        self._check(self.lib.kvx_ctx_sm_count(self.handle, ctypes.byref(out)), "kvx_ctx_sm_count")
        return out.value

    def check(self) -> None:
        self._check(self.lib.kvx_ctx_check(self.handle), "kvx_ctx_check")

    # -- memory ---------------------------------------------------------------------------------
    def mem_alloc(self, nbytes: int) -> int:
        out = c_vp()
        self._check(self.lib.kvx_mem_alloc(self.handle, nbytes, ctypes.byref(out)), "kvx_mem_alloc")
        return out.value

    def mem_free(self, ptr: int) -> None:
        self._check(self.lib.kvx_mem_free(self.handle, ptr), "kvx_mem_free")

    def mem_alloc_vmm(self, nbytes: int, piece_bytes: int = 0) -> int:
        """Device memory from CUDA virtual memory management, exportable as file descriptors (kvx_mem_alloc_vmm)."""
        out = c_vp()
        self._check(self.lib.kvx_mem_alloc_vmm(self.handle, nbytes, piece_bytes, ctypes.byref(out)), "kvx_mem_alloc_vmm")
        return out.value

    def mem_free_vmm(self, ptr: int) -> None:
        self._check(self.lib.kvx_mem_free_vmm(self.handle, ptr), "kvx_mem_free_vmm")

    def import_pool_fds(self, blob: bytes, fds) -> Pool:
        """Map a peer's VMM pool from its record and the file descriptors received with it (kvx_pool_import_fds)."""
        import struct

        arr = (c_i32 * len(fds))(*fds)
        h = c_vp()
        self._check(self.lib.kvx_pool_import_fds(self.handle, blob, len(blob), arr, len(fds), ctypes.byref(h)), "kvx_pool_import_fds")
        (_, regions, N, H, D, E, layout, _, _, _, nb, _) = struct.unpack_from("<10I2Q", blob, 0)
        return Pool(self, h.value, regions=regions, block_tokens=N, heads=H, head_dim=D, elem_bytes=E, layout=layout,
                    num_blocks=nb, imported=True)

    def mem_export(self, ptr, nbytes: int) -> bytes:
        need = c_sz(0)
        self.lib.kvx_mem_export(self.handle, _ptr(ptr), nbytes, None, ctypes.byref(need))
        buf = ctypes.create_string_buffer(need.value)
        self._check(self.lib.kvx_mem_export(self.handle, _ptr(ptr), nbytes, buf, ctypes.byref(need)), "kvx_mem_export")
        return buf.raw[: need.value]

    def mem_import(self, blob: bytes) -> int:
        out = c_vp()
        self._check(self.lib.kvx_mem_import(self.handle, blob, len(blob), ctypes.byref(out)), "kvx_mem_import")
        return out.value

    def mem_unmap(self, ptr: int) -> None:
        self._check(self.lib.kvx_mem_unmap(self.handle, ptr), "kvx_mem_unmap")

    # -- pools ----------------------------------------------------------------------------------
    def register_pool(self, bases: Sequence[int], *, num_blocks: int, heads: int, head_dim: int, layout: int,
                      block_tokens: int = 16, elem_bytes: int = 2, block_stride: int = 0, tp_rank: int = 0,
                      tp_size: int = 1, head_stride: int = 0) -> Pool:
        arr = (c_vp * len(bases))(*[int(b) for b in bases])
        d = PoolDesc(ctypes.sizeof(PoolDesc), len(bases), block_tokens, heads, head_dim, elem_bytes, layout, tp_rank,
                     tp_size, head_stride, num_blocks, block_stride, ctypes.cast(arr, ctypes.POINTER(c_vp)))
        h = c_vp()
        self._check(self.lib.kvx_pool_register(self.handle, ctypes.byref(d), ctypes.byref(h)), "kvx_pool_register")
        return Pool(self, h.value, regions=len(bases), block_tokens=block_tokens, heads=heads, head_dim=head_dim,
                    elem_bytes=elem_bytes, layout=layout, num_blocks=num_blocks)

    def import_pool(self, blob: bytes) -> Pool:
        h = c_vp()
        self._check(self.lib.kvx_pool_import(self.handle, blob, len(blob), ctypes.byref(h)), "kvx_pool_import")
        import struct

        (_, regions, N, H, D, E, layout, _, _, _, nb, _) = struct.unpack_from("<10I2Q", blob, 0)
        return Pool(self, h.value, regions=regions, block_tokens=N, heads=H, head_dim=D, elem_bytes=E, layout=layout,
                    num_blocks=nb, imported=True)

    # -- flags ----------------------------------------------------------------------------------
    def flags(self):
        ptr, cnt = c_vp(), c_u32()
        self._check(self.lib.kvx_ctx_flags(self.handle, ctypes.byref(ptr), ctypes.byref(cnt)), "kvx_ctx_flags")
        return ptr.value, cnt.value

    def flag_reset(self, flag: int, value: int = 0, stream=None) -> None:
        self._check(self.lib.kvx_flag_reset(self.handle, flag, value, _stream(stream)), "kvx_flag_reset")

    def flag_wait(self, flag: int, value: int, timeout_ms: int = 2000, stream=None) -> None:
        self._check(self.lib.kvx_flag_wait(self.handle, flag, value, timeout_ms, _stream(stream)), "kvx_flag_wait")

    def flag_poll(self, flag: int, value: int) -> bool:
        rc = self.lib.kvx_flag_poll(self.handle, flag, value)
        if rc == ENOTREADY:
            return False
        self._check(rc, "kvx_flag_poll")
        return True

    # -- transfers ------------------------------------------------------------------------------
    def gather(self, src: Pool, d_table, n: int, staging, xf: Optional[Xform] = None, stream=None) -> None:
        x = xf.c() if xf else None
        self._check(self.lib.kvx_gather(self.handle, src.handle, _ptr(d_table), n, _ptr(staging),
                                        ctypes.byref(x) if x else None, _stream(stream)), "kvx_gather")

    def scatter(self, staging, dst: Pool, d_table, n: int, xf: Optional[Xform] = None, stream=None) -> None:
        x = xf.c() if xf else None
        self._check(self.lib.kvx_scatter(self.handle, _ptr(staging), dst.handle, _ptr(d_table), n,
                                         ctypes.byref(x) if x else None, _stream(stream)), "kvx_scatter")

    def copy(self, src, dst, nbytes: int, flag: Optional[int] = None, flag_val: int = 0, stream=None) -> None:
        self._check(self.lib.kvx_copy(self.handle, _ptr(src), _ptr(dst), nbytes, flag, flag_val, _stream(stream)),
                    "kvx_copy")

    def handoff_fused(self, src: Pool, d_src_table, dst: Pool, d_dst_table, n: int, xf: Optional[Xform] = None,
                      flag: Optional[int] = None, flag_val: int = 0, stream=None) -> None:
        x = xf.c() if xf else None
        self._check(self.lib.kvx_handoff_fused(self.handle, src.handle, _ptr(d_src_table), dst.handle,
                                               _ptr(d_dst_table), n, ctypes.byref(x) if x else None, flag, flag_val,
                                               _stream(stream)), "kvx_handoff_fused")

    def handoff_host(self, src: Pool, h_src_table, dst: Pool, h_dst_table, n: int, xf: Optional[Xform] = None,
                     flag: Optional[int] = None, flag_val: int = 1, h_done=None, stream=None) -> None:
        x = xf.c() if xf else None
        self._check(self.lib.kvx_handoff_host(self.handle, src.handle, _ptr(h_src_table), dst.handle,
                                              _ptr(h_dst_table), n, ctypes.byref(x) if x else None, flag, flag_val,
                                              _ptr(h_done), _stream(stream)), "kvx_handoff_host")


    def queue(self, src: Pool, dst: Pool, xf: Optional[Xform] = None, *, max_ctas: int = 0, idle_us: int = 0) -> "Queue":
        """Start a resident receiver for the edge src → dst (kvx_queue_start)."""
        return Queue(self, src, dst, xf, max_ctas=max_ctas, idle_us=idle_us)

    def handoff_batch(self, src: Pool, dst: Pool, requests, xf: Optional[Xform] = None, stream=None) -> None:
        """Several requests between the same two pools in one launch (kvx_handoff_batch).  `requests` is a
        sequence of (h_src_ids, h_dst_ids, flag, flag_val, h_done): int32 numpy arrays (or None for an empty
        request), a device/peer flag address or None, the value, a pinned host word (tensor / address) or None."""
        import numpy as np

        keep, arr = [], (BatchItem * len(requests))()
        for k, (s_ids, d_ids, flag, val, h_done) in enumerate(requests):
            n = 0 if d_ids is None else len(d_ids)
            if n:
                s_ids = np.ascontiguousarray(s_ids, dtype=np.int32)
                d_ids = np.ascontiguousarray(d_ids, dtype=np.int32)
                if len(s_ids) != n:
                    raise ValueError("a batched request needs equally long block-id lists")
                keep += [s_ids, d_ids]
            arr[k] = BatchItem(_ptr(s_ids) if n else None, _ptr(d_ids) if n else None, n, 0, flag, int(val), _ptr(h_done))
        x = xf.c() if xf else None
        self._check(self.lib.kvx_handoff_batch(self.handle, src.handle, dst.handle, arr, len(requests),
                                               ctypes.byref(x) if x else None, _stream(stream)), "kvx_handoff_batch")


class Queue:
    """kvx_queue: the resident receiver of one P→D edge.  A persistent kernel serves requests from a queue in
    pinned host memory: submit() costs a few hundred nanoseconds and never launches while the kernel is resident;
    back-to-back requests run into each other without draining the pipeline.  The kernel parks (exits) after
    `idle_us` without work and comes back with the next submit; park() makes it leave now — do that before
    anything that waits for the whole device (torch.cuda.synchronize, cudaFree) if you cannot afford to wait out
    the idle time."""

    def __init__(self, ctx: "Context", src: Pool, dst: Pool, xf: Optional[Xform] = None, *, max_ctas: int = 0,
                 idle_us: int = 0):
        self.ctx = ctx
        x = xf.c() if xf else None
        h = c_vp()
        ctx._check(ctx.lib.kvx_queue_start(ctx.handle, src.handle, dst.handle, ctypes.byref(x) if x else None, max_ctas,
                                           idle_us, ctypes.byref(h)), "kvx_queue_start")
        self.handle = h.value

    def submit(self, h_src_table, h_dst_table, n: Optional[int] = None, flag: Optional[int] = None, flag_val: int = 1,
               h_done=None) -> int:
        """Hand over one request (int32 numpy arrays / pinned tensors of block ids); returns its ticket."""
        if n is None:
            n = 0 if h_dst_table is None else len(h_dst_table)
        t = c_u64()
        self.ctx._check(self.ctx.lib.kvx_queue_submit(self.handle, _ptr(h_src_table), _ptr(h_dst_table), n, flag, flag_val,
                                                      _ptr(h_done), ctypes.byref(t)), "kvx_queue_submit")
        return t.value

    @staticmethod
    def pack(requests):
        """A scheduler step's requests as a kvx_batch_item array for submit_many: a sequence of (h_src_ids, h_dst_ids,
        flag or None, flag_val, pinned h_done or None).  Returns (array, keepalive)."""
        import numpy as np

        keep, arr = [], (BatchItem * len(requests))()
        for k, (s_ids, d_ids, flag, val, h_done) in enumerate(requests):
            n = 0 if d_ids is None else len(d_ids)
            if n:
                s_ids = np.ascontiguousarray(s_ids, dtype=np.int32)
                d_ids = np.ascontiguousarray(d_ids, dtype=np.int32)
                keep += [s_ids, d_ids]
            arr[k] = BatchItem(_ptr(s_ids) if n else None, _ptr(d_ids) if n else None, n, 0, flag, int(val), _ptr(h_done))
        return arr, keep

    def submit_many(self, packed, count: Optional[int] = None) -> int:
        """kvx_queue_submit_batch: all requests of a step with one call; returns the LAST ticket."""
        arr = packed[0] if isinstance(packed, tuple) else packed
        t = c_u64()
        self.ctx._check(self.ctx.lib.kvx_queue_submit_batch(self.handle, arr, len(arr) if count is None else count, ctypes.byref(t)),
                        "kvx_queue_submit_batch")
        return t.value

    def wait(self, ticket: int, timeout_ms: int = 5000) -> None:
        self.ctx._check(self.ctx.lib.kvx_queue_wait(self.handle, ticket, timeout_ms), "kvx_queue_wait")

    def poll(self, ticket: int) -> bool:
        rc = self.ctx.lib.kvx_queue_poll(self.handle, ticket)
        if rc == ENOTREADY:
            return False
        self.ctx._check(rc, "kvx_queue_poll")
        return True

    def park(self) -> None:
        self.ctx._check(self.ctx.lib.kvx_queue_park(self.handle), "kvx_queue_park")

    @property
    def launches(self) -> int:
        n, c = c_u64(), c_u32()
        self.ctx._check(self.ctx.lib.kvx_queue_launch_count(self.handle, ctypes.byref(n), ctypes.byref(c)), "kvx_queue_launch_count")
        return n.value

    @property
    def ctas(self) -> int:
        n, c = c_u64(), c_u32()
        self.ctx._check(self.ctx.lib.kvx_queue_launch_count(self.handle, ctypes.byref(n), ctypes.byref(c)), "kvx_queue_launch_count")
        return c.value

    def stop(self) -> None:
        if self.handle:
            self.ctx._check(self.ctx.lib.kvx_queue_stop(self.handle), "kvx_queue_stop")
            self.handle = None

    def __enter__(self) -> "Queue":
        return self

    def __exit__(self, *exc) -> None:
        self.stop()


STATIC_MAX_BLOCKS = 4096
STATIC_MAX_REGIONS = 512


def pool_desc(bases: Sequence[int], *, num_blocks: int, heads: int, head_dim: int, layout: int, block_tokens: int = 16,
              elem_bytes: int = 2, block_stride: int = 0, tp_rank: int = 0, tp_size: int = 1, head_stride: int = 0):
    """A kvx_pool_desc for entry points that take descriptors rather than registered pools.  Returns
    (desc, keepalive): keep `keepalive` referenced for as long as `desc` is in use."""
    arr = (c_vp * len(bases))(*[int(b) for b in bases])
    d = PoolDesc(ctypes.sizeof(PoolDesc), len(bases), block_tokens, heads, head_dim, elem_bytes, layout, tp_rank,
                 tp_size, head_stride, num_blocks, block_stride, ctypes.cast(arr, ctypes.POINTER(c_vp)))
    return d, arr


def handoff_static(device: int, src_desc, h_src_table, dst_desc, h_dst_table, n: int, xf: Optional[Xform] = None,
                   flag: Optional[int] = None, flag_val: int = 1, stream=None) -> int:
    """kvx_handoff_static: the fused handoff for callers that may not allocate device memory (engines
    whose HBM is fully budgeted to the KV pool, request-path threads that must not trigger
    cudaMalloc's device-wide synchronisation).  No context, no registered pools: descriptors and host
    block-id lists in, return code out (0 = launched).  At most STATIC_MAX_BLOCKS blocks and
    STATIC_MAX_REGIONS regions per call; calls on one device must not overlap on the GPU."""
    x = xf.c() if xf else None
    return _lib.load().kvx_handoff_static(device, ctypes.byref(src_desc), _ptr(h_src_table), ctypes.byref(dst_desc),
                                          _ptr(h_dst_table), n, ctypes.byref(x) if x else None, flag, flag_val,
                                          _stream(stream))
