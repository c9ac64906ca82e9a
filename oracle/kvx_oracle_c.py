"""ctypes binding of oracle/kvx_oracle.c (checker and timed CPU baseline; test infrastructure only)."""
from __future__ import annotations

import ctypes
import os
import subprocess
from typing import Optional, Sequence

import numpy as np

from . import kvx_oracle as o

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "libkvx_oracle.so")


class _Side(ctypes.Structure):
    _fields_ = [
        ("base", ctypes.POINTER(ctypes.c_void_p)),
        ("table", ctypes.POINTER(ctypes.c_int32)),
        ("base0", ctypes.c_void_p),
        ("block_stride", ctypes.c_uint64),
        ("num_blocks", ctypes.c_uint64),
        ("layout", ctypes.c_uint32),
        ("heads", ctypes.c_uint32),
        ("head_off", ctypes.c_uint32),
        ("block_tokens", ctypes.c_uint32),
    ]


def build(force: bool = False) -> str:
    if force or not os.path.exists(_SO):
        subprocess.check_call(["make", "-C", _HERE], stdout=subprocess.DEVNULL)
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = ctypes.CDLL(build())
        _lib.kvxo_xfer.restype = ctypes.c_int
        _lib.kvxo_xfer.argtypes = [ctypes.POINTER(_Side), ctypes.POINTER(_Side)] + [ctypes.c_uint32] * 6
    return _lib


def _paged(pool: o.Pool, table: np.ndarray, head_off: int, keep: list) -> _Side:
    base_addr = pool.buf.ctypes.data
    arr = (ctypes.c_void_p * pool.regions)(*[base_addr + off for off in pool.region_off])
    tab = np.ascontiguousarray(table, dtype=np.int32)
    keep += [arr, tab]
    return _Side(
        ctypes.cast(arr, ctypes.POINTER(ctypes.c_void_p)),
        tab.ctypes.data_as(ctypes.POINTER(ctypes.c_int32)),
        None,
        pool.block_stride,
        pool.num_blocks,
        pool.layout,
        pool.heads,
        head_off,
        pool.block_tokens,
    )


def _contig(image: np.ndarray, layout: int, heads: int) -> _Side:
    return _Side(None, None, image.ctypes.data, 0, 0, layout, heads, 0, 0)


def handoff(src: o.Pool, src_table: Sequence[int], dst: o.Pool, dst_table: Sequence[int],
            xf: Optional[o.Xform] = None, threads: int = 1) -> int:
    """Direct paged→paged move with the C loop (same semantics as kvx_oracle.handoff, equal-length tables;
    with unequal block sizes the granule is the smaller block, see kvx_oracle.handoff_regroup)."""
    so, do, hc, _ = o._resolve(xf, src.heads, dst.heads, src.layout)
    keep: list = []
    s = _paged(src, np.asarray(src_table), so, keep)
    d = _paged(dst, np.asarray(dst_table), do, keep)
    N = min(src.block_tokens, dst.block_tokens)
    n = min(len(src_table) * (src.block_tokens // N), len(dst_table) * (dst.block_tokens // N))
    return lib().kvxo_xfer(ctypes.byref(s), ctypes.byref(d), src.regions, n, N, hc, src.row_bytes, threads)


def gather(src: o.Pool, table: Sequence[int], xf: Optional[o.Xform] = None, threads: int = 1) -> np.ndarray:
    so, _, hc, sl = o._resolve(xf, src.heads, None, src.layout)
    n = len(table)
    image = np.empty((src.regions, n, src.block_tokens * hc * src.row_bytes), dtype=np.uint8)
    keep: list = []
    s = _paged(src, np.asarray(table), so, keep)
    d = _contig(image, sl, hc)
    rc = lib().kvxo_xfer(ctypes.byref(s), ctypes.byref(d), src.regions, n, src.block_tokens, hc, src.row_bytes, threads)
    if rc:
        raise RuntimeError(f"kvxo_xfer rc={rc}")
    return image


def scatter(image: np.ndarray, dst: o.Pool, table: Sequence[int], xf: Optional[o.Xform] = None, threads: int = 1) -> None:
    _, do, hc, sl = o._resolve(xf, None, dst.heads, dst.layout)
    n = len(table)
    image = np.ascontiguousarray(image)
    keep: list = []
    s = _contig(image, sl, hc)
    d = _paged(dst, np.asarray(table), do, keep)
    rc = lib().kvxo_xfer(ctypes.byref(s), ctypes.byref(d), dst.regions, n, dst.block_tokens, hc, dst.row_bytes, threads)
    if rc:
        raise RuntimeError(f"kvxo_xfer rc={rc}")
