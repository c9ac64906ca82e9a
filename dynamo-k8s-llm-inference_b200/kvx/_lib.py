"""ctypes binding of libkvx.so (include/kvx.h).  Loading fails loudly when the library is missing:
there is no Python or CPU fallback for any compute entry."""
from __future__ import annotations

import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# KVX_LIB_PATH points at another build of the same library (A/B runs of kernel variants under tools/).
LIB_PATH = os.environ.get("KVX_LIB_PATH") or os.path.join(os.path.dirname(_HERE), "lib", "libkvx.so")

c_i32, c_u32, c_i64, c_u64 = ctypes.c_int32, ctypes.c_uint32, ctypes.c_int64, ctypes.c_uint64
c_vp, c_sz = ctypes.c_void_p, ctypes.c_size_t
P = ctypes.POINTER


class PoolDesc(ctypes.Structure):
    """struct kvx_pool_desc (include/kvx.h)."""

    _fields_ = [
        ("struct_size", c_u32), ("regions", c_u32), ("block_tokens", c_u32), ("heads", c_u32),
        ("head_dim", c_u32), ("elem_bytes", c_u32), ("layout", c_u32), ("tp_rank", c_u32),
        ("tp_size", c_u32), ("head_stride", c_u32), ("num_blocks", c_u64), ("block_stride", c_u64),
        ("base", P(c_vp)),
    ]


class XformC(ctypes.Structure):
    """struct kvx_xform (include/kvx.h)."""

    _fields_ = [("src_head_off", c_u32), ("dst_head_off", c_u32), ("heads", c_u32), ("stage_layout", c_u32)]


class BatchItem(ctypes.Structure):
    """struct kvx_batch_item (include/kvx.h)."""

    _fields_ = [("h_src_table", c_vp), ("h_dst_table", c_vp), ("n", c_u32), ("reserved", c_u32), ("flag", c_vp),
                ("flag_val", c_u64), ("h_done", c_vp)]


# name → (restype, argtypes): every symbol include/kvx.h declares
PROTOTYPES = {
    "kvx_version": (c_u32, []),
    "kvx_strerror": (ctypes.c_char_p, [c_i32]),
    "kvx_last_error": (ctypes.c_char_p, [c_vp]),
    "kvx_ctx_create": (c_i32, [c_i32, P(c_vp)]),
    "kvx_ctx_destroy": (c_i32, [c_vp]),
    "kvx_ctx_set_option": (c_i32, [c_vp, c_i32, c_i64]),
    "kvx_ctx_get_option": (c_i32, [c_vp, c_i32, P(c_i64)]),
    "kvx_ctx_launch_count": (c_i32, [c_vp, P(c_u64)]),
    "kvx_ctx_sm_count": (c_i32, [c_vp, P(c_i32)]),
    "kvx_mem_alloc": (c_i32, [c_vp, c_u64, P(c_vp)]),
    "kvx_mem_free": (c_i32, [c_vp, c_vp]),
    "kvx_mem_export": (c_i32, [c_vp, c_vp, c_u64, c_vp, P(c_sz)]),
    "kvx_mem_import": (c_i32, [c_vp, c_vp, c_sz, P(c_vp)]),
    "kvx_mem_unmap": (c_i32, [c_vp, c_vp]),
    "kvx_pool_register": (c_i32, [c_vp, P(PoolDesc), P(c_vp)]),
    "kvx_pool_release": (c_i32, [c_vp]),
    "kvx_pool_chunk_bytes": (c_i32, [c_vp, P(c_u64)]),
    "kvx_pool_export": (c_i32, [c_vp, c_vp, P(c_sz)]),
    "kvx_pool_import": (c_i32, [c_vp, c_vp, c_sz, P(c_vp)]),
    "kvx_pool_exportable": (c_i32, [c_vp, P(c_i32)]),
    "kvx_pool_region_base": (c_i32, [c_vp, c_u32, P(c_vp)]),
    "kvx_pool_slice": (c_i32, [c_vp, c_u32, c_u32, P(c_vp)]),
    "kvx_mem_alloc_vmm": (c_i32, [c_vp, c_u64, c_u64, P(c_vp)]),
    "kvx_mem_free_vmm": (c_i32, [c_vp, c_vp]),
    "kvx_pool_export_fds": (c_i32, [c_vp, c_vp, P(c_sz), P(c_i32), P(c_u32)]),
    "kvx_pool_import_fds": (c_i32, [c_vp, c_vp, c_sz, P(c_i32), c_u32, P(c_vp)]),
    "kvx_ctx_flags": (c_i32, [c_vp, P(c_vp), P(c_u32)]),
    "kvx_flag_reset": (c_i32, [c_vp, c_vp, c_u64, c_vp]),
    "kvx_flag_wait": (c_i32, [c_vp, c_vp, c_u64, c_u32, c_vp]),
    "kvx_flag_poll": (c_i32, [c_vp, c_vp, c_u64]),
    "kvx_ctx_check": (c_i32, [c_vp]),
    "kvx_gather": (c_i32, [c_vp, c_vp, c_vp, c_u32, c_vp, P(XformC), c_vp]),
    "kvx_copy": (c_i32, [c_vp, c_vp, c_vp, c_u64, c_vp, c_u64, c_vp]),
    "kvx_scatter": (c_i32, [c_vp, c_vp, c_vp, c_vp, c_u32, P(XformC), c_vp]),
    "kvx_handoff_fused": (c_i32, [c_vp, c_vp, c_vp, c_vp, c_vp, c_u32, P(XformC), c_vp, c_u64, c_vp]),
    "kvx_handoff_host": (c_i32, [c_vp, c_vp, c_vp, c_vp, c_vp, c_u32, P(XformC), c_vp, c_u64, c_vp, c_vp]),
    "kvx_handoff_batch": (c_i32, [c_vp, c_vp, c_vp, P(BatchItem), c_u32, P(XformC), c_vp]),
    "kvx_queue_start": (c_i32, [c_vp, c_vp, c_vp, P(XformC), c_u32, c_u32, P(c_vp)]),
    "kvx_queue_submit": (c_i32, [c_vp, c_vp, c_vp, c_u32, c_vp, c_u64, c_vp, P(c_u64)]),
    "kvx_queue_submit_batch": (c_i32, [c_vp, P(BatchItem), c_u32, P(c_u64)]),
    "kvx_queue_wait": (c_i32, [c_vp, c_u64, c_u32]),
    "kvx_queue_poll": (c_i32, [c_vp, c_u64]),
    "kvx_queue_park": (c_i32, [c_vp]),
    "kvx_queue_launch_count": (c_i32, [c_vp, P(c_u64), P(c_u32)]),
    "kvx_queue_stop": (c_i32, [c_vp]),
    "kvx_handoff_static": (c_i32, [c_i32, P(PoolDesc), c_vp, P(PoolDesc), c_vp, c_u32, P(XformC), c_vp, c_u64, c_vp]),
}

_lib = None


def load() -> ctypes.CDLL:
    """dlopen libkvx.so and bind every prototype.  Raises if the library was not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(kvx has no CPU fallback)"
        )
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in PROTOTYPES.items():
        fn = getattr(lib, name)  # AttributeError = header and library disagree
        fn.restype = res
        fn.argtypes = args
    if lib.kvx_version() >> 16 != 1:
        raise RuntimeError(f"libkvx ABI {lib.kvx_version():#x} is not 1.x")
    _lib = lib
    return lib
