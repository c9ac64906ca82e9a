"""CPU-side checks of the drop-in boundary: the library loads, exports every symbol the header
declares, and refuses to compute without a GPU (no CPU fallback)."""
import ctypes
import os
import re

import pytest

from conftest import ROOT


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "kvx.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(kvx_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_exported():
    from kvx import _lib

    lib = _lib.load()
    names = _declared_symbols()
    assert len(names) >= 25
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/kvx.h but not exported by libkvx.so"
    # and the binding covers exactly the header
    assert sorted(_lib.PROTOTYPES) == names


def test_version_and_strerror():
    from kvx import _lib

    lib = _lib.load()
    assert lib.kvx_version() == 0x00010000
    assert lib.kvx_strerror(0) == b"ok"
    assert b"unsupported" in lib.kvx_strerror(-4)
    assert b"unknown" in lib.kvx_strerror(-99)


def test_struct_sizes_match_header():
    from kvx._lib import BatchItem, PoolDesc, XformC

    assert ctypes.sizeof(PoolDesc) == 64
    assert ctypes.sizeof(XformC) == 16
    assert ctypes.sizeof(BatchItem) == 48
    assert PoolDesc.head_stride.offset == 36 and PoolDesc.num_blocks.offset == 40


def test_no_cpu_fallback():
    import torch

    if torch.cuda.is_available():
        pytest.skip("GPU present")
    import kvx

    with pytest.raises(kvx.KvxError) as e:
        kvx.Context(0)
    assert e.value.code == -2
    lib = kvx._lib.load()
    # null-context calls fail with EINVAL, they never compute
    assert lib.kvx_gather(None, None, None, 1, None, None, None) == -1
    assert lib.kvx_copy(None, None, None, 16, None, 0, None) == -1


def test_product_never_touches_oracle():
    """The oracle is a checker: nothing under the package may import, link or mention it."""
    pkg = os.path.join(ROOT, "dynamo-k8s-llm-inference_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp", ".sh")):
                src = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in src, (dirpath, f)
