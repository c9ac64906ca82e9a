"""A host program in plain C drives libkvx.so through include/kvx.h alone (no Python, no torch, no
CUDA headers): the binding a non-Python engine would write.  tests/c_client/kvx_c_client.c."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB_DIR = os.path.join(ROOT, "dynamo-k8s-llm-inference_b200", "lib")
ORACLE = os.path.join(ROOT, "oracle", "_build", "libkvx_oracle.so")
CUDART = os.environ.get("KVX_CUDART", "/usr/local/cuda/lib64/libcudart.so")


@pytest.fixture(scope="module")
def client(tmp_path_factory):
    import __graft_entry__ as g

    g.build_lib(force=False)
    g.build_oracle(force=False)
    exe = str(tmp_path_factory.mktemp("c_client") / "kvx_c_client")
    subprocess.check_call(["gcc", "-std=c11", "-O2", "-Wall", "-Wextra", "-Werror", "-I", os.path.join(ROOT, "include"),
                           "-o", exe, os.path.join(ROOT, "tests", "c_client", "kvx_c_client.c"),
                           "-L", LIB_DIR, "-lkvx", "-ldl", f"-Wl,-rpath,{LIB_DIR}"])
    return exe


def test_c_client_builds_and_fails_loudly_without_a_device(client):
    """The header compiles as C11 with -Wall -Wextra -Werror, the library links from C, and without a
    CUDA device kvx_ctx_create reports KVX_ECUDA (exit code 3) instead of falling back to the CPU."""
    import torch

    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present; the GPU test covers this binary")
    r = subprocess.run([client, ORACLE, CUDART], capture_output=True, text=True, timeout=120)
    assert r.returncode == 3, r.stderr
    assert "kvx_ctx_create" in r.stderr


@pytest.mark.gpu
def test_c_client_bit_exact(client):
    r = subprocess.run([client, ORACLE, CUDART], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "C client OK" in r.stdout
