"""N>1 coverage: world_size-2 gloo run of the host-side control plane on CPU, and (GPU) the real
two-process P→D transfer over CUDA IPC / NVLink."""
import os
import subprocess
import sys

import numpy as np
import pytest

from conftest import ROOT


def _torchrun(extra, port, timeout=240):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "tests", "pair_worker.py"), *extra]
    return subprocess.run(cmd, capture_output=True, text=True, timeout=timeout)


def test_control_plane_world2_gloo():
    r = _torchrun(["--cpu-dry-run"], 29611)
    assert r.returncode == 0, r.stdout + r.stderr
    assert r.stdout.count("dry-run ok") == 2


def test_block_allocator_and_alignment():
    from kvx.workers import BlockAllocator, align_tables

    a = BlockAllocator(17, seed=1)
    x = a.allocate("a", 10)
    y = a.allocate("b", 6)
    assert 0 not in x and 0 not in y and not set(x) & set(y) and a.free_blocks == 0
    with pytest.raises(MemoryError):
        a.allocate("c", 1)
    a.free("a")
    assert a.free_blocks == 10
    with pytest.raises(KeyError):
        a.allocate("b", 1)
    s, d = align_tables(np.arange(1, 9, dtype=np.int32), np.array([4, 2, 9], dtype=np.int32))
    assert s.tolist() == [6, 7, 8] and d.tolist() == [4, 2, 9]
    s, d = align_tables(np.arange(1, 9, dtype=np.int32), np.array([], dtype=np.int32))
    assert len(s) == 0 and len(d) == 0
    with pytest.raises(ValueError):
        align_tables(np.arange(3), np.arange(4))


def test_tp_plan_matches_golden(golden_dir):
    import json

    from kvx.tp import plan_edges

    rows = json.load(open(os.path.join(golden_dir, "tp_mapping.json")))["rows"]
    for row in rows:
        H = row["total_kv_heads"]
        edges = [e for e in plan_edges(row["p_tp"], row["d_tp"], H) if e.d_rank == row["d_rank"]]
        assert [e.p_rank for e in edges] == row["source_ranks"], row
        hp, hd = max(1, H // row["p_tp"]), max(1, H // row["d_tp"])
        for e in edges:
            if row["d_tp"] > row["p_tp"]:
                assert e.src_head_off == row["rank_offset_factor"] * hd and e.dst_head_off == 0 and e.heads == hd
            else:
                assert e.dst_head_off == row["rank_to_slot"][str(e.p_rank)] * hp and e.src_head_off == 0


@pytest.mark.gpu
def test_two_process_handoff_over_ipc():
    import torch

    if not torch.cuda.is_available():
        pytest.skip("no GPU")
    r = _torchrun([], 29612)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "pull ok" in r.stdout and "pair ok" in r.stdout and "batch ok" in r.stdout
    if torch.cuda.device_count() >= 2:   # with two devices every byte crossed NVLink: the ranks say which GPU they used
        assert "rank 0 uses cuda:0" in r.stdout and "rank 1 uses cuda:1" in r.stdout, r.stdout[-2000:]


@pytest.mark.gpu
def test_in_process_peer_handoff():
    """P pool on cuda:0, D pool on cuda:1 in one process: import falls back to peer access."""
    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import kvx
    from kvx import synth
    from kvx.shapes import KVShape
    from kvx.workers import DecodeWorker, PrefillWorker
    from oracle import kvx_oracle as o

    shape = KVShape("t", 2, 8, 128)
    src_host = o.make_pool(2, 8, 128, 30, o.NHD, seed=3)
    p = PrefillWorker(0, shape, 30, kvx.NHD)
    d = DecodeWorker(1, shape, 33, kvx.HND)
    synth.upload(p.kv, src_host.buf)
    synth.fill_canary(d.kv)
    d.expect(0)
    torch.cuda.synchronize(0); torch.cuda.synchronize(1)
    p.add_remote(0, d.metadata())
    ts, td = o.make_table(30, 20, 1), o.make_table(33, 20, 2)
    p.push(0, ts, td, slot=0, value=5)
    d.wait(0, 5, 5000)
    torch.cuda.synchronize(0); torch.cuda.synchronize(1)
    want = o.make_pool(2, 8, 128, 33, o.HND, fill="canary")
    o.handoff(src_host, ts, want, td)
    assert np.array_equal(d.kv.to_numpy(), want.buf)
    # and the pull direction into fresh slots, releasing the prefill blocks
    synth.fill_canary(d.kv)
    p.expect(1)
    torch.cuda.synchronize(0); torch.cuda.synchronize(1)
    d.add_remote(0, p.metadata())
    d.pull(0, ts, td, notify_slot=1, value=6)
    p.wait(1, 6, 5000)
    torch.cuda.synchronize(0); torch.cuda.synchronize(1)
    assert np.array_equal(d.kv.to_numpy(), want.buf)
    p.close(); d.close()


@pytest.mark.gpu
def test_bootstrap_port_pair(tmp_path):
    """Two independent processes (no torchrun): handshake over the bootstrap port, pull over CUDA IPC, release,
    lease bookkeeping on the prefill side — the whole control + data path of one request."""
    script = os.path.join(ROOT, "tests", "bootstrap_pair.py")
    portfile = str(tmp_path / "port")
    p = subprocess.Popen([sys.executable, script, "prefill", portfile], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    d = subprocess.Popen([sys.executable, script, "decode", portfile], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    try:
        d_out, _ = d.communicate(timeout=240)
        p_out, _ = p.communicate(timeout=60)
    finally:
        for proc in (p, d):
            if proc.poll() is None:
                proc.kill()
    assert d.returncode == 0 and "decode ok" in d_out, d_out[-3000:]
    assert p.returncode == 0 and "prefill ok" in p_out, p_out[-3000:]


@pytest.mark.gpu
@pytest.mark.parametrize("allocator", ["kvx_mem_alloc_vmm", "torch_expandable_segments"])
def test_vmm_pool_shared_as_file_descriptors(allocator):
    """A KV pool in CUDA virtual-memory-management memory (no CUDA IPC handle) crosses to another process as file
    descriptors over a Unix-domain socket, and a request is pulled out of it bit for bit — with the library's own VMM
    allocator, and with the pool held in a torch tensor under PYTORCH_CUDA_ALLOC_CONF=expandable_segments:True, which is
    how an engine holds its KV cache."""
    import socket

    env = dict(os.environ)
    if allocator == "torch_expandable_segments":
        env["PYTORCH_CUDA_ALLOC_CONF"] = "expandable_segments:True"

    a, b = socket.socketpair(socket.AF_UNIX, socket.SOCK_STREAM)
    script = os.path.join(ROOT, "tests", "vmm_pair.py")
    p = subprocess.Popen([sys.executable, script, "prefill", str(a.fileno())], pass_fds=[a.fileno()], stdout=subprocess.PIPE,
                         stderr=subprocess.STDOUT, text=True, env=env)
    d = subprocess.Popen([sys.executable, script, "decode", str(b.fileno())], pass_fds=[b.fileno()], stdout=subprocess.PIPE,
                         stderr=subprocess.STDOUT, text=True)
    a.close(); b.close()
    try:
        d_out, _ = d.communicate(timeout=240)
        p_out, _ = p.communicate(timeout=60)
    finally:
        for proc in (p, d):
            if proc.poll() is None:
                proc.kill()
    assert d.returncode == 0 and "decode ok" in d_out, d_out[-3000:]
    assert p.returncode == 0 and "prefill ok" in p_out, p_out[-3000:]
