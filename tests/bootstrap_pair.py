"""Two plain OS processes (no torchrun, no torch.distributed): a prefill worker serves its handshake record on a
bootstrap port, a decode worker connects, maps the prefill pool over CUDA IPC, pulls one request announced over
the same connection and releases the blocks.  Used by tests/test_multiproc.py::test_bootstrap_port_pair.

  python bootstrap_pair.py prefill <portfile>      python bootstrap_pair.py decode <portfile>
"""
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "dynamo-k8s-llm-inference_b200"))

import kvx  # noqa: E402
from kvx import bootstrap as b, control as c, synth  # noqa: E402
from kvx.shapes import KVShape  # noqa: E402
from kvx.workers import DecodeWorker, PrefillWorker, TransferParams  # noqa: E402
from oracle import kvx_oracle as o  # noqa: E402

SHAPE = KVShape("test", 2, 8, 64)
KW = dict(regions=SHAPE.regions, head_dim=64, elem_bytes=2, total_kv_heads=8)
NB_P, NB_D, N = 48, 40, 19


def prefill(portfile):
    dev = 0
    torch.cuda.set_device(dev)
    w = PrefillWorker(dev, SHAPE, NB_P, kvx.HND)
    synth.upload(w.kv, o.make_pool(SHAPE.layers, 8, 64, NB_P, o.HND, seed=123).buf)
    w.expect(5)
    torch.cuda.synchronize()
    leases = c.LeaseTable(lease_seconds=60)
    ids = w.blocks.allocate("req-1", N)
    leases.grant("req-1")
    peers = []
    srv = b.BootstrapServer(c.encode_metadata(w.metadata(), **KW), lambda rec, addr: peers.append(c.decode_metadata(rec, expect_compat=c.compat_hash(abi=0x00010000, **KW))),
                            on_frame=lambda frame, addr: c.encode_params(TransferParams("req-1", ids, 0, 1)))
    with open(portfile + ".tmp", "w") as f:
        f.write(str(srv.address[1]))
    os.replace(portfile + ".tmp", portfile)
    deadline = time.time() + 120
    while time.time() < deadline and not w.poll(5, 77):   # release flag written by the decode worker's kernel
        time.sleep(0.01)
    assert w.poll(5, 77), "the decode worker never released the blocks"
    assert leases.release("req-1")
    w.blocks.free("req-1")
    assert peers and peers[0].role == "decode" and not srv.errors, srv.errors
    srv.close()
    w.close()
    print("prefill ok")


def decode(portfile):
    dev = torch.cuda.device_count() - 1
    torch.cuda.set_device(dev)
    w = DecodeWorker(dev, SHAPE, NB_D, kvx.NHD)
    synth.fill_canary(w.kv)
    torch.cuda.synchronize()
    deadline = time.time() + 120
    while not os.path.exists(portfile) and time.time() < deadline:
        time.sleep(0.05)
    port = int(open(portfile).read())
    cli = b.BootstrapClient("127.0.0.1", port, c.encode_metadata(w.metadata(), **KW))
    w.add_remote(0, c.decode_metadata(cli.peer_record, expect_compat=c.compat_hash(abi=0x00010000, **KW)))
    params = c.decode_params(cli.send(c.encode_heartbeat(["req-1"]), expect_reply=True))   # "what do you hold for me?"
    local = w.blocks.allocate(params.request_id, N)
    out = c.guarded_pull(w, 0, params, local, notify_slot=5, value=77)
    torch.cuda.synchronize()
    w.ctx.check()
    assert out.ok, out.detail
    want = o.make_pool(SHAPE.layers, 8, 64, NB_D, o.NHD, fill="canary")
    o.handoff(o.make_pool(SHAPE.layers, 8, 64, NB_P, o.HND, seed=123), params.remote_block_ids, want, local)
    assert np.array_equal(w.kv.to_numpy(), want.buf), "bytes on the decode GPU differ from the oracle"
    cli.close()
    w.close()
    print("decode ok")


if __name__ == "__main__":
    {"prefill": prefill, "decode": decode}[sys.argv[1]](sys.argv[2])
