"""Two plain processes, one Unix-domain socket: a prefill worker whose KV pool lives in CUDA virtual-memory-management
memory (several physical allocations mapped back to back, the shape of an expandable segment) hands the pool to a decode
worker as file descriptors (SCM_RIGHTS); the decode worker maps it and pulls a request out of it.
  python tests/vmm_pair.py prefill <fd-of-socket>  |  python tests/vmm_pair.py decode <fd-of-socket>"""
import os
import socket
import struct
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "dynamo-k8s-llm-inference_b200"))

import kvx  # noqa: E402
from kvx import synth  # noqa: E402
from oracle import kvx_oracle as o  # noqa: E402

L, H, D, NB, N_REQ = 3, 8, 128, 700, 40   # pool = 6 regions x 700 blocks x 32 KiB = 131 MiB: many 2 MiB-granular pieces


def send_msg(sock, payload: bytes, fds=()):
    socket.send_fds(sock, [struct.pack("<I", len(payload))], list(fds))
    sock.sendall(payload)


def recv_msg(sock, max_fds=0):
    hdr, fds, _, _ = socket.recv_fds(sock, 4, max_fds)
    n = struct.unpack("<I", hdr)[0]
    buf = b""
    while len(buf) < n:
        buf += sock.recv(n - len(buf))
    return buf, list(fds)


def main():
    role, sock = sys.argv[1], socket.socket(fileno=int(sys.argv[2]))
    dev = (0 if role == "prefill" else 1) % torch.cuda.device_count()
    torch.cuda.set_device(dev)
    ctx = kvx.Context(dev)
    src_host = o.make_pool(L, H, D, NB, o.HND, seed=21)
    ts, td = o.make_table(NB, N_REQ, 1), o.make_table(NB + 3, N_REQ, 2)
    if role == "prefill":
        nbytes = src_host.buf.nbytes
        engine_pool = os.environ.get("PYTORCH_CUDA_ALLOC_CONF", "").find("expandable_segments:True") >= 0
        if engine_pool:   # the KV pool as an engine allocates it: a torch tensor in an expandable segment
            keep = torch.from_numpy(src_host.buf).cuda()
            ptr = keep.data_ptr()
        else:
            ptr = ctx.mem_alloc_vmm(nbytes, piece_bytes=16 << 20)   # 9 physical allocations, contiguous in the address space
            synth.as_torch(ptr, nbytes, dev).copy_(torch.from_numpy(src_host.buf))
        torch.cuda.synchronize()
        pool = ctx.register_pool([ptr + off for off in src_host.region_off], num_blocks=NB, heads=H, head_dim=D, layout=kvx.HND)
        assert not pool.exportable_across_processes      # no CUDA IPC handle for this memory ...
        blob, fds = pool.export_fds()                    # ... but file descriptors
        assert engine_pool or len(fds) == 9, fds
        send_msg(sock, blob, fds)
        for fd in fds:
            os.close(fd)
        done, _ = recv_msg(sock)
        assert done == b"landed"
        pool.release()
        if not engine_pool:
            ctx.mem_free_vmm(ptr)
        print(f"prefill ok on cuda:{dev}: {len(fds)} allocations shared as file descriptors", flush=True)
    else:
        blob, fds = recv_msg(sock, max_fds=64)
        peer = ctx.import_pool_fds(blob, fds)
        for fd in fds:
            os.close(fd)
        dst_host = o.make_pool(L, H, D, NB + 3, o.NHD, fill="canary")
        t = torch.from_numpy(dst_host.buf.copy()).cuda()
        mine = ctx.register_pool([t.data_ptr() + off for off in dst_host.region_off], num_blocks=NB + 3, heads=H, head_dim=D, layout=kvx.NHD)
        ctx.handoff_host(peer, ts, mine, td, N_REQ)
        torch.cuda.synchronize()
        ctx.check()
        o.handoff(src_host, ts, dst_host, td)
        assert np.array_equal(t.cpu().numpy(), dst_host.buf), "bytes pulled out of the VMM pool differ from the oracle"
        peer.release(); mine.release()
        send_msg(sock, b"landed")
        print(f"decode ok on cuda:{dev}", flush=True)
    ctx.close()


if __name__ == "__main__":
    main()
