#!/usr/bin/env python
"""The smallest complete use of the library from Python: two workers, one request.

  python examples/one_request.py            # one GPU is enough (both workers on it); with two, the pull crosses NVLink

A prefill worker owns an HND pool holding a finished prompt's KV, a decode worker owns an NHD pool.  They shake
hands (each maps the other's pool and flags), the decode worker allocates fresh blocks and pulls the prompt's
blocks — layout transform in flight — and the prefill worker sees its blocks released.
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "dynamo-k8s-llm-inference_b200"))

import torch

import kvx
from kvx import synth
from kvx.workers import DecodeWorker, PrefillWorker


def main() -> int:
    if not torch.cuda.is_available():
        print("needs a CUDA device: kvx has no CPU fallback", file=sys.stderr)
        return 1
    shape = kvx.KV_SHAPES["meta-llama/Llama-3.1-8B"]
    isl = 4096
    n = shape.blocks_for(isl)
    d_dev = 1 if torch.cuda.device_count() > 1 else 0
    prefill = PrefillWorker(0, shape, 2 * n + 1, kvx.HND)       # vLLM's NIXL connector keeps HND pools
    decode = DecodeWorker(d_dev, shape, 3 * n + 1, kvx.NHD)     # this decode engine wants NHD
    synth.fill_random(prefill.kv, 1)                             # stands in for the prefill forward pass
    synth.fill_canary(decode.kv)
    prefill.expect(0)                                            # release flag slot 0 starts at 0
    torch.cuda.synchronize()

    decode.add_remote(0, prefill.metadata())                     # handshake (kvx.bootstrap does this over TCP)
    src_ids = prefill.blocks.allocate("req-1", n)                # the blocks prefill wrote
    dst_ids = decode.blocks.allocate("req-1", n)                 # fresh slots on the decode side

    torch.cuda.set_device(d_dev)
    stream = torch.cuda.current_stream(d_dev)
    done = torch.zeros(1, dtype=torch.int64).pin_memory()        # completion word, written by the kernel
    warm = decode.blocks.allocate("warm-up", n)                  # first launch loads the kernel: not timed
    decode.pull(0, src_ids, warm, stream=stream)
    torch.cuda.synchronize(d_dev)
    decode.blocks.free("warm-up")
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record(stream)
    decode.pull(0, src_ids, dst_ids, notify_slot=0, value=1, h_done=done, stream=stream)
    t1.record(stream)
    t1.synchronize()
    decode.ctx.check()

    nbytes = n * shape.regions * shape.chunk_bytes()
    ms = t0.elapsed_time(t1)
    assert int(done[0]) == 1 and prefill.poll(0, 1)              # landed on D, released on P
    # spot check: block 0 of layer 0's K, token 3, head 5 — HND on P, NHD on D
    N, H, D = shape.block_tokens, shape.kv_heads, shape.head_dim
    p = prefill.kv.tensor().view(torch.int16).view(shape.regions, -1, H, N, D)[0, int(src_ids[0]), 5, 3].cpu()
    d = decode.kv.tensor().view(torch.int16).view(shape.regions, -1, N, H, D)[0, int(dst_ids[0]), 3, 5].cpu()
    assert torch.equal(p, d)
    print(f"handed off {n} blocks ({nbytes / 2**20:.0f} MiB) in {ms:.3f} ms = {nbytes / ms / 1e6:.0f} GB/s "
          f"({'NVLink' if d_dev else 'one GPU'}); prefill blocks released")
    prefill.blocks.free("req-1")
    decode.close()
    prefill.close()
    return 0


if __name__ == "__main__":
    sys.exit(main())
