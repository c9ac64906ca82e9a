"""Two-process P→D check (run under torchrun, world_size 2).  Rank 0 = prefill worker on GPU 0, rank 1 =
decode worker on GPU 1 (or both on GPU 0 when only one device is visible).  The decode rank verifies the
bytes that landed against the oracle.  With --cpu-dry-run no GPU is touched: only the control records
travel (gloo), which is what the CPU test exercises."""
import argparse
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "dynamo-k8s-llm-inference_b200"))

from oracle import kvx_oracle as o  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cpu-dry-run", action="store_true")
    ap.add_argument("--p-tp", type=int, default=1)
    args = ap.parse_args()
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from kvx.shapes import KVShape
    from kvx.tp import plan_edges
    from kvx.workers import BlockAllocator, TransferParams, align_tables

    shape = KVShape("test", 2, 8, 64)
    n, nb_p, nb_d = 21, 40, 50
    # control plane: decode allocates, prefill announces its block ids (kv_transfer_params)
    if rank == 0:
        alloc = BlockAllocator(nb_p, seed=5)
        src_ids = alloc.allocate("req-0", n)
        params = TransferParams("req-0", src_ids, 0, 1)
    else:
        alloc = BlockAllocator(nb_d, seed=6)
        dst_ids = alloc.allocate("req-0", n - 4)  # partial prefix hit: 4 blocks already cached on D
        params = None
    box = [params]
    dist.broadcast_object_list(box, src=0)
    params = box[0]
    if args.cpu_dry_run:
        if rank == 1:
            s, d = align_tables(params.remote_block_ids, dst_ids)
            assert len(s) == len(d) == n - 4 and np.array_equal(s, params.remote_block_ids[4:])
            assert 0 not in d and len(set(d.tolist())) == len(d)
        edges = plan_edges(4, 2, 8)
        assert [(e.p_rank, e.d_rank, e.dst_head_off) for e in edges] == [(0, 0, 0), (1, 0, 2), (2, 1, 0), (3, 1, 2)]
        dist.barrier()
        print(f"rank {rank} dry-run ok")
        dist.destroy_process_group()
        return

    import kvx
    from kvx import synth
    from kvx.workers import DecodeWorker, PrefillWorker

    dev = rank % torch.cuda.device_count()   # two devices when the box has them: then every byte crosses NVLink
    torch.cuda.set_device(dev)
    print(f"rank {rank} uses cuda:{dev} of {torch.cuda.device_count()} visible ({torch.cuda.get_device_properties(dev).uuid})", flush=True)
    src_host = o.make_pool(shape.layers, 8, 64, nb_p, o.HND, seed=77)
    if rank == 0:
        w = PrefillWorker(dev, shape, nb_p, kvx.HND)
        synth.upload(w.kv, src_host.buf)
    else:
        w = DecodeWorker(dev, shape, nb_d, kvx.NHD)
        synth.fill_canary(w.kv)
    w.expect(3)
    torch.cuda.synchronize()
    got = [None, None]
    dist.all_gather_object(got, w.metadata())
    w.add_remote(0, got[1 - rank])
    ids = [dst_ids if rank == 1 else None]
    dist.broadcast_object_list(ids, src=1)
    want = o.make_pool(shape.layers, 8, 64, nb_d, o.NHD, fill="canary")
    o.handoff(src_host, params.remote_block_ids, want, ids[0])

    # 1) pull: the decode rank reads the prefill pool over IPC/NVLink and releases the blocks
    if rank == 1:
        w.pull(0, params.remote_block_ids, dst_ids, notify_slot=3, value=9)
        torch.cuda.synchronize()
        w.ctx.check()
        assert np.array_equal(w.kv.to_numpy(), want.buf), "pull: bytes on the decode GPU differ from the oracle"
        print("pull ok")
    dist.barrier()
    if rank == 0:
        w.wait(3, 9, 5000)  # release notification written by the decode rank's kernel
        torch.cuda.synchronize()
        w.ctx.check()
        assert w.poll(3, 9)
        alloc.free("req-0")  # every consumer released the blocks: give them back
    dist.barrier()

    # 2) push: same bytes the other way round (prefill GPU stores into the decode pool)
    if rank == 1:
        synth.fill_canary(w.kv)
        w.expect(4)
        torch.cuda.synchronize()
    dist.barrier()
    if rank == 0:
        w.push(0, params.remote_block_ids, ids[0], slot=4, value=11)
        torch.cuda.synchronize()
        w.ctx.check()
    dist.barrier()
    if rank == 1:
        w.wait(4, 11, 5000)
        torch.cuda.synchronize()
        w.ctx.check()
        assert np.array_equal(w.kv.to_numpy(), want.buf), "push: bytes on the decode GPU differ from the oracle"
        print("pair ok")
    dist.barrier()

    # 3) a scheduler step's worth of requests in one launch: three requests (one with a prefix hit), each
    #    releasing its prefill blocks through its own flag in the prefill worker's memory
    counts = [5, 9, 3]
    if rank == 0:
        reqs_src = [alloc.allocate(f"b{k}", c) for k, c in enumerate(counts)]
    else:
        synth.fill_canary(w.kv)
        reqs_dst = [alloc.allocate(f"b{k}", c - (2 if k == 1 else 0)) for k, c in enumerate(counts)]
        torch.cuda.synchronize()
    for k in range(3):
        if rank == 0:
            w.expect(10 + k)
    torch.cuda.synchronize()
    box = [reqs_src if rank == 0 else None]
    dist.broadcast_object_list(box, src=0)
    if rank == 1:
        words = torch.zeros(3, dtype=torch.int64).pin_memory()
        w.pull_many(0, [(box[0][k], reqs_dst[k], 10 + k, 50 + k, words[k:k + 1]) for k in range(3)])
        torch.cuda.synchronize()
        w.ctx.check()
        want3 = o.make_pool(shape.layers, 8, 64, nb_d, o.NHD, fill="canary")
        for k in range(3):
            s_ids, d_ids = align_tables(box[0][k], reqs_dst[k])
            o.handoff(src_host, s_ids, want3, d_ids)
        assert np.array_equal(w.kv.to_numpy(), want3.buf), "pull_many: bytes on the decode GPU differ from the oracle"
        assert words.tolist() == [50, 51, 52]
    dist.barrier()
    if rank == 0:
        for k in range(3):
            assert w.poll(10 + k, 50 + k), f"request {k} was not released"
        print("batch ok")
    dist.barrier()
    w.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
