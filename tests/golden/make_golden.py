"""Generates tests/golden/*.npz|json from the on-box vLLM 0.22 helpers (run in the authoring container;
the fixtures are committed, this script is how they were made).

vLLM is NOT the reference (the reference holds no code for this path); it is the code
the reference's manifests select at run time (examples/deploy/vllm/disagg.yaml:37,57) and the only
on-box authority for what "NHD/HND", "block id" and "heterogeneous-TP slot" mean in bytes.

  layout_hnd_to_nhd.npz   raw HND blocks landed in an NHD-shaped cache, then
                          kv_postprocess_layout_on_receive (kv_connector/utils.py:255-277)
  tp_mapping.json         compute_tp_mapping (v1/nixl/tp_mapping.py:65-142) known answers
"""
import json
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))


def layout_fixture():
    from vllm.distributed.kv_transfer.kv_connector.utils import kv_postprocess_layout_on_receive

    rng = np.random.Generator(np.random.PCG64(20261106))
    L, H, N, D = 2, 4, 16, 8  # int16 elements: row = 16 bytes
    nb_p, nb_d, n = 11, 13, 5
    regions = 2 * L
    # P pool: every region [nb_p][H][N][D] (HND), random 16-bit patterns
    p = rng.integers(0, 65536, size=(regions, nb_p, H, N, D), dtype=np.uint16)
    p_table = (1 + rng.permutation(nb_p - 1)[:n]).astype(np.int32)
    d_table = (1 + rng.permutation(nb_d - 1)[:n]).astype(np.int32)
    # D pool: every region [nb_d][N][H][D] (NHD), canary
    d = np.full((regions, nb_d, N, H, D), 0xC0DE, dtype=np.uint16)
    expect = d.copy()
    for r in range(regions):
        cache = torch.from_numpy(expect[r].view(np.int16))  # NHD-shaped cache of region r
        # the transport lands P's raw (HND-ordered) block bytes in D's blocks ...
        flat = cache.view(nb_d, -1)
        flat[torch.from_numpy(d_table.astype(np.int64))] = torch.from_numpy(
            p[r].view(np.int16).reshape(nb_p, -1)[p_table.astype(np.int64)]
        )
        # ... and vLLM permutes them in place to the local layout
        kv_postprocess_layout_on_receive(cache, torch.from_numpy(d_table.astype(np.int64)))
    np.savez_compressed(
        os.path.join(HERE, "layout_hnd_to_nhd.npz"),
        p_pool=p, p_table=p_table, d_table=d_table, d_expect=expect,
        shape=np.array([L, H, N, D, nb_p, nb_d, n], dtype=np.int64),
    )
    print("layout_hnd_to_nhd.npz written")


def blocksize_fixture():
    """P blocks of 4 tokens (HND) land in D blocks of 16 tokens: D kept HND
    (kv_postprocess_blksize_on_receive, utils.py:224-252) and D in NHD
    (kv_postprocess_blksize_and_layout_on_receive, utils.py:280-299)."""
    from vllm.distributed.kv_transfer.kv_connector.utils import (
        kv_postprocess_blksize_and_layout_on_receive, kv_postprocess_blksize_on_receive)

    rng = np.random.Generator(np.random.PCG64(20261107))
    L, H, D, Np, ratio = 1, 4, 8, 4, 4
    Nd = Np * ratio
    nb_p, nb_d, n_d = 40, 9, 3          # 3 decode blocks = 12 prefill blocks
    n_p = n_d * ratio
    regions = 2 * L
    p = rng.integers(0, 65536, size=(regions, nb_p, H, Np, D), dtype=np.uint16)      # HND small blocks
    p_table = (1 + rng.permutation(nb_p - 1)[:n_p]).astype(np.int32)
    d_table = (1 + rng.permutation(nb_d - 1)[:n_d]).astype(np.int32)
    idx = torch.from_numpy(d_table.astype(np.int64))
    out = {}
    for name in ("hnd", "nhd"):
        d = np.full((regions, nb_d, Nd * H * D), 0xC0DE, dtype=np.uint16)
        for r in range(regions):
            flat = torch.from_numpy(d[r].view(np.int16))
            # raw landing: the k prefill blocks of one decode block, back to back
            src = torch.from_numpy(p[r].view(np.int16).reshape(nb_p, -1)[p_table.astype(np.int64)]).reshape(n_d, -1)
            flat[idx] = src
            if name == "hnd":
                # logical view [nb, N, H, D] whose physical order is HND
                cache = flat.view(nb_d, H, Nd, D).permute(0, 2, 1, 3)
                kv_postprocess_blksize_on_receive(cache, idx, ratio)
            else:
                cache = flat.view(nb_d, Nd, H, D)
                kv_postprocess_blksize_and_layout_on_receive(cache, idx, ratio)
        out[name] = d
    np.savez_compressed(os.path.join(HERE, "blocksize_ratio.npz"), p_pool=p, p_table=p_table, d_table=d_table,
                        d_expect_hnd=out["hnd"], d_expect_nhd=out["nhd"],
                        shape=np.array([L, H, D, Np, ratio, nb_p, nb_d, n_p], dtype=np.int64))
    print("blocksize_ratio.npz written")


def tp_fixture():
    from vllm.distributed.kv_transfer.kv_connector.utils import TransferTopology
    from vllm.distributed.kv_transfer.kv_connector.v1.nixl.tp_mapping import compute_tp_mapping
    from vllm.v1.kv_cache_interface import FullAttentionSpec
    import inspect

    rows = []
    cases = [(2, 4, 8), (4, 2, 8), (2, 1, 8), (1, 4, 8), (1, 1, 8), (8, 2, 8), (4, 4, 8), (1, 2, 8), (8, 8, 8), (2, 8, 8),
             # tp > kv heads: KV is replicated across ranks (kv_connector/utils.py:525-538)
             (16, 8, 8), (8, 16, 8), (16, 16, 8), (4, 16, 8), (16, 4, 8), (4, 8, 2), (8, 4, 2)]
    sig = inspect.signature(TransferTopology)
    for d_tp, p_tp, heads in cases:
        for d_rank in range(d_tp):
            kwargs = {}
            for name in sig.parameters:
                if name == "tp_rank": kwargs[name] = d_rank
                elif name == "tp_size": kwargs[name] = d_tp
                elif name == "total_num_kv_heads": kwargs[name] = heads
                elif name == "is_mla": kwargs[name] = False
                elif name == "block_size": kwargs[name] = 16
                elif name == "engine_id": kwargs[name] = "d"
                elif name == "remote_tp_size": kwargs[name] = {}
                elif name == "remote_block_size": kwargs[name] = {}
                elif name == "is_mamba": kwargs[name] = False
                elif name == "attn_backends": kwargs[name] = []
                elif name == "tensor_shape": kwargs[name] = None
            try:
                topo = TransferTopology(**kwargs)
            except Exception as e:  # constructor differs: build a duck-typed stand-in
                class _T: pass
                topo = _T(); topo.tp_rank = d_rank; topo.tp_size = d_tp; topo.total_num_kv_heads = heads; topo.is_mla = False
            m = compute_tp_mapping(topo, p_tp, (FullAttentionSpec,))
            rows.append({
                "d_tp": d_tp, "p_tp": p_tp, "d_rank": d_rank, "total_kv_heads": heads,
                "source_ranks": list(m.all_source_ranks),
                "rank_to_slot": {str(k): int(v) for k, v in m.rank_to_attention_slot.items()},
                "rank_offset_factor": int(m.rank_offset_factor),
            })
    with open(os.path.join(HERE, "tp_mapping.json"), "w") as f:
        json.dump({"source": "vllm 0.22.0 compute_tp_mapping", "rows": rows}, f, indent=1)
    print("tp_mapping.json written,", len(rows), "rows")


if __name__ == "__main__":
    layout_fixture()
    blocksize_fixture()
    tp_fixture()
