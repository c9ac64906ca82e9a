"""Generates tests/golden/*.npz|json from the on-box vLLM 0.22 helpers (run in the authoring container;
the fixtures are committed, this script is how they were made).

vLLM is NOT the reference (the reference holds no code for this path); it is the code
the reference's manifests select at run time (examples/deploy/vllm/disagg.yaml:37,57) and the only
on-box authority for what "NHD/HND", "block id" and "heterogeneous-TP slot" mean in bytes.

  layout_hnd_to_nhd.npz   raw HND blocks landed in an NHD-shaped cache, then
                          kv_postprocess_layout_on_receive (kv_connector/utils.py:255-277)
  tp_mapping.json         compute_tp_mapping (v1/nixl/tp_mapping.py:65-142) known answers
  addressing.json         what the connector's own descriptor code says a request touches, byte for byte: the methods of
                          NixlConnectorWorker (v1/nixl/worker.py) called unbound on a stand-in `self` that carries only
                          the attributes they read — _compute_desc_ids (:89-151), _build_fa_local incl. the blocks-first
                          K/V virtual-region split (:1099-1133), _build_fa_remote (heterogeneous-TP byte offsets,
                          :1135-1182), _build_local_splits_from_plan (:152-192), _apply_prefix_caching (tail alignment)
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


def addressing_fixture():
    import types

    from vllm.distributed.kv_transfer.kv_connector.v1.nixl import worker as W
    from vllm.distributed.kv_transfer.kv_connector.v1.nixl.tp_mapping import compute_tp_mapping
    from vllm.v1.kv_cache_interface import FullAttentionSpec

    C = W.NixlConnectorWorker

    def stand_in(base_count, block_len, num_blocks, blocks_first, regions):
        ns = types.SimpleNamespace(num_regions=regions, block_len_per_layer=[block_len] * base_count, _has_mamba=False,
                                   num_blocks=num_blocks, device_id=0, _group_spec_types=(FullAttentionSpec,),
                                   transfer_topo=types.SimpleNamespace(is_kv_layout_blocks_first=blocks_first))
        ns.get_backend_aware_kv_block_len = types.MethodType(C.get_backend_aware_kv_block_len, ns)
        return ns

    rng = np.random.Generator(np.random.PCG64(20261108))
    out = {"source": "vllm 0.22.0 NixlConnectorWorker methods (unbound, stand-in self)", "local": [], "hetero_tp": [], "prefix": []}
    BASE = 0x7F0000000000
    L, N, H, D, E = 3, 16, 8, 64, 2
    chunk = N * H * D * E
    # --- local descriptors + descriptor ids of a request, split K/V pools and blocks-first joint pools
    for arrangement in ("split", "blocks_first"):
        nb = 23
        if arrangement == "split":   # (2, nb, N, H, D) per layer: K and V are regions of their own
            bases = [BASE + r * nb * chunk for r in range(2 * L)]
            ns = stand_in(2 * L, chunk, nb, False, 2 * L)
        else:                        # (nb, 2, N, H, D) per layer: one registered region per layer, two virtual ones
            bases = [BASE + l * nb * 2 * chunk for l in range(L)]
            ns = stand_in(L, 2 * chunk, nb, True, 2 * L)
        descs = C._build_fa_local(ns, bases, 1)
        ids = (1 + rng.permutation(nb - 1)[:9]).astype(np.int64)
        sel = C._compute_desc_ids(ns, [ids.tolist()], nb, None, 1)
        out["local"].append({"arrangement": arrangement, "layers": L, "block_tokens": N, "heads": H, "head_dim": D, "elem_bytes": E,
                             "num_blocks": nb, "base": BASE, "block_ids": ids.tolist(), "desc_ids": [int(x) for x in sel],
                             "touched": [[int(descs[i][0]) - BASE, int(descs[i][1])] for i in sel]})
    # --- heterogeneous TP: byte ranges a decode rank reads from each prefill rank's block, and where they land
    for d_tp, p_tp in ((2, 4), (4, 2), (1, 4), (4, 1), (2, 1), (1, 2)):
        for blocks_first in (False, True):
            hp, hd = H // p_tp, H // d_tp
            chunk_p, chunk_d = N * hp * D * E, N * hd * D * E
            nb_p, nb_d = 7, 5
            for d_rank in range(d_tp):
                topo = types.SimpleNamespace(tp_rank=d_rank, tp_size=d_tp, total_num_kv_heads=H, is_mla=False)
                plan = compute_tp_mapping(topo, p_tp, (FullAttentionSpec,))
                k = len(plan.source_ranks_per_group[0])
                if blocks_first:
                    me = stand_in(L, 2 * chunk_d, nb_d, True, 2 * L)
                    l_bases = [BASE + l * nb_d * 2 * chunk_d for l in range(L)]
                    r_bases = [BASE + l * nb_p * 2 * chunk_p for l in range(L)]
                    meta = types.SimpleNamespace(kv_caches_base_addr=r_bases, num_blocks=nb_p, block_lens=[2 * chunk_p] * L, device_id=1)
                else:
                    me = stand_in(2 * L, chunk_d, nb_d, False, 2 * L)
                    l_bases = [BASE + r * nb_d * chunk_d for r in range(2 * L)]
                    r_bases = [BASE + r * nb_p * chunk_p for r in range(2 * L)]
                    meta = types.SimpleNamespace(kv_caches_base_addr=r_bases, num_blocks=nb_p, block_lens=[chunk_p] * (2 * L), device_id=1)
                local = C._build_fa_local(me, l_bases, 1)
                remote = C._build_fa_remote(me, plan, meta, 1)
                handles = list(C._build_local_splits_from_plan(me, plan, local, len(local))) if k > 1 else [local]
                row = {"d_tp": d_tp, "p_tp": p_tp, "d_rank": d_rank, "blocks_first": blocks_first, "layers": L, "block_tokens": N,
                       "total_kv_heads": H, "head_dim": D, "elem_bytes": E, "nb_p": nb_p, "nb_d": nb_d, "base": BASE,
                       "source_ranks": list(plan.all_source_ranks),
                       # per (region, block): [offset in the prefill rank's pool, length]
                       "remote": [[int(a) - BASE, int(n_)] for a, n_, _ in remote],
                       # per source rank, per (region, block): [offset in this decode rank's pool, length]
                       "local_per_source": [[[int(a) - BASE, int(n_)] for a, n_, _ in h] for h in handles]}
                out["hetero_tp"].append(row)
    # --- prefix-cache hits on the decode side: which remote blocks are read
    # (an empty local list never reaches this helper: a full prefix hit only notifies, worker.py:2166-2185)
    for n_remote, n_local in ((9, 9), (9, 4), (9, 1), (17, 16), (1, 1)):
        ns = types.SimpleNamespace(_has_mamba=False)
        remote_ids = (1 + rng.permutation(40)[:n_remote]).tolist()
        local_ids = (1 + rng.permutation(40)[:n_local]).tolist()
        l_out, r_out = C._apply_prefix_caching(ns, [list(local_ids)], [list(remote_ids)], 1)
        out["prefix"].append({"remote": remote_ids, "local": local_ids, "remote_read": list(r_out[0]), "local_written": list(l_out[0])})
    with open(os.path.join(HERE, "addressing.json"), "w") as f:
        json.dump(out, f)
    print("addressing.json written:", len(out["local"]), "local,", len(out["hetero_tp"]), "hetero-TP,", len(out["prefix"]), "prefix rows")


if __name__ == "__main__":
    addressing_fixture()
    if "--only-addressing" in sys.argv:
        sys.exit(0)
    layout_fixture()
    blocksize_fixture()
    tp_fixture()
