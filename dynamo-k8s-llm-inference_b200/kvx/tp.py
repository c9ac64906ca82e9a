"""TP reshard planning: which prefill rank feeds which decode rank, and the head slice each edge moves.

Host logic of H3.  The reference sets TP only through manifest args
(examples/deploy/sglang/disagg.yaml:41-42,76-77 `--tp`); the mapping rule is the one its images apply,
vLLM's compute_tp_mapping (v1/nixl/tp_mapping.py:65-142): with P_tp > D_tp one decode rank pulls from
k = P_tp/D_tp prefill ranks and each lands in head slot s; with D_tp > P_tp a decode rank reads a head
slice of one prefill rank at rank_offset_factor.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List

from .api import Xform


@dataclass(frozen=True)
class TPEdge:
    p_rank: int
    d_rank: int
    src_head_off: int  # in P rank's local heads
    dst_head_off: int  # in D rank's local heads
    heads: int

    def xform(self, stage_layout: int) -> Xform:
        return Xform(self.src_head_off, self.dst_head_off, self.heads, stage_layout)


def heads_per_rank(total_kv_heads: int, tp: int) -> int:
    # tp > kv heads replicates KV (vllm kv_connector/utils.py:525-538): one head per rank
    return max(1, total_kv_heads // tp)


def plan_edges(p_tp: int, d_tp: int, total_kv_heads: int) -> List[TPEdge]:
    """All (P rank → D rank) edges of one request, each with its head slice."""
    if p_tp <= 0 or d_tp <= 0 or max(p_tp, d_tp) % min(p_tp, d_tp):
        raise ValueError(f"TP sizes must divide: P={p_tp} D={d_tp}")
    hp, hd = heads_per_rank(total_kv_heads, p_tp), heads_per_rank(total_kv_heads, d_tp)
    edges: List[TPEdge] = []
    for j in range(d_tp):
        if d_tp >= p_tp:
            ratio = d_tp // p_tp
            p = j * p_tp // d_tp
            if d_tp > total_kv_heads:
                # replicated KV on D: global head of this rank, relative to P rank's first head
                off = j * total_kv_heads // d_tp - p * total_kv_heads // p_tp
            else:
                off = (j % ratio) * hd
            edges.append(TPEdge(p, j, off, 0, hd))
        else:
            k = p_tp // d_tp
            seen = {}
            for q in range(j * k, (j + 1) * k):
                g = q * total_kv_heads // p_tp  # first global head held by P rank q
                if g in seen:
                    continue  # replicated head: read it once (np.unique dedup in the reference images' code)
                slot = len(seen)
                seen[g] = slot
                edges.append(TPEdge(q, j, 0, slot * hp, min(hp, hd)))
    return edges


def sources_of(edges: List[TPEdge], d_rank: int) -> List[TPEdge]:
    return [e for e in edges if e.d_rank == d_rank]
