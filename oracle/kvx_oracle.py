"""CPU oracle for the prefill→decode paged-KV handoff.  TEST INFRASTRUCTURE ONLY.

PARITY UNPINNED BY THE REFERENCE: emolinaro/dynamo-k8s-llm-inference holds no code, tests or golden
vectors for this path.  The algorithm restated here is the one
`BASELINE.json.north_star` names; the byte-level meaning of its words follows the transfer code the
reference's manifests select at run time (examples/deploy/vllm/disagg.yaml:28,37,57 →
`nvcr.io/nvidia/ai-dynamo/vllm-runtime:0.8.1` → vLLM's NIXL connector).  That code is a third-party
dependency absent from /root/reference; the copy cited below as "[vllm]" is vLLM 0.22.0 as installed
on this box (vllm/distributed/kv_transfer/kv_connector/...), used as the layout authority, and the
fixtures under tests/golden/ were generated from its helpers by tests/golden/make_golden.py.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import
this module; the product (dynamo-k8s-llm-inference_b200/) never does.

Model
-----
A pool is raw bytes: `buf` (uint8, flat) + one byte offset per region + a block stride, exactly the
record [vllm] v1/nixl/worker.py:1099-1182 registers (base address per region, block_len, num_blocks).
A chunk = one block of one region = N tokens x H heads x D*E bytes, rows ordered
  NHD: [N][H][D]   ([vllm] v1/attention/backends/flash_attn.py:140-168, stride order (0,1,2,3,4))
  HND: [H][N][D]   (same file, stride order (0,1,3,2,4))
All functions move opaque bytes; equality is on integers, never on floats.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import List, Optional, Sequence, Tuple

import numpy as np

NHD = 0
HND = 1
CANARY16 = 0xC0DE


@dataclass
class Xform:
    """Head slice moved by one edge (TP reshard) and the staging image's row order."""

    src_head_off: int = 0
    dst_head_off: int = 0
    heads: int = 0  # 0 = min(src heads, dst heads)
    stage_layout: Optional[int] = None  # None = the pool's own layout


@dataclass
class Pool:
    """Host model of one rank's paged KV pool (mirrors kvx_pool_desc in include/kvx.h)."""

    buf: np.ndarray  # uint8, flat
    region_off: List[int]  # byte offset of every region's block 0 inside buf
    num_blocks: int
    block_stride: int  # bytes between block ids inside a region
    block_tokens: int
    heads: int
    head_dim: int
    elem_bytes: int
    layout: int
    tp_rank: int = 0
    tp_size: int = 1

    @property
    def regions(self) -> int:
        return len(self.region_off)

    @property
    def row_bytes(self) -> int:
        return self.head_dim * self.elem_bytes

    @property
    def chunk_bytes(self) -> int:
        return self.block_tokens * self.heads * self.row_bytes

    def region_view(self, r: int) -> np.ndarray:
        """[num_blocks, chunk_bytes] strided view of region r (writes go through to buf)."""
        start = self.region_off[r]
        return np.lib.stride_tricks.as_strided(
            self.buf[start:], shape=(self.num_blocks, self.chunk_bytes), strides=(self.block_stride, 1), writeable=True
        )


def make_pool(
    layers: int,
    heads: int,
    head_dim: int,
    num_blocks: int,
    layout: int,
    *,
    block_tokens: int = 16,
    elem_bytes: int = 2,
    arrangement: str = "split",
    fill: str = "random",
    seed: int = 0,
    tp_rank: int = 0,
    tp_size: int = 1,
) -> Pool:
    """Allocate a host pool.

    arrangement "split": every layer's K and V are separate regions, each [num_blocks][chunk]
      (FlashAttention (2, nb, N, H, D): [vllm] flash_attn.py:140-168);
    arrangement "blocks_first": one allocation per layer [num_blocks][2][chunk], K and V become two
      virtual regions with stride 2*chunk ([vllm] v1/nixl/worker.py:946-955,1121-1132).
    fill "random": uniform random 16-bit patterns (NaN/Inf/denormal bf16 included on purpose,
      SURVEY.md §8d); fill "canary": 0xC0DE words.
    """
    chunk = block_tokens * heads * head_dim * elem_bytes
    regions = 2 * layers
    total = regions * num_blocks * chunk
    if fill == "random":
        rng = np.random.Generator(np.random.PCG64(seed))
        words = rng.integers(0, 65536, size=total // 2, dtype=np.uint16)
        buf = words.view(np.uint8).copy()
    elif fill == "canary":
        buf = np.full(total // 2, CANARY16, dtype=np.uint16).view(np.uint8).copy()
    else:
        raise ValueError(fill)
    if arrangement == "split":
        region_off = [r * num_blocks * chunk for r in range(regions)]
        stride = chunk
    elif arrangement == "blocks_first":
        region_off = []
        for l in range(layers):
            base = l * num_blocks * 2 * chunk
            region_off += [base, base + chunk]
        stride = 2 * chunk
    else:
        raise ValueError(arrangement)
    return Pool(buf, region_off, num_blocks, stride, block_tokens, heads, head_dim, elem_bytes, layout, tp_rank, tp_size)


def make_table(num_blocks: int, n: int, seed: int, contiguous: bool = False) -> np.ndarray:
    """Block table per SURVEY.md §8d: unique, scattered, never block 0 (vLLM's null block)."""
    if n > num_blocks - 1:
        raise ValueError("table longer than the pool")
    if contiguous:
        return np.arange(1, n + 1, dtype=np.int32)
    rng = np.random.Generator(np.random.PCG64(seed))
    return (1 + rng.permutation(num_blocks - 1)[:n]).astype(np.int32)


# ------------------------------------------------------------------------------------------------
# row-level helpers
# ------------------------------------------------------------------------------------------------
def _rows(chunks: np.ndarray, layout: int, N: int, H: int, R: int) -> np.ndarray:
    """[n, chunk_bytes] → logical [n, N(token), H(head), R] view/copy."""
    n = chunks.shape[0]
    if layout == NHD:
        return chunks.reshape(n, N, H, R)
    return chunks.reshape(n, H, N, R).transpose(0, 2, 1, 3)


def _pack(rows: np.ndarray, layout: int) -> np.ndarray:
    """logical [n, N, H, R] → [n, chunk_bytes] in `layout`."""
    n = rows.shape[0]
    if layout == NHD:
        return np.ascontiguousarray(rows).reshape(n, -1)
    return np.ascontiguousarray(rows.transpose(0, 2, 1, 3)).reshape(n, -1)


def _resolve(xf: Optional[Xform], src_heads: Optional[int], dst_heads: Optional[int], dflt_layout: int):
    xf = xf or Xform()
    so, do, hc = xf.src_head_off, xf.dst_head_off, xf.heads
    if hc == 0:
        cands = []
        if src_heads is not None:
            cands.append(src_heads - so)
        if dst_heads is not None:
            cands.append(dst_heads - do)
        hc = min(cands)
    if hc <= 0:
        raise ValueError("no heads to move")
    if src_heads is not None and so + hc > src_heads:
        raise ValueError("source head slice out of range")
    if dst_heads is not None and do + hc > dst_heads:
        raise ValueError("destination head slice out of range")
    sl = dflt_layout if xf.stage_layout is None else xf.stage_layout
    return so, do, hc, sl


def touched(pool: Pool, table: Sequence[int], head_off: int = 0, heads: Optional[int] = None) -> List[tuple]:
    """The byte ranges (offset into pool.buf, length) one side of a transfer touches, region-major then in table
    order: descriptor (region r, block id b) = region base + b * block stride ([vllm] v1/nixl/worker.py:113-115 for the
    id, :1108-1133 for the address, blocks-first K/V as two virtual regions), and for a head slice of an HND chunk the
    `heads` heads from `head_off` on, which are contiguous: + head_off * N * row, heads * N * row long ([vllm]
    worker.py:1159-1168 on the remote side, :152-192 on the local side)."""
    if pool.layout != HND and (head_off != 0 or (heads is not None and heads != pool.heads)):
        raise ValueError("a head slice is one contiguous range only in HND chunks")
    heads = pool.heads if heads is None else heads
    head_bytes = pool.block_tokens * pool.row_bytes
    out = []
    for r in range(pool.regions):
        for b in table:
            out.append((pool.region_off[r] + int(b) * pool.block_stride + head_off * head_bytes, heads * head_bytes))
    return out


def _check_table(table: np.ndarray, num_blocks: int) -> np.ndarray:
    t = np.asarray(table, dtype=np.int64)
    if t.size and (t.min() < 0 or t.max() >= num_blocks):
        raise IndexError("block id out of range")
    return t


# ------------------------------------------------------------------------------------------------
# H1+H2(+H3): gather with in-flight layout transform / head slice
# ------------------------------------------------------------------------------------------------
def gather(src: Pool, table: Sequence[int], xf: Optional[Xform] = None) -> np.ndarray:
    """Staging image uint8 [regions, n, N*heads*R] in xf.stage_layout, heads compacted.

    Block lookup: address = region base + block_id * block_len ([vllm] v1/nixl/worker.py:1108-1119,
    descriptor id = region*num_blocks + block_id, worker.py:113-115).  Layout permute: the
    NHD↔HND exchange of the two middle axes that [vllm] kv_connector/utils.py:255-277 performs on
    the receive side, done here on the way out.  Head slice: worker.py:152-192 / 1159-1168.
    """
    so, _, hc, sl = _resolve(xf, src.heads, None, src.layout)
    t = _check_table(table, src.num_blocks)
    N, R = src.block_tokens, src.row_bytes
    out = np.empty((src.regions, len(t), N * hc * R), dtype=np.uint8)
    for r in range(src.regions):
        chunks = src.region_view(r)[t]  # fancy index → copy [n, chunk]
        rows = _rows(chunks, src.layout, N, src.heads, R)[:, :, so : so + hc, :]
        out[r] = _pack(rows, sl)
    return out


# ------------------------------------------------------------------------------------------------
# H4: peer-to-peer copy (bytes are bytes)
# ------------------------------------------------------------------------------------------------
def copy(staging: np.ndarray) -> np.ndarray:
    """The transport moves the image unchanged ([vllm] v1/nixl/worker.py:2220-2233)."""
    return staging.copy()


# ------------------------------------------------------------------------------------------------
# H5(+H3): scatter into freshly allocated slots
# ------------------------------------------------------------------------------------------------
def scatter(staging: np.ndarray, dst: Pool, table: Sequence[int], xf: Optional[Xform] = None) -> None:
    """Write staging rows into dst's blocks `table`; touches only those blocks (and, with a head
    slice, only those heads).  Local descriptor side of the READ: [vllm] worker.py:2208-2213; slot
    offset for P_tp > D_tp: worker.py:1159-1168 with rank_to_attention_slot (tp_mapping.py:119-127).
    """
    _, do, hc, sl = _resolve(xf, None, dst.heads, dst.layout)
    t = _check_table(table, dst.num_blocks)
    if len(np.unique(t)) != len(t):
        raise ValueError("destination table must not repeat a block")
    N, R = dst.block_tokens, dst.row_bytes
    if staging.shape != (dst.regions, len(t), N * hc * R):
        raise ValueError(f"staging shape {staging.shape} does not match")
    for r in range(dst.regions):
        view = dst.region_view(r)
        rows_in = _rows(staging[r], sl, N, hc, R)
        cur = view[t]  # copy of the destination chunks
        rows_out = _rows(cur, dst.layout, N, dst.heads, R).copy()
        rows_out[:, :, do : do + hc, :] = rows_in
        view[t] = _pack(rows_out, dst.layout)


def handoff(src: Pool, src_table: Sequence[int], dst: Pool, dst_table: Sequence[int], xf: Optional[Xform] = None) -> None:
    """H1-H5 composed: what one P rank → D rank edge of a request does."""
    if src.block_tokens != dst.block_tokens:
        return handoff_regroup(src, src_table, dst, dst_table, xf)
    so, do, hc, _ = _resolve(xf, src.heads, dst.heads, src.layout)
    st_src, st_dst = list(src_table), list(dst_table)
    if len(st_dst) < len(st_src):
        # partial prefix hit: D already holds the head of the prompt, take P's LAST len(D) blocks
        # ([vllm] v1/nixl/worker.py:2307-2316)
        st_src = st_src[len(st_src) - len(st_dst) :]
    if len(st_dst) != len(st_src):
        raise ValueError("destination table longer than source table")
    if len(st_dst) == 0:
        return  # full prefix hit: notify only ([vllm] worker.py:2166-2185)
    image = gather(src, st_src, Xform(so, 0, hc, dst.layout))
    image = copy(image)
    scatter(image, dst, st_dst, Xform(0, do, hc, dst.layout))


def handoff_regroup(src: Pool, src_table: Sequence[int], dst: Pool, dst_table: Sequence[int],
                    xf: Optional[Xform] = None, n_small: Optional[int] = None) -> None:
    """Handoff between pools whose block sizes differ by an integer ratio k (P 16-token blocks into
    D 32/64-token blocks or the reverse).  The granule is the smaller block; granule i lives in big
    block table[i // k] at tokens [(i % k) * N, +N).  This is the regrouping [vllm]
    kv_connector/utils.py:224-252 (same layout) and :280-299 (HND → NHD) perform after landing k
    remote blocks back to back in one local block.  n_small defaults to k * len(big table), capped by
    the small side's table."""
    so, do, hc, _ = _resolve(xf, src.heads, dst.heads, src.layout)
    Ns, Nd = src.block_tokens, dst.block_tokens
    N = min(Ns, Nd)
    if max(Ns, Nd) % N:
        raise ValueError("block sizes must divide")
    ks, kd = Ns // N, Nd // N
    ts, td = _check_table(src_table, src.num_blocks), _check_table(dst_table, dst.num_blocks)
    if n_small is None:
        n_small = min(len(ts) * ks, len(td) * kd)
    if n_small > len(ts) * ks or n_small > len(td) * kd:
        raise ValueError("tables too short for n_small granules")
    R = src.row_bytes
    for r in range(src.regions):
        sview, dview = src.region_view(r), dst.region_view(r)
        for i in range(n_small):
            sb, ss = int(ts[i // ks]), i % ks
            db, dsub = int(td[i // kd]), i % kd
            srows = _rows(sview[sb : sb + 1], src.layout, Ns, src.heads, R)[0, ss * N : (ss + 1) * N, so : so + hc, :]
            cur = dview[db : db + 1].copy()
            drows = _rows(cur, dst.layout, Nd, dst.heads, R).copy()
            drows[0, dsub * N : (dsub + 1) * N, do : do + hc, :] = srows
            dview[db] = _pack(drows, dst.layout)[0]


# ------------------------------------------------------------------------------------------------
# H3: TP mapping (restates [vllm] v1/nixl/tp_mapping.py:65-142 for attention KV, non-MLA)
# ------------------------------------------------------------------------------------------------
@dataclass
class Edge:
    p_rank: int
    d_rank: int
    xform: Xform


@dataclass
class TPMap:
    source_ranks: Tuple[int, ...]
    rank_to_slot: dict
    rank_offset_factor: int
    tp_ratio: int
    edges: List[Edge] = field(default_factory=list)


def tp_mapping(d_tp: int, p_tp: int, d_rank: int, total_kv_heads: int) -> TPMap:
    """Which P ranks D rank `d_rank` pulls from, where each lands, and which head slice it reads."""
    if d_tp >= p_tp:
        if d_tp % p_tp:
            raise ValueError("tp sizes must divide")
        ranks = [d_rank * p_tp // d_tp]
        tp_ratio = d_tp // p_tp
    else:
        if p_tp % d_tp:
            raise ValueError("tp sizes must divide")
        k = p_tp // d_tp
        start = d_rank * k
        heads = [(start + j) * total_kv_heads // p_tp for j in range(k)]
        ranks, seen = [], set()
        for j, h in enumerate(heads):  # GQA dedup: first rank per distinct head
            if h not in seen:
                seen.add(h)
                ranks.append(start + j)
        tp_ratio = -k
    head_to_slot = {}
    for i, r in enumerate(ranks):
        head_to_slot[r * total_kv_heads // p_tp] = i
    rank_to_slot = {r: head_to_slot.get(r * total_kv_heads // p_tp, 0) for r in sorted(set(ranks))}
    if d_tp <= p_tp:
        rof = 0
    elif d_tp > total_kv_heads:
        local_head = d_rank * total_kv_heads // d_tp
        p_start = ranks[0] * total_kv_heads // p_tp
        rof = local_head - p_start
    else:
        rof = d_rank % (d_tp // p_tp)
    hp = max(1, total_kv_heads // p_tp)  # heads per P rank
    hd = max(1, total_kv_heads // d_tp)  # heads per D rank
    edges = []
    for r in ranks:
        if d_tp > p_tp:
            edges.append(Edge(r, d_rank, Xform(src_head_off=rof * hd if d_tp <= total_kv_heads else rof, dst_head_off=0, heads=hd)))
        else:
            edges.append(Edge(r, d_rank, Xform(src_head_off=0, dst_head_off=rank_to_slot[r] * hp, heads=min(hp, hd))))
    return TPMap(tuple(ranks), rank_to_slot, rof, tp_ratio, edges)
