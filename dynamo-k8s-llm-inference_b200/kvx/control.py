"""Control plane of the handoff as a wire format plus the prefill-side block lease.

What travels between workers besides KV bytes:
  * once per worker pair, a handshake record — the counterpart of NixlAgentMetadata
    (vllm v1/nixl/metadata.py:41-134): pool geometry, peer-mapping blobs, and a compatibility hash
    that both sides must agree on before any block moves;
  * per request, the prefill side's block ids and identity — the counterpart of kv_transfer_params
    (vllm v1/nixl/scheduler.py:664-674);
  * completion: a release per (request, consumer) — here a flag slot written by the consumer's kernel —
    and heartbeats that keep a lease alive while a request waits in the decode queue.
The prefill worker must not free a finished request's blocks until every consumer has read them, and
must not keep them forever if a consumer dies: `LeaseTable` restates that rule
(vllm v1/nixl/worker.py:1771-1788 expiry, 1792-1845 consumer counting, 1847-1866 heartbeats).
"""
from __future__ import annotations

import hashlib
import struct
from collections import OrderedDict
from dataclasses import dataclass
from typing import Callable, Dict, Iterable, List, Optional, Set

import msgpack
import numpy as np

from .workers import TransferParams, WorkerMetadata

WIRE_VERSION = 1
MAGIC = b"KVXC"


class WireError(ValueError):
    pass


def compat_hash(*, regions: int, head_dim: int, elem_bytes: int, total_kv_heads: int, abi: int) -> str:
    """Both ends must produce the same digest or the handshake is refused (the role vLLM's
    compatibility hash plays, metadata.py:60-134): anything that changes the meaning of a byte."""
    h = hashlib.sha256()
    h.update(struct.pack("<6I", WIRE_VERSION, abi >> 16, regions, head_dim, elem_bytes, total_kv_heads))
    return h.hexdigest()[:16]


def _frame(kind: int, body: dict) -> bytes:
    payload = msgpack.packb(body, use_bin_type=True)
    return MAGIC + struct.pack("<HHI", WIRE_VERSION, kind, len(payload)) + payload


def _unframe(buf: bytes, kind: int) -> dict:
    if len(buf) < 12 or buf[:4] != MAGIC:
        raise WireError("not a kvx control record")
    ver, k, n = struct.unpack_from("<HHI", buf, 4)
    if ver != WIRE_VERSION:
        raise WireError(f"wire version {ver}, expected {WIRE_VERSION}")
    if k != kind:
        raise WireError(f"record kind {k}, expected {kind}")
    if len(buf) != 12 + n:
        raise WireError("truncated record")
    return msgpack.unpackb(buf[12:], raw=False)


def encode_metadata(meta: WorkerMetadata, *, regions: int, head_dim: int, elem_bytes: int, total_kv_heads: int,
                    abi: int = 0x00010000) -> bytes:
    body = dict(vars(meta))
    body["compat"] = compat_hash(regions=regions, head_dim=head_dim, elem_bytes=elem_bytes,
                                 total_kv_heads=total_kv_heads, abi=abi)
    return _frame(1, body)


def decode_metadata(buf: bytes, *, expect_compat: Optional[str] = None) -> WorkerMetadata:
    body = _unframe(buf, 1)
    compat = body.pop("compat")
    if expect_compat is not None and compat != expect_compat:
        raise WireError(f"incompatible peer: compat hash {compat} != {expect_compat}")
    return WorkerMetadata(**body)


def encode_params(p: TransferParams) -> bytes:
    ids = np.ascontiguousarray(p.remote_block_ids, dtype=np.int32)
    return _frame(2, {"request_id": p.request_id, "remote_rank": p.remote_rank, "tp_size": p.tp_size,
                      "ids": ids.tobytes()})


def decode_params(buf: bytes) -> TransferParams:
    b = _unframe(buf, 2)
    return TransferParams(b["request_id"], np.frombuffer(b["ids"], dtype=np.int32).copy(), b["remote_rank"], b["tp_size"])


def encode_heartbeat(request_ids: Iterable[str]) -> bytes:
    return _frame(3, {"reqs": list(request_ids)})


def decode_heartbeat(buf: bytes) -> List[str]:
    return list(_unframe(buf, 3)["reqs"])


@dataclass
class _Lease:
    expires: float
    consumers: int
    released: int = 0


class LeaseTable:
    """Prefill-side bookkeeping of finished requests whose blocks a decode worker has yet to read.

    grant()     after prefill: blocks stay allocated until `consumers` releases arrive or the lease ends
    release()   one consumer finished reading (flag observed); returns True when the blocks may be freed
    heartbeat() the decode side still wants these requests: push their expiry out
    expire()    requests whose lease ran out — their blocks are reclaimed and late readers are refused
    """

    def __init__(self, lease_seconds: float = 30.0, extension_seconds: float = 30.0,
                 clock: Callable[[], float] = None):
        import time

        self._clock = clock or time.perf_counter
        self._lease, self._ext = lease_seconds, extension_seconds
        self._leases: "OrderedDict[str, _Lease]" = OrderedDict()
        self.expired_total = 0

    def __contains__(self, request_id: str) -> bool:
        return request_id in self._leases

    def __len__(self) -> int:
        return len(self._leases)

    def grant(self, request_id: str, consumers: int = 1) -> None:
        if request_id in self._leases:
            raise KeyError(f"request {request_id} already leased")
        if consumers < 1:
            raise ValueError("a lease needs at least one consumer")
        self._leases[request_id] = _Lease(self._clock() + self._lease, consumers)

    def release(self, request_id: str) -> bool:
        lease = self._leases.get(request_id)
        if lease is None:
            # read after expiry (or never granted): the reader may hold invalid blocks
            return False
        lease.released += 1
        if lease.released >= lease.consumers:
            del self._leases[request_id]
            return True
        return False

    def heartbeat(self, request_ids: Iterable[str]) -> None:
        new_expiry = self._clock() + self._ext
        for rid in request_ids:
            lease = self._leases.get(rid)
            if lease is not None:
                lease.expires = max(lease.expires, new_expiry)

    def expire(self) -> List[str]:
        now = self._clock()
        gone = [rid for rid, l in self._leases.items() if now >= l.expires]
        for rid in gone:
            del self._leases[rid]
        self.expired_total += len(gone)
        return gone


def consumers_per_producer(p_tp: int, d_tp: int) -> int:
    """How many decode ranks read from one prefill rank (vllm worker.py:1825-1844)."""
    return d_tp // p_tp if d_tp > p_tp else 1


@dataclass
class TransferOutcome:
    request_id: str
    ok: bool
    invalid_block_ids: np.ndarray  # local blocks whose contents cannot be trusted (recompute them)
    detail: str = ""


def guarded_pull(worker, key: int, params: TransferParams, local_ids, **kw) -> TransferOutcome:
    """Run one pull and turn any failure into "these local blocks are invalid" instead of an exception:
    the scheduler recomputes them (vllm worker.py:1915-1931 marks failed transfers' blocks invalid)."""
    from .api import KvxError

    local = np.asarray(local_ids, dtype=np.int32)
    try:
        worker.pull(key, params.remote_block_ids, local, **kw)
        return TransferOutcome(params.request_id, True, np.empty(0, np.int32))
    except (KvxError, KeyError, ValueError) as e:
        return TransferOutcome(params.request_id, False, local.copy(), str(e))
