"""Control-plane wire format and the prefill-side lease (host logic, CPU)."""
import numpy as np
import pytest

from kvx import control as c
from kvx.workers import TransferParams, WorkerMetadata


def _meta():
    return WorkerMetadata("decode", 1, 2, 0, 4097, 16, 4, b"\x00pool\xff" * 9, b"flags" * 3, 4096)


def test_metadata_round_trip_and_compat():
    kw = dict(regions=160, head_dim=128, elem_bytes=2, total_kv_heads=8)
    blob = c.encode_metadata(_meta(), **kw)
    good = c.compat_hash(abi=0x00010000, **kw)
    assert c.decode_metadata(blob, expect_compat=good) == _meta()
    with pytest.raises(c.WireError, match="incompatible"):
        c.decode_metadata(blob, expect_compat=c.compat_hash(abi=0x00010000, **{**kw, "head_dim": 64}))
    # an ABI minor bump keeps the hash, a major bump changes it
    assert c.compat_hash(abi=0x00010007, **kw) == good != c.compat_hash(abi=0x00020000, **kw)


def test_framing_errors():
    blob = c.encode_params(TransferParams("r1", np.array([5, 9, 2], np.int32), 3, 4))
    p = c.decode_params(blob)
    assert p.request_id == "r1" and p.remote_block_ids.tolist() == [5, 9, 2] and (p.remote_rank, p.tp_size) == (3, 4)
    with pytest.raises(c.WireError):
        c.decode_params(blob[:-1])
    with pytest.raises(c.WireError):
        c.decode_params(b"XXXX" + blob[4:])
    with pytest.raises(c.WireError):
        c.decode_metadata(blob)  # wrong record kind
    assert c.decode_heartbeat(c.encode_heartbeat(["a", "b"])) == ["a", "b"]


def test_lease_release_counts_consumers():
    now = [0.0]
    t = c.LeaseTable(lease_seconds=10, extension_seconds=10, clock=lambda: now[0])
    t.grant("r", consumers=c.consumers_per_producer(2, 4))  # TP2 prefill read by TP4 decode: 2 readers each
    assert not t.release("r") and "r" in t
    assert t.release("r") and "r" not in t
    assert not t.release("r")  # late / duplicate release is refused
    with pytest.raises(ValueError):
        t.grant("z", consumers=0)
    t.grant("dup")
    with pytest.raises(KeyError):
        t.grant("dup")


def test_lease_expiry_and_heartbeat():
    now = [0.0]
    t = c.LeaseTable(lease_seconds=10, extension_seconds=10, clock=lambda: now[0])
    t.grant("a"); t.grant("b")
    now[0] = 8.0
    t.heartbeat(c.decode_heartbeat(c.encode_heartbeat(["b", "unknown"])))  # b now lives until 18
    now[0] = 11.0
    assert t.expire() == ["a"] and "b" in t and t.expired_total == 1
    assert not t.release("a")  # a reader that arrives after expiry is told its blocks are invalid
    now[0] = 18.5
    assert t.expire() == ["b"] and len(t) == 0


def test_guarded_pull_reports_invalid_blocks():
    class Boom:
        def pull(self, *a, **k):
            raise ValueError("destination table longer than source table")

    class Fine:
        def pull(self, *a, **k):
            return None

    params = TransferParams("r9", np.array([1, 2, 3], np.int32), 0, 1)
    out = c.guarded_pull(Boom(), 0, params, [7, 8, 9, 10])
    assert not out.ok and out.invalid_block_ids.tolist() == [7, 8, 9, 10] and "longer" in out.detail
    out = c.guarded_pull(Fine(), 0, params, [7, 8, 9])
    assert out.ok and out.invalid_block_ids.size == 0


def test_bootstrap_handshake_over_tcp():
    """Two workers shake hands over the bootstrap port (the manifests' --disaggregation-bootstrap-port role):
    records cross intact, a per-request record follows on the same connection, an incompatible peer is refused."""
    from kvx import bootstrap as b

    kw = dict(regions=64, head_dim=128, elem_bytes=2, total_kv_heads=8)
    good = c.compat_hash(abi=0x00010000, **kw)
    p_meta = WorkerMetadata("prefill", 0, 1, 1, 4097, 16, 8, b"P" * 5000, b"pf" * 40, 4096)
    seen, requests = [], []

    def on_peer(rec, addr):
        seen.append(c.decode_metadata(rec, expect_compat=good))   # raises WireError for a stranger

    def on_frame(frame, addr):
        requests.append(c.decode_params(frame))
        return c.encode_heartbeat([requests[-1].request_id])

    srv = b.BootstrapServer(c.encode_metadata(p_meta, **kw), on_peer, on_frame=on_frame)
    try:
        host, port = srv.address
        cli = b.BootstrapClient(host, port, c.encode_metadata(_meta(), **kw))
        assert c.decode_metadata(cli.peer_record, expect_compat=good) == p_meta
        assert seen == [_meta()]
        reply = cli.send(c.encode_params(TransferParams("req-7", np.array([3, 1, 4, 1, 5], np.int32), 0, 1)), expect_reply=True)
        assert c.decode_heartbeat(reply) == ["req-7"]
        assert requests[0].remote_block_ids.tolist() == [3, 1, 4, 1, 5]
        cli.close()
        # a peer built for another head_dim: the server refuses it and sends nothing back
        with pytest.raises(c.WireError, match="handshake"):
            b.BootstrapClient(host, port, c.encode_metadata(_meta(), **{**kw, "head_dim": 64}), timeout=3.0)
        assert len(seen) == 1 and any("incompatible" in e for e in srv.errors)
        # garbage on the port is an error, not a crash
        import socket

        with socket.create_connection((host, port), timeout=3) as s:
            s.sendall(b"GET / HTTP/1.1\r\n\r\n")
        cli2 = b.BootstrapClient(host, port, c.encode_metadata(_meta(), **kw))
        assert len(seen) == 2
        cli2.close()
    finally:
        srv.close()
