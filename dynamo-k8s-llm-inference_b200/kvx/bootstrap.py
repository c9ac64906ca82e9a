"""Handshake over a plain TCP socket: the role of the bootstrap port in the reference's manifests
(`--disaggregation-bootstrap-port "12345"`, examples/deploy/sglang/disagg.yaml:49-50,84-85) and of the NIXL
side channel the vLLM images open per engine pair (vllm v1/nixl/worker.py handshake, 1224-1340).

A worker SERVES its own handshake record (kvx.control.encode_metadata) on a port; a peer connects, sends its
record and gets the server's back — one exchange per engine pair, before any block moves.  The same
connection then carries per-request records (kv_transfer_params, heartbeats) as further frames.  Frames are
the self-delimiting control records of kvx.control (magic, version, kind, length), so the stream needs no
extra framing.  Nothing here touches the GPU: the blobs inside the records are what kvx_pool_import maps.
"""
from __future__ import annotations

import socket
import struct
import threading
from typing import Callable, List, Optional, Tuple

from .control import MAGIC, WireError

_HDR = 12  # MAGIC(4) + version u16 + kind u16 + length u32
MAX_FRAME = 64 << 20


def _recv_exact(sock: socket.socket, n: int) -> bytes:
    buf = bytearray()
    while len(buf) < n:
        part = sock.recv(n - len(buf))
        if not part:
            raise WireError("peer closed the connection mid-record")
        buf += part
    return bytes(buf)


def recv_frame(sock: socket.socket) -> bytes:
    """One whole control record from the stream (header validated here, body by kvx.control.decode_*)."""
    hdr = _recv_exact(sock, _HDR)
    if hdr[:4] != MAGIC:
        raise WireError("not a kvx control record")
    (n,) = struct.unpack_from("<I", hdr, 8)
    if n > MAX_FRAME:
        raise WireError(f"record of {n} bytes exceeds the {MAX_FRAME}-byte limit")
    return hdr + _recv_exact(sock, n)


def send_frame(sock: socket.socket, frame: bytes) -> None:
    sock.sendall(frame)


class BootstrapServer:
    """Serves this worker's handshake record; `on_peer(peer_record, address)` is called for every peer that
    shakes hands (typically: decode_metadata + Context.import_pool).  `on_frame(frame, address)` — optional —
    receives every further record a peer sends on the same connection and may return a reply frame."""

    def __init__(self, my_record: bytes, on_peer: Callable[[bytes, Tuple[str, int]], None], *, host: str = "127.0.0.1",
                 port: int = 0, on_frame: Optional[Callable[[bytes, Tuple[str, int]], Optional[bytes]]] = None):
        self._rec, self._on_peer, self._on_frame = my_record, on_peer, on_frame
        self._srv = socket.create_server((host, port))
        self._srv.settimeout(0.2)
        self.address = self._srv.getsockname()
        self._stop = threading.Event()
        self._threads: List[threading.Thread] = []
        self.errors: List[str] = []
        self._accept = threading.Thread(target=self._accept_loop, daemon=True)
        self._accept.start()

    def _accept_loop(self) -> None:
        while not self._stop.is_set():
            try:
                conn, addr = self._srv.accept()
            except socket.timeout:
                continue
            except OSError:
                return
            t = threading.Thread(target=self._serve, args=(conn, addr), daemon=True)
            t.start()
            self._threads.append(t)

    def _serve(self, conn: socket.socket, addr) -> None:
        with conn:
            try:
                conn.settimeout(30.0)
                peer = recv_frame(conn)
                self._on_peer(peer, addr)      # may raise WireError (incompatible peer): nothing is sent back
                send_frame(conn, self._rec)
                while self._on_frame is not None and not self._stop.is_set():
                    try:
                        frame = recv_frame(conn)
                    except (WireError, socket.timeout, OSError):
                        return
                    reply = self._on_frame(frame, addr)
                    if reply is not None:
                        send_frame(conn, reply)
            except Exception as e:  # a bad peer must not take the server down
                self.errors.append(f"{addr}: {e}")

    def close(self) -> None:
        self._stop.set()
        try:
            self._srv.close()
        except OSError:
            pass
        self._accept.join(timeout=2)


class BootstrapClient:
    """Connects to a peer's bootstrap port, exchanges handshake records, then sends per-request records."""

    def __init__(self, host: str, port: int, my_record: bytes, *, timeout: float = 10.0):
        self.sock = socket.create_connection((host, port), timeout=timeout)
        try:
            send_frame(self.sock, my_record)
            self.peer_record = recv_frame(self.sock)
        except (WireError, OSError) as e:
            self.sock.close()
            raise WireError(f"handshake with {host}:{port} failed: {e}") from e

    def send(self, frame: bytes, expect_reply: bool = False) -> Optional[bytes]:
        send_frame(self.sock, frame)
        return recv_frame(self.sock) if expect_reply else None

    def close(self) -> None:
        self.sock.close()
