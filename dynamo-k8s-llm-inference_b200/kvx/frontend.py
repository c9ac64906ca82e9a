"""OpenAI-compatible Frontend stand-in.

The reference's clients talk to the Dynamo Frontend pod over HTTP: `chat.sh:84-99` and
`examples/dgdr/trtllm/multi_convos_parallel.sh:21-48` POST `{model, messages, temperature, max_tokens}` to
`/v1/chat/completions` and read `.choices[0].message.content` (or `.error`); `README.md:133-138` curls
`/v1/models`; `run-benchmarks.sh` hands the same URL to aiperf.  This module answers those calls with a
stand-in that runs, per request, what the disaggregated deployment runs between prefill and first token:
allocate blocks on both workers, pull the prompt's KV from the prefill pool into the decode pool (layout
transform in flight), release the prefill blocks — and exports the metrics the reference's Grafana
dashboard plots (grafana-dynamo-dashboard-configmap.yaml:121,214,307,400,493,504) plus the transfer
metrics the images export (vllm:nixl_*).  No model runs: completions are synthetic text.

  python -m kvx.frontend --model Qwen/Qwen3-0.6B --port 8000 [--manifest disagg.yaml]
  API_URL=http://127.0.0.1:8000/v1/chat/completions MODEL=Qwen/Qwen3-0.6B ./chat.sh
"""
from __future__ import annotations

import argparse
import json
import threading
import time
import uuid
from http.server import BaseHTTPRequestHandler, ThreadingHTTPServer
from typing import Dict, Optional, Tuple

from prometheus_client import CollectorRegistry, Counter, Histogram, generate_latest

from .shapes import KV_SHAPES, KVShape

_BUCKETS = (0.0001, 0.00025, 0.0005, 0.001, 0.0025, 0.005, 0.01, 0.025, 0.05, 0.1, 0.25, 0.5, 1.0, 2.5, 5.0)


def count_tokens(messages) -> int:
    """Prompt length without a tokenizer: one token per 4 characters of all message contents."""
    chars = 0
    for m in messages:
        c = m.get("content", "")
        if isinstance(c, list):  # content parts
            c = " ".join(str(p.get("text", "")) for p in c if isinstance(p, dict))
        chars += len(str(c)) + 4
    return max(1, chars // 4)


class InProcessHandoff:
    """One prefill worker and one decode worker in this process (GPUs `p_device`, `d_device`; they may be
    the same device).  `handoff(isl)` moves one prompt's KV and returns (seconds, bytes, descriptors)."""

    def __init__(self, shape: KVShape, *, p_device: int = 0, d_device: Optional[int] = None, p_layout: int = 1,
                 d_layout: int = 0, max_isl: int = 32768, resident_receiver: bool = False):
        import torch

        from . import synth
        from .workers import DecodeWorker, PrefillWorker

        if not torch.cuda.is_available():
            raise RuntimeError("the kvx frontend needs a CUDA device: there is no CPU fallback")
        if d_device is None:
            d_device = 1 if torch.cuda.device_count() > 1 else p_device
        self.torch, self.shape, self.max_isl = torch, shape, max_isl
        nb = 3 * shape.blocks_for(max_isl) + 1
        self.p = PrefillWorker(p_device, shape, nb, p_layout)
        self.d = DecodeWorker(d_device, shape, nb, d_layout)
        synth.fill_random(self.p.kv, 0xD15A66)
        synth.fill_canary(self.d.kv)
        self.d.add_remote(0, self.p.metadata())
        self.lock = threading.Lock()
        self.seq = 0
        self.d_device = d_device
        # chat-sized prompts are a few blocks: the decode worker then receives through its resident receiver
        # (kvx_queue_*: no kernel launch per request) instead of one launch per request
        self.resident_receiver = resident_receiver

    def handoff(self, isl: int) -> Tuple[float, int, int]:
        torch = self.torch
        n = self.shape.blocks_for(min(isl, self.max_isl))
        with self.lock:
            self.seq += 1
            rid = f"req-{self.seq}"
            src = self.p.blocks.allocate(rid, n)  # what prefill filled
            dst = self.d.blocks.allocate(rid, n)  # fresh decode slots
            try:
                torch.cuda.set_device(self.d_device)
                if self.resident_receiver:
                    t0 = time.perf_counter()
                    ticket = self.d.pull_queued(0, src, dst, notify_slot=self.seq, value=self.seq)
                    self.d.pull_queue(0).wait(ticket, 10000)
                    secs = time.perf_counter() - t0   # submit -> completion word seen (there is no launch to bracket)
                else:
                    stream = torch.cuda.current_stream(self.d_device)
                    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    a.record(stream)
                    self.d.pull(0, src, dst, notify_slot=self.seq, value=self.seq, stream=stream)
                    b.record(stream)
                    b.synchronize()
                    secs = a.elapsed_time(b) / 1e3
                self.d.ctx.check()
                if not self.p.poll(self.seq, self.seq):  # release written by the kernel's last CTA
                    raise RuntimeError("prefill blocks were not released")
            finally:
                self.p.blocks.free(rid)
                self.d.blocks.free(rid)
        chunk = self.shape.chunk_bytes()
        return secs, n * self.shape.regions * chunk, n * self.shape.regions

    def close(self) -> None:
        self.d.close()
        self.p.close()


class Frontend:
    """HTTP surface + metrics; `backend` is anything with handoff(isl) -> (seconds, bytes, descriptors)."""

    def __init__(self, model: str, backend, *, prefill_ms_per_ktoken: float = 0.0, itl_ms: float = 0.0,
                 namespace: str = "default"):
        if model not in KV_SHAPES:
            raise ValueError(f"unknown model '{model}'")
        self.model, self.backend = model, backend
        self.prefill_ms_per_ktoken, self.itl_ms, self.ns = prefill_ms_per_ktoken, itl_ms, namespace
        r = self.registry = CollectorRegistry()
        lab = ["namespace", "model"]
        self.m_req = Counter("dynamo_frontend_requests", "requests", lab + ["status"], registry=r)
        self.m_ttft = Histogram("dynamo_frontend_time_to_first_token_seconds", "TTFT", lab, buckets=_BUCKETS, registry=r)
        self.m_itl = Histogram("dynamo_frontend_inter_token_latency_seconds", "ITL", lab, buckets=_BUCKETS, registry=r)
        self.m_dur = Histogram("dynamo_frontend_request_duration_seconds", "duration", lab, buckets=_BUCKETS, registry=r)
        self.m_isl = Histogram("dynamo_frontend_input_sequence_tokens", "ISL", lab, buckets=(64, 256, 1024, 4096, 16384, 65536), registry=r)
        self.m_osl = Histogram("dynamo_frontend_output_sequence_tokens", "OSL", lab, buckets=(16, 64, 256, 1024, 4096), registry=r)
        self.m_xfer = Histogram("vllm:nixl_xfer_time_seconds", "KV transfer time", lab, buckets=_BUCKETS, registry=r)
        self.m_bytes = Histogram("vllm:nixl_bytes_transferred", "KV bytes per transfer", lab,
                                 buckets=tuple(float(1 << s) for s in range(20, 34, 2)), registry=r)
        self.m_desc = Histogram("vllm:nixl_num_descriptors", "chunks per transfer", lab,
                                buckets=(256, 1024, 4096, 16384, 65536, 262144), registry=r)
        self.m_fail = Counter("vllm:nixl_num_failed_transfers", "failed transfers", lab, registry=r)

    # -- request handling ---------------------------------------------------------------------------
    def models(self) -> Dict:
        return {"object": "list", "data": [{"id": self.model, "object": "model", "owned_by": "kvx"}]}

    def chat(self, body: Dict) -> Tuple[int, Dict]:
        t0 = time.perf_counter()
        lab = (self.ns, self.model)
        model, messages = body.get("model"), body.get("messages")
        if model != self.model:
            self.m_req.labels(*lab, "error").inc()
            return 404, {"error": {"message": f"model '{model}' not found; serving '{self.model}'", "type": "invalid_request_error"}}
        if not isinstance(messages, list) or not messages:
            self.m_req.labels(*lab, "error").inc()
            return 400, {"error": {"message": "'messages' must be a non-empty list", "type": "invalid_request_error"}}
        isl = count_tokens(messages)
        osl = max(1, min(int(body.get("max_tokens") or 16), 64))
        if self.prefill_ms_per_ktoken > 0:
            time.sleep(self.prefill_ms_per_ktoken * isl / 1e6)  # stands in for the prefill forward pass
        try:
            secs, nbytes, ndesc = self.backend.handoff(isl)
        except Exception as e:  # a failed transfer = failed request, reported like the reference clients expect
            self.m_fail.labels(*lab).inc()
            self.m_req.labels(*lab, "error").inc()
            return 500, {"error": {"message": f"KV handoff failed: {e}", "type": "internal_error"}}
        ttft = time.perf_counter() - t0
        if self.itl_ms > 0:
            time.sleep(self.itl_ms * (osl - 1) / 1e3)
        text = (f"kvx synthetic completion: {isl} prompt tokens, KV handoff {nbytes / 2**20:.1f} MiB in "
                f"{secs * 1e3:.3f} ms ({nbytes / secs / 1e9:.1f} GB/s).")
        self.m_ttft.labels(*lab).observe(ttft)
        self.m_itl.labels(*lab).observe(self.itl_ms / 1e3)
        self.m_isl.labels(*lab).observe(isl)
        self.m_osl.labels(*lab).observe(osl)
        self.m_xfer.labels(*lab).observe(secs)
        self.m_bytes.labels(*lab).observe(nbytes)
        self.m_desc.labels(*lab).observe(ndesc)
        self.m_dur.labels(*lab).observe(time.perf_counter() - t0)
        self.m_req.labels(*lab, "success").inc()
        return 200, {
            "id": "chatcmpl-" + uuid.uuid4().hex[:24], "object": "chat.completion", "created": int(time.time()), "model": self.model,
            "choices": [{"index": 0, "message": {"role": "assistant", "content": text}, "finish_reason": "stop"}],
            "usage": {"prompt_tokens": isl, "completion_tokens": osl, "total_tokens": isl + osl},
            "kv_transfer": {"xfer_time_seconds": secs, "bytes_transferred": nbytes, "num_descriptors": ndesc,
                            "time_to_first_token_seconds": ttft},
        }

    def metrics(self) -> bytes:
        return generate_latest(self.registry)

    # -- HTTP ---------------------------------------------------------------------------------------
    def handler(self):
        fe = self

        class H(BaseHTTPRequestHandler):
            protocol_version = "HTTP/1.1"

            def log_message(self, *a):
                pass

            def _send(self, code: int, payload: bytes, ctype: str = "application/json"):
                self.send_response(code)
                self.send_header("Content-Type", ctype)
                self.send_header("Content-Length", str(len(payload)))
                self.end_headers()
                self.wfile.write(payload)

            def do_GET(self):
                if self.path.rstrip("/") == "/v1/models":
                    self._send(200, json.dumps(fe.models()).encode())
                elif self.path.startswith("/metrics"):
                    self._send(200, fe.metrics(), "text/plain; version=0.0.4")
                elif self.path in ("/health", "/live"):
                    self._send(200, b'{"status":"healthy"}')
                else:
                    self._send(404, b'{"error":{"message":"not found"}}')

            def do_POST(self):
                if self.path.rstrip("/") != "/v1/chat/completions":
                    return self._send(404, b'{"error":{"message":"not found"}}')
                try:
                    body = json.loads(self.rfile.read(int(self.headers.get("Content-Length", "0"))) or b"{}")
                except (ValueError, json.JSONDecodeError):
                    return self._send(400, b'{"error":{"message":"invalid JSON"}}')
                code, resp = fe.chat(body)
                if code == 200 and body.get("stream"):
                    chunk = {"id": resp["id"], "object": "chat.completion.chunk", "created": resp["created"], "model": resp["model"],
                             "choices": [{"index": 0, "delta": {"role": "assistant", "content": resp["choices"][0]["message"]["content"]},
                                          "finish_reason": "stop"}]}
                    data = f"data: {json.dumps(chunk)}\n\ndata: [DONE]\n\n".encode()
                    return self._send(200, data, "text/event-stream")
                self._send(code, json.dumps(resp).encode())

        return H

    def serve(self, host: str = "127.0.0.1", port: int = 8000) -> ThreadingHTTPServer:
        srv = ThreadingHTTPServer((host, port), self.handler())
        threading.Thread(target=srv.serve_forever, daemon=True).start()
        return srv


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(description=__doc__.split("\n")[0])
    ap.add_argument("--model", default="Qwen/Qwen3-0.6B")  # chat.sh:4
    ap.add_argument("--host", default="127.0.0.1")
    ap.add_argument("--port", type=int, default=8000)
    ap.add_argument("--manifest", help="DynamoGraphDeployment yaml: worker layouts come from its backend")
    ap.add_argument("--prefill-ms-per-ktoken", type=float, default=0.0)
    ap.add_argument("--itl-ms", type=float, default=0.0)
    ap.add_argument("--resident-receiver", action="store_true",
                    help="the decode worker receives through kvx_queue_* (a resident kernel) instead of one launch per request")
    args = ap.parse_args(argv)
    p_layout, d_layout = 1, 0
    if args.manifest:
        from .manifest import load_manifest

        topo = load_manifest(args.manifest)
        if not topo.disaggregated:
            raise SystemExit("the manifest has no prefill/decode split")
        p_layout, d_layout = topo.role("prefill")[0].layout, topo.role("decode")[0].layout
    backend = InProcessHandoff(KV_SHAPES[args.model], p_layout=p_layout, d_layout=d_layout, resident_receiver=args.resident_receiver)
    fe = Frontend(args.model, backend, prefill_ms_per_ktoken=args.prefill_ms_per_ktoken, itl_ms=args.itl_ms)
    srv = fe.serve(args.host, args.port)
    print(f"kvx frontend on http://{args.host}:{srv.server_address[1]}/v1/chat/completions (model {args.model})", flush=True)
    try:
        while True:
            time.sleep(3600)
    except KeyboardInterrupt:
        srv.shutdown()
        backend.close()
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
