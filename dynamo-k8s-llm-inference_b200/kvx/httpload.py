"""HTTP load for `run-benchmarks.sh -u http://...`: the reference runner hands its endpoint URL to an
external load generator (run-benchmarks.sh:56-68 → aiperf, not installable here).  This stand-in drives an
OpenAI-compatible endpoint with the shape of the reference's own load script
(examples/dgdr/trtllm/multi_convos_parallel.sh: NUM_CONVOS three-turn conversations, at most CONCURRENCY
in flight, a conversation fails when any turn returns an error object; env knobs with the same names and
defaults, :11-13) and records what the dashboard plots per request: time to first token (the first
server-sent chunk with KVX_STREAM=1, else the whole reply) and, when the server reports it (`kvx.frontend`
does, in non-streamed replies), the KV-transfer time and bytes.
"""
from __future__ import annotations

import json
import os
import statistics
import threading
import time
import urllib.error
import urllib.request
from typing import Dict, List, Optional

TURNS = ("Give me a short overview of the topic.", "Now expand on the second point.", "Thanks! End with a one-line summary.")


def _endpoint(url: str) -> str:
    u = url.rstrip("/")
    return u if u.endswith("/chat/completions") else (u + "/chat/completions" if u.endswith("/v1") else u + "/v1/chat/completions")


def _read_sse(r, t0: float) -> Dict:
    """Server-sent events of a streamed completion: time to the first chunk, text of all chunks."""
    first, text = None, []
    for raw in r:
        line = raw.decode("utf-8", "replace").strip()
        if not line.startswith("data:"):
            continue
        data = line[5:].strip()
        if first is None:
            first = time.perf_counter() - t0
        if data == "[DONE]":
            break
        try:
            delta = json.loads(data)["choices"][0].get("delta", {})
        except (ValueError, KeyError, IndexError):
            return {"error": {"message": f"bad stream chunk: {data[:80]}"}}
        text.append(delta.get("content") or "")
    if first is None:
        return {"error": {"message": "empty stream"}}
    return {"choices": [{"message": {"role": "assistant", "content": "".join(text)}}], "_ttft_s": first}


def chat_once(endpoint: str, model: str, messages: List[Dict], timeout: float, stream: bool = False) -> Dict:
    payload = {"model": model, "messages": messages, "temperature": 0, "max_tokens": 64}
    if stream:
        payload["stream"] = True
    body = json.dumps(payload).encode()
    req = urllib.request.Request(endpoint, body, {"Content-Type": "application/json"})
    t0 = time.perf_counter()
    try:
        with urllib.request.urlopen(req, timeout=timeout) as r:
            if stream and "text/event-stream" in (r.headers.get("Content-Type") or ""):
                resp = _read_sse(r, t0)
            else:
                resp = json.loads(r.read())
    except urllib.error.HTTPError as e:
        try:
            resp = json.loads(e.read())
        except Exception:
            resp = {"error": {"message": f"HTTP {e.code}"}}
    except Exception as e:  # connection refused, timeout, bad JSON
        resp = {"error": {"message": str(e)}}
    resp["_latency_s"] = time.perf_counter() - t0
    return resp


def run_conversations(url: str, model: str, *, num_convos: Optional[int] = None, concurrency: Optional[int] = None,
                      prompt_tokens: Optional[int] = None, timeout: float = 60.0, stream: Optional[bool] = None) -> Dict:
    num_convos = int(os.environ.get("NUM_CONVOS", 10)) if num_convos is None else num_convos
    concurrency = int(os.environ.get("CONCURRENCY", 5)) if concurrency is None else concurrency
    prompt_tokens = int(os.environ.get("KVX_PROMPT_TOKENS", 1024)) if prompt_tokens is None else prompt_tokens
    stream = (os.environ.get("KVX_STREAM", "0") not in ("", "0")) if stream is None else stream
    endpoint = _endpoint(url)
    sem = threading.Semaphore(max(1, concurrency))
    records: List[Dict] = []
    failed: List[int] = []
    lock = threading.Lock()

    def convo(cid: int) -> None:
        with sem:
            msgs = [{"role": "system", "content": "You are a concise assistant."},
                    {"role": "user", "content": f"Conversation {cid}. " + "context " * max(0, prompt_tokens * 4 // 8)}]
            ok = True
            for turn, text in enumerate(TURNS):
                msgs.append({"role": "user", "content": text})
                resp = chat_once(endpoint, model, msgs, timeout, stream)
                if "error" in resp or not resp.get("choices"):
                    ok = False
                    rec = {"conversation": cid, "turn": turn, "error": str(resp.get("error", "no choices"))[:200],
                           "latency_s": resp["_latency_s"]}
                else:
                    msgs.append({"role": "assistant", "content": resp["choices"][0]["message"]["content"]})
                    kt = resp.get("kv_transfer") or {}
                    rec = {"conversation": cid, "turn": turn, "latency_s": resp["_latency_s"],
                           "ttft_s": resp.get("_ttft_s", resp["_latency_s"]),   # streamed: first chunk; else the whole reply
                           # the server's count when it reports one, else the same estimate kvx.frontend uses
                           "prompt_tokens": (resp.get("usage") or {}).get("prompt_tokens") or
                                            max(1, sum(len(str(m.get("content", ""))) + 4 for m in msgs[:-1]) // 4),
                           "xfer_time_seconds": kt.get("xfer_time_seconds"), "bytes_transferred": kt.get("bytes_transferred")}
                with lock:
                    records.append(rec)
                if not ok:
                    break
            if not ok:
                with lock:
                    failed.append(cid)

    t0 = time.perf_counter()
    threads = [threading.Thread(target=convo, args=(c,)) for c in range(1, num_convos + 1)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    wall = time.perf_counter() - t0
    good = [r for r in records if "error" not in r]
    lat = sorted(r["ttft_s"] for r in good)
    xfer = [r for r in good if r.get("xfer_time_seconds")]

    def pct(v, p):
        return v[min(len(v) - 1, int(p * len(v)))] if v else None

    summary = {
        "endpoint": endpoint, "model": model, "conversations": num_convos, "concurrency": concurrency, "turns_per_conversation": len(TURNS),
        "requests": len(records), "failed_conversations": sorted(failed), "wall_seconds": wall, "streamed": bool(stream),
        "requests_per_second": len(good) / wall if wall > 0 else None,
        "time_to_first_token_seconds": {"p50": pct(lat, 0.5), "p90": pct(lat, 0.9), "max": lat[-1] if lat else None,
                                        "mean": statistics.mean(lat) if lat else None},
    }
    if xfer:
        tot_b, tot_t = sum(r["bytes_transferred"] for r in xfer), sum(r["xfer_time_seconds"] for r in xfer)
        summary["kv_transfer"] = {"requests": len(xfer), "bytes_transferred": tot_b, "xfer_time_seconds": tot_t,
                                  "throughput_GBps": tot_b / tot_t / 1e9 if tot_t > 0 else None}
    return {"summary": summary, "requests": records}


def run_http_benchmark(url: str, model: str, run_dir: str) -> Dict:
    os.makedirs(run_dir, exist_ok=True)
    out = run_conversations(url, model)
    with open(os.path.join(run_dir, "results.json"), "w") as f:
        json.dump(out, f, indent=1)
    return out["summary"]
