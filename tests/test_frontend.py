"""OpenAI-compatible Frontend stand-in: the HTTP contract chat.sh / multi_convos_parallel.sh / aiperf rely on,
and the metric names the reference's Grafana dashboard queries."""
import json
import urllib.error
import urllib.request

import pytest

from kvx.frontend import Frontend, count_tokens


class FakeBackend:
    """Host-logic double: the HTTP layer is tested on CPU; the GPU test below uses the real workers."""

    def __init__(self):
        self.calls = []

    def handoff(self, isl):
        self.calls.append(isl)
        if isl > 100000:
            raise RuntimeError("boom")
        return 0.001, isl * 1000, 56 * isl


def _post(url, body):
    req = urllib.request.Request(url, json.dumps(body).encode(), {"Content-Type": "application/json", "Authorization": "Bearer dummy"})
    try:
        with urllib.request.urlopen(req) as r:
            return r.status, json.loads(r.read())
    except urllib.error.HTTPError as e:
        return e.code, json.loads(e.read())


@pytest.fixture()
def server():
    be = FakeBackend()
    fe = Frontend("Qwen/Qwen3-0.6B", be)
    srv = fe.serve("127.0.0.1", 0)
    yield f"http://127.0.0.1:{srv.server_address[1]}", be
    srv.shutdown()


def test_chat_completion_contract(server):
    base, be = server
    msgs = [{"role": "system", "content": "You are terse."}, {"role": "user", "content": "hello " * 200}]
    code, resp = _post(base + "/v1/chat/completions", {"model": "Qwen/Qwen3-0.6B", "messages": msgs, "temperature": 0, "max_tokens": 512})
    assert code == 200 and "error" not in resp
    assert resp["choices"][0]["message"]["role"] == "assistant" and resp["choices"][0]["message"]["content"]
    assert resp["usage"]["prompt_tokens"] == count_tokens(msgs) == be.calls[0]
    assert resp["kv_transfer"]["bytes_transferred"] == be.calls[0] * 1000
    with urllib.request.urlopen(base + "/v1/models") as r:
        assert json.loads(r.read())["data"][0]["id"] == "Qwen/Qwen3-0.6B"


def test_errors_are_json_error_objects(server):
    base, _ = server
    code, resp = _post(base + "/v1/chat/completions", {"model": "other", "messages": [{"role": "user", "content": "x"}]})
    assert code == 404 and "message" in resp["error"]
    code, resp = _post(base + "/v1/chat/completions", {"model": "Qwen/Qwen3-0.6B", "messages": []})
    assert code == 400 and "error" in resp
    code, resp = _post(base + "/v1/chat/completions", {"model": "Qwen/Qwen3-0.6B", "messages": [{"role": "user", "content": "y" * 500000}]})
    assert code == 500 and "KV handoff failed" in resp["error"]["message"]


def test_dashboard_metric_names(server):
    base, _ = server
    for _ in range(3):
        _post(base + "/v1/chat/completions", {"model": "Qwen/Qwen3-0.6B", "messages": [{"role": "user", "content": "hi there"}]})
    with urllib.request.urlopen(base + "/metrics") as r:
        text = r.read().decode()
    for name in ("dynamo_frontend_requests_total", "dynamo_frontend_time_to_first_token_seconds_sum",
                 "dynamo_frontend_time_to_first_token_seconds_count", "dynamo_frontend_inter_token_latency_seconds_sum",
                 "dynamo_frontend_request_duration_seconds_count", "dynamo_frontend_input_sequence_tokens_sum",
                 "dynamo_frontend_output_sequence_tokens_count", "vllm:nixl_xfer_time_seconds_sum",
                 "vllm:nixl_bytes_transferred_sum", "vllm:nixl_num_descriptors_count"):
        assert name in text, name
    assert 'dynamo_frontend_requests_total{model="Qwen/Qwen3-0.6B",namespace="default",status="success"} 3.0' in text


def test_streaming_is_sse(server):
    base, _ = server
    req = urllib.request.Request(base + "/v1/chat/completions", json.dumps(
        {"model": "Qwen/Qwen3-0.6B", "messages": [{"role": "user", "content": "hi"}], "stream": True}).encode(),
        {"Content-Type": "application/json"})
    with urllib.request.urlopen(req) as r:
        body = r.read().decode()
        assert r.headers["Content-Type"] == "text/event-stream"
    assert body.startswith("data: {") and body.rstrip().endswith("data: [DONE]")


@pytest.mark.gpu
@pytest.mark.parametrize("resident_receiver", [False, True], ids=["launch_per_request", "resident_receiver"])
def test_frontend_runs_real_handoffs(resident_receiver):
    """Three conversations' worth of requests through the real prefill/decode workers, with a launch per request and
    through the decode worker's resident receiver."""
    import kvx
    from kvx.frontend import InProcessHandoff

    be = InProcessHandoff(kvx.KV_SHAPES["meta-llama/Llama-3.2-1B-Instruct"], max_isl=4096, resident_receiver=resident_receiver)
    fe = Frontend("meta-llama/Llama-3.2-1B-Instruct", be)
    srv = fe.serve("127.0.0.1", 0)
    base = f"http://127.0.0.1:{srv.server_address[1]}"
    try:
        for words in (40, 900, 3000):
            code, resp = _post(base + "/v1/chat/completions", {"model": "meta-llama/Llama-3.2-1B-Instruct",
                                                                "messages": [{"role": "user", "content": "token " * words}], "max_tokens": 8})
            assert code == 200, resp
            kv = resp["kv_transfer"]
            n = min(256, -(-resp["usage"]["prompt_tokens"] // 16))  # prompts are clipped to max_isl
            assert kv["bytes_transferred"] == n * 32 * 16384 and kv["num_descriptors"] == n * 32 and kv["xfer_time_seconds"] > 0
        # blocks were freed on both sides after every request
        assert be.p.blocks.free_blocks == be.d.blocks.free_blocks == 3 * 256
    finally:
        srv.shutdown()
        be.close()


def test_runner_drives_an_http_endpoint(server, tmp_path, monkeypatch):
    """run-benchmarks.sh -u http://... : bounded-concurrency three-turn conversations against the endpoint,
    results under <out>/<name>/results.json, exit 1 when any conversation fails (as the reference's script)."""
    import os

    from kvx import cli

    base, be = server
    monkeypatch.setenv("NUM_CONVOS", "4")
    monkeypatch.setenv("CONCURRENCY", "2")
    monkeypatch.setenv("KVX_PROMPT_TOKENS", "64")
    rc = cli.main(["-u", base + "/v1/chat/completions", "-m", "Qwen/Qwen3-0.6B", "-o", str(tmp_path), "-b", "http1"])
    assert rc == 0
    res = json.load(open(os.path.join(tmp_path, "http1", "results.json")))
    s = res["summary"]
    assert s["requests"] == 12 and not s["failed_conversations"] and len(be.calls) == 12
    assert s["kv_transfer"]["requests"] == 12 and s["time_to_first_token_seconds"]["p50"] > 0
    prompts = [r["prompt_tokens"] for r in res["requests"] if r["conversation"] == 1]
    assert prompts == sorted(prompts) and prompts[0] < prompts[-1]      # the conversation grows turn by turn
    # a wrong model name is an error object on every first turn: the run fails
    rc = cli.main(["-u", base, "-m", "meta-llama/Llama-3.1-8B", "-o", str(tmp_path), "-b", "http2"])
    assert rc == 1
    assert len(json.load(open(os.path.join(tmp_path, "http2", "results.json")))["summary"]["failed_conversations"]) == 4


@pytest.mark.gpu
def test_runner_against_real_frontend(tmp_path, monkeypatch):
    """The whole chain the reference describes, on one box: runner (-u http://...) -> Frontend stand-in ->
    prefill/decode workers -> a real KV handoff per request."""
    import os

    import kvx
    from kvx import cli
    from kvx.frontend import InProcessHandoff

    be = InProcessHandoff(kvx.KV_SHAPES["Qwen/Qwen3-0.6B"], max_isl=8192)
    srv = Frontend("Qwen/Qwen3-0.6B", be).serve("127.0.0.1", 0)
    monkeypatch.setenv("NUM_CONVOS", "6")
    monkeypatch.setenv("CONCURRENCY", "3")
    monkeypatch.setenv("KVX_PROMPT_TOKENS", "2048")
    try:
        rc = cli.main(["-u", f"http://127.0.0.1:{srv.server_address[1]}", "-m", "Qwen/Qwen3-0.6B", "-o", str(tmp_path), "-b", "chain"])
        assert rc == 0
        s = json.load(open(os.path.join(tmp_path, "chain", "results.json")))["summary"]
        assert s["requests"] == 18 and not s["failed_conversations"]
        assert s["kv_transfer"]["bytes_transferred"] > 18 * 128 * 56 * 32768 and s["kv_transfer"]["throughput_GBps"] > 100
        assert be.p.blocks.free_blocks == be.d.blocks.free_blocks
    finally:
        srv.shutdown()
        be.close()


@pytest.mark.gpu
def test_example_script_runs():
    import os
    import subprocess
    import sys

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, os.path.join(root, "examples", "one_request.py")], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "prefill blocks released" in r.stdout, r.stdout + r.stderr


def test_runner_streams_for_ttft(server, tmp_path, monkeypatch):
    """KVX_STREAM=1: the load asks for server-sent events and takes time-to-first-token from the first chunk."""
    import os

    from kvx import cli

    base, be = server
    for k, v in (("NUM_CONVOS", "2"), ("CONCURRENCY", "2"), ("KVX_PROMPT_TOKENS", "32"), ("KVX_STREAM", "1")):
        monkeypatch.setenv(k, v)
    assert cli.main(["-u", base, "-m", "Qwen/Qwen3-0.6B", "-o", str(tmp_path), "-b", "sse", "-p"]) == 0
    assert os.path.exists(os.path.join(tmp_path, "sse", "time_to_first_token_ms.svg"))
    res = json.load(open(os.path.join(tmp_path, "sse", "results.json")))
    assert res["summary"]["streamed"] and res["summary"]["requests"] == 6 and not res["summary"]["failed_conversations"]
    assert all(0 < r["ttft_s"] <= r["latency_s"] for r in res["requests"])
