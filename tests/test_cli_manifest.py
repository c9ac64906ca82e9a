"""Outer boundary kept from the reference: the run-benchmarks.sh flag surface and the manifest-driven
prefill/decode topology."""
import json
import os
import subprocess
import sys

import pytest

from conftest import ROOT

RUNNER = os.path.join(ROOT, "run-benchmarks.sh")

DISAGG = """
apiVersion: nvidia.com/v1alpha1
kind: DynamoGraphDeployment
metadata: {name: t-disagg}
spec:
  services:
    Frontend: {componentType: frontend, replicas: 1}
    Dec:
      componentType: worker
      subComponentType: decode
      replicas: 2
      resources: {limits: {gpu: "2"}}
      extraPodSpec:
        mainContainer:
          image: example/sglang-runtime:1
          command: [python3, -m, dynamo.sglang]
          args: [--model-path, meta-llama/Llama-3.1-70B, --page-size, "16", --tp, "2", --disaggregation-mode, decode]
    Pre:
      componentType: worker
      replicas: 1
      resources: {limits: {gpu: "4"}}
      extraPodSpec:
        mainContainer:
          command: [python3, -m, dynamo.vllm]
          args: [--model, meta-llama/Llama-3.1-70B, --is-prefill-worker, --tensor-parallel-size=4]
"""


def _run(*args, env=None):
    return subprocess.run([RUNNER, *args], capture_output=True, text=True, env=dict(os.environ, **(env or {})))


def test_usage_and_exit_codes_match_reference_runner():
    r = _run("-h")
    assert r.returncode == 0 and r.stderr.startswith("Usage: run-benchmarks.sh -u <api-url> -m <model> -o <output-dir> -b <benchmark-name>")
    assert r.stdout == ""
    r = _run("-u", "kvx://local", "-m", "Qwen/Qwen3-0.6B")
    assert r.returncode == 1 and r.stderr.startswith("Missing required options.\nUsage:")
    r = _run("-q")
    assert r.returncode == 1 and r.stderr.startswith("Unknown option: -q\nUsage:")
    r = _run("-u", "a", "-m")
    assert r.returncode == 1 and r.stderr.startswith("Missing value for -m\nUsage:")
    r = _run()
    assert r.returncode == 1 and "Missing required options." in r.stderr


def test_bad_endpoint_and_model_fail_cleanly(tmp_path):
    r = _run("-u", "ftp://10.0.0.1:30080", "-m", "Qwen/Qwen3-0.6B", "-o", str(tmp_path), "-b", "x")
    assert r.returncode == 1 and "unsupported endpoint scheme" in r.stderr
    # an HTTP endpoint nobody listens on: every conversation fails, the runner exits 1 (no hang)
    r = _run("-u", "http://127.0.0.1:1", "-m", "Qwen/Qwen3-0.6B", "-o", str(tmp_path), "-b", "dead", env={"NUM_CONVOS": "2"})
    assert r.returncode == 1 and '"failed_conversations": [1, 2]' in r.stdout
    r = _run("-u", "kvx://local", "-m", "no/such-model", "-o", str(tmp_path), "-b", "x")
    assert r.returncode == 1 and "unknown model" in r.stderr


def test_getopts_semantics():
    from kvx.cli import parse_args

    assert parse_args(["-ukvx://l", "-m", "M", "-o", "O", "-b", "B", "-p"])[0] == ("kvx://l", "M", "O", "B", True)
    assert parse_args(["-pu", "U", "-m", "M", "-o", "O", "-b", "B"])[0] == ("U", "M", "O", "B", True)
    assert parse_args(["-u", "U", "-m", "M", "-o", "O", "-b", "B", "extra", "-p"])[0] == ("U", "M", "O", "B", False)


def test_manifest_topology(tmp_path):
    from kvx import HND, NHD
    from kvx.manifest import load_manifest
    from kvx.harness import rank_plans

    p = tmp_path / "m.yaml"
    p.write_text(DISAGG)
    t = load_manifest(str(p))
    assert t.disaggregated and t.name == "t-disagg" and t.gpus_needed() == 8
    dec, pre = t.role("decode")[0], t.role("prefill")[0]
    assert (dec.replicas, dec.gpus, dec.tp, dec.page_size, dec.backend, dec.layout) == (2, 2, 2, 16, "sglang", NHD)
    assert (pre.replicas, pre.gpus, pre.tp, pre.backend, pre.layout, pre.model) == (1, 4, 4, "vllm", HND, "meta-llama/Llama-3.1-70B")
    plans = rank_plans(t, 8)
    assert [(x.role, x.replica, x.tp_rank, x.device) for x in plans] == [
        ("prefill", 0, 0, 0), ("prefill", 0, 1, 1), ("prefill", 0, 2, 2), ("prefill", 0, 3, 3),
        ("decode", 0, 0, 4), ("decode", 0, 1, 5), ("decode", 1, 0, 6), ("decode", 1, 1, 7)]
    (tmp_path / "bad.yaml").write_text("kind: ConfigMap\n")
    with pytest.raises(ValueError):
        load_manifest(str(tmp_path / "bad.yaml"))


def test_reference_manifests_if_present():
    """The reference's own disagg manifests parse to 1 prefill + 1 decode worker, one GPU each (only where
    /root/reference is mounted; never on the GPU box)."""
    base = "/root/reference/examples/deploy"
    if not os.path.isdir(base):
        pytest.skip("reference not mounted")
    from kvx.manifest import load_manifest

    for be in ("vllm", "sglang", "trtllm"):
        t = load_manifest(f"{base}/{be}/disagg.yaml")
        assert t.backend == be and t.disaggregated and t.gpus_needed() == 2
        assert all(w.replicas == 1 and w.gpus == 1 and w.tp == 1 and w.page_size == 16 for w in t.workers)
        assert not load_manifest(f"{base}/{be}/agg.yaml").disaggregated
        if be == "sglang":   # the one manifest that names the handshake port and the transfer backend
            assert {w.bootstrap_port for w in t.workers} == {12345} and {w.transfer_backend for w in t.workers} == {"nixl"}
        else:
            assert all(w.bootstrap_port is None for w in t.workers)


def test_build_job_from_url(tmp_path):
    from kvx.cli import build_job

    p = tmp_path / "m.yaml"
    p.write_text(DISAGG)
    topo, cfg = build_job(f"kvx://local?manifest={p}&isl=1024,4096&requests=9&concurrency=2&stages=3", "meta-llama/Llama-3.1-70B")
    assert topo.name == "t-disagg" and len(cfg.isl) == 9 and set(cfg.isl) <= {1024, 4096} and cfg.concurrency == 2
    assert cfg.options == {"stages": 3}
    topo, cfg = build_job("kvx://local?prefill=4&decode=4&isl=8192&requests=3", "meta-llama/Llama-3.1-8B")
    assert [w.replicas for w in topo.workers] == [4, 4] and cfg.isl == [8192] * 3 and cfg.batch == 1
    _, cfg = build_job("kvx://local?prefill=1&decode=1&isl=128&requests=16&batch=8&max_ctas=32", "meta-llama/Llama-3.1-8B")
    assert cfg.batch == 8 and cfg.options == {"max_ctas": 32}


def test_plots_from_results(tmp_path):
    from kvx.plot import plot_dir

    d = tmp_path / "bench"
    d.mkdir()
    res = {"summary": {"topology": {"name": "t"}}, "decode_ranks": [{"requests": [
        {"isl": 1024, "throughput_GBps": 500.0, "time_to_first_token_added_ms": 0.3},
        {"isl": 8192, "throughput_GBps": 780.0, "time_to_first_token_added_ms": 1.4}]}]}
    (d / "results.json").write_text(json.dumps(res))
    out = plot_dir(str(tmp_path))
    assert len(out) == 2 and all(open(p).read().startswith("<svg") for p in out)


@pytest.mark.gpu
def test_runner_end_to_end_on_gpu(tmp_path):
    """The whole outer path: runner flags → topology → worker processes → pull stream → results + plots."""
    p = tmp_path / "m.yaml"
    p.write_text(DISAGG.replace('replicas: 2', 'replicas: 1').replace('gpu: "2"', 'gpu: "1"').replace('"2", --disagg', '"1", --disagg')
                 .replace('gpu: "4"', 'gpu: "1"').replace("--tensor-parallel-size=4", "--tensor-parallel-size=1")
                 .replace("Llama-3.1-70B", "Llama-3.2-1B-Instruct"))
    out = tmp_path / "results"
    r = _run("-u", f"kvx://local?manifest={p}&isl=1024,2048&requests=6&concurrency=2", "-m", "meta-llama/Llama-3.2-1B-Instruct",
             "-o", str(out), "-b", "deployment-kvx", "-p")
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3000:]
    res = json.load(open(out / "deployment-kvx" / "results.json"))
    s = res["summary"]
    assert s["requests"] == 6 and s["verified_bit_exact"] is True and s["prefill_blocks_released"] is True
    assert os.path.exists(out / "deployment-kvx" / "throughput_GBps.svg")


@pytest.mark.gpu
def test_runner_tp_reshard_on_gpu(tmp_path):
    """Config #4 in small: TP4 prefill → TP2 decode (ranks share GPUs when fewer than 6 are visible)."""
    out = tmp_path / "results"
    r = _run("-u", "kvx://local?prefill=1&decode=1&ptp=4&dtp=2&isl=512&requests=3&layout=nhd", "-m", "meta-llama/Llama-3.2-1B-Instruct",
             "-o", str(out), "-b", "tp")
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3000:]
    s = json.load(open(out / "tp" / "results.json"))["summary"]
    assert s["verified_bit_exact"] is True and s["ranks"] == 6


def test_build_job_stream_options():
    from kvx.cli import build_job

    _, cfg = build_job("kvx://local?prefill=4&decode=4&isl=1024,32768&requests=12&arrival=bursty&gap_ms=0.5&route=sched",
                       "meta-llama/Llama-3.1-8B")
    assert cfg.route == "sched" and len(cfg.ready_ms) == 12 and cfg.ready_ms == sorted(cfg.ready_ms)
    _, cfg = build_job("kvx://local?prefill=2&decode=2&isl=1024,2048&requests=5&per_pair_seed=1", "meta-llama/Llama-3.1-8B")
    assert len(cfg.per_replica_isl) == 2 and cfg.per_replica_isl[0] != cfg.per_replica_isl[1]
    _, cfg = build_job("kvx://local?isl=8192&requests=2&layerwise=8&prefill_ms=20&corrupt=0,1", "meta-llama/Llama-3.1-8B")
    assert cfg.layerwise == 8 and cfg.prefill_ms == 20.0 and cfg.corrupt == (0, 1)
    with pytest.raises(ValueError):
        build_job("kvx://local?arrival=poisson", "meta-llama/Llama-3.1-8B")


M1B = "meta-llama/Llama-3.2-1B-Instruct"


@pytest.mark.gpu
def test_runner_verifies_every_request_and_fails_on_one_bad_byte(tmp_path):
    """Every request of every decode rank is compared bit for bit when it lands, and the whole decode pool at the
    end; one flipped byte in one landed block of one request turns the run into exit code 1."""
    out = tmp_path / "r"
    url = "kvx://local?prefill=1&decode=1&isl=256,1024&requests=12&concurrency=2"
    r = _run("-u", url, "-m", M1B, "-o", str(out), "-b", "good")
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3000:]
    s = json.load(open(out / "good" / "results.json"))["summary"]
    assert s["verified_bit_exact"] is True and s["requests_verified"] == s["requests"] == 12
    assert s["canary_blocks_checked"] > 0 and s["blocks_checked"] > 0 and s["releases_observed"] == s["releases_expected"] == 12
    r = _run("-u", url + "&corrupt=0,7", "-m", M1B, "-o", str(out), "-b", "bad")
    assert r.returncode == 1, r.stdout[-2000:]
    s = json.load(open(out / "bad" / "results.json"))["summary"]
    assert s["verified_bit_exact"] is False and s["requests_verified"] == 11
    # verification switched off by the caller: reported as not verified, with a warning, never as verified
    r = _run("-u", url + "&verify=0", "-m", M1B, "-o", str(out), "-b", "off")
    assert r.returncode == 0 and "NOT checked" in r.stderr
    s = json.load(open(out / "off" / "results.json"))["summary"]
    assert s["verified_bit_exact"] is None and s["verification"].startswith("off")


@pytest.mark.gpu
def test_runner_nhd_prefill_pools_are_verified_too(tmp_path):
    """SGLang-style token-major prefill pools (backend=sglang → NHD on both sides) and NHD prefill → HND decode."""
    out = tmp_path / "r"
    for name, extra in (("nhd_nhd", ""), ("nhd_hnd", "&layout=hnd")):
        r = _run("-u", f"kvx://local?prefill=1&decode=1&backend=sglang&isl=512&requests=4{extra}", "-m", M1B, "-o", str(out), "-b", name)
        assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3000:]
        s = json.load(open(out / name / "results.json"))["summary"]
        assert s["verified_bit_exact"] is True and s["requests_verified"] == 4


@pytest.mark.gpu
def test_runner_routes_and_release_accounting(tmp_path):
    """2P2D: spread routing (each decode replica reads both prefill replicas) and the router's bookings on a
    bursty stream; every prefill rank must see one release per request it served, from the right consumer."""
    out = tmp_path / "r"
    r = _run("-u", "kvx://local?prefill=2&decode=2&isl=256,512&requests=6&route=spread&concurrency=2", "-m", M1B, "-o", str(out), "-b", "spread")
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3000:]
    res = json.load(open(out / "spread" / "results.json"))
    s = res["summary"]
    assert s["verified_bit_exact"] is True and s["requests_verified"] == 12 and s["releases_observed"] == s["releases_expected"] == 12
    assert {q["prefill_replica"] for q in res["decode_ranks"][0]["requests"]} == {0, 1}
    r = _run("-u", "kvx://local?prefill=2&decode=2&isl=256,2048&requests=10&arrival=bursty&gap_ms=0.2&route=sched", "-m", M1B,
             "-o", str(out), "-b", "sched")
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3000:]
    s = json.load(open(out / "sched" / "results.json"))["summary"]
    assert s["verified_bit_exact"] is True and s["requests_verified"] == s["requests"] == 10 and s["arrival_driven"] is True


@pytest.mark.gpu
def test_runner_layerwise_handoff(tmp_path):
    """Layer-wise handoff through the worker API (publish_layers / pull_layers): 4 layer groups behind a
    synthetic prefill, also with TP fan-in (each decode rank gated by two prefill ranks' flags)."""
    out = tmp_path / "r"
    for name, extra in (("lw", ""), ("lw_tp", "&ptp=4&dtp=2")):
        r = _run("-u", f"kvx://local?prefill=1&decode=1&isl=512&requests=3&layerwise=4&prefill_ms=2{extra}", "-m", M1B, "-o", str(out), "-b", name)
        assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3000:]
        res = json.load(open(out / name / "results.json"))
        assert res["summary"]["verified_bit_exact"] is True and res["summary"]["layerwise"] == 4
        assert all(q["added_after_prefill_ms"] is not None and q["added_after_prefill_ms"] >= 0 for q in res["decode_ranks"][0]["requests"])


@pytest.mark.gpu
def test_runner_through_the_resident_receiver(tmp_path):
    """queue=1: the decode worker pulls every request through kvx_queue_* (no launch per request); every request and the
    whole pool are verified as in every other run, and one flipped byte still fails the run."""
    out = tmp_path / "r"
    url = "kvx://local?prefill=1&decode=1&isl=128,512,2048&requests=14&concurrency=2&queue=1"
    r = _run("-u", url, "-m", M1B, "-o", str(out), "-b", "q")
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3000:]
    s = json.load(open(out / "q" / "results.json"))["summary"]
    assert s["resident_receiver"] is True and s["verified_bit_exact"] is True and s["requests_verified"] == s["requests"] == 14
    assert s["releases_observed"] == s["releases_expected"] == 14
    r = _run("-u", url + "&corrupt=0,9", "-m", M1B, "-o", str(out), "-b", "qbad")
    assert r.returncode == 1
