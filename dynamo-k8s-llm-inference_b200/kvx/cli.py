"""Benchmark runner with the reference runner's interface (run-benchmarks.sh:4-38 in the reference):

  -u <api-url> -m <model> -o <output-dir> -b <benchmark-name> [-p] [-h]

Same required flags, same messages, same exit codes.  `-u` is either an OpenAI-compatible HTTP endpoint, as
in the reference (driven by kvx/httpload.py with the shape of the reference's multi_convos_parallel.sh), or a
transport URL:  kvx://local?manifest=<DynamoGraphDeployment yaml>[&isl=8192][&requests=64][&concurrency=1]
           kvx://local?prefill=4&decode=4[&ptp=4&dtp=2][&isl=1024,2048,...][&layout=nhd][&route=spread][&batch=8]
           ...&arrival=bursty[&gap_ms=1.0][&burst=4]&route=sched   one bursty stream, booked by kvx/router.py
           ...&layerwise=8&prefill_ms=20                            hand over in 8 layer groups behind a synthetic prefill
           ...&queue=1                                              pull through the resident receiver (kvx_queue_*)
Every request is verified bit for bit unless verify=0 is given; an unverified run says so and is never "verified".
Results land under <output-dir>/<benchmark-name>/results.json; -p adds SVG plots from that tree.
"""
from __future__ import annotations

import json
import os
import sys
from typing import List, Optional
from urllib.parse import parse_qs, urlparse

USAGE = """Usage: run-benchmarks.sh -u <api-url> -m <model> -o <output-dir> -b <benchmark-name>
  -u  API endpoint URL
  -m  Model name
  -o  Output directory (relative to ./dynamo or absolute)
  -b  Benchmark name
  -p  Generate plots after benchmark run
"""


def usage() -> None:
    sys.stderr.write(USAGE)


def parse_args(argv: List[str]):
    """getopts ":u:m:o:b:ph" semantics: options stop at the first non-option word."""
    opts = {"u": "", "m": "", "o": "", "b": ""}
    plots = False
    i = 0
    while i < len(argv):
        a = argv[i]
        if a == "--":
            break
        if not a.startswith("-") or a == "-":
            break
        j = 1
        while j < len(a):
            c = a[j]
            if c in opts:
                val = a[j + 1:] if j + 1 < len(a) else (argv[i + 1] if i + 1 < len(argv) else None)
                if val is None:
                    sys.stderr.write(f"Missing value for -{c}\n")
                    usage()
                    return None, 1
                if j + 1 >= len(a):
                    i += 1
                opts[c] = val
                break
            if c == "p":
                plots = True
            elif c == "h":
                usage()
                return None, 0
            else:
                sys.stderr.write(f"Unknown option: -{c}\n")
                usage()
                return None, 1
            j += 1
        i += 1
    if not all(opts.values()):
        sys.stderr.write("Missing required options.\n")
        usage()
        return None, 1
    return (opts["u"], opts["m"], opts["o"], opts["b"], plots), None


def build_job(url: str, model: str):
    """Topology + stream configuration from the transport URL."""
    from .harness import StreamConfig, draw_isls
    from .manifest import load_manifest, synthetic_topology
    from .shapes import KV_SHAPES

    u = urlparse(url)
    if u.scheme != "kvx":
        raise ValueError(f"unsupported endpoint scheme '{u.scheme}': expected kvx://local?...")
    if model not in KV_SHAPES:
        raise ValueError(f"unknown model '{model}': known KV shapes are {sorted(KV_SHAPES)}")
    q = {k: v[-1] for k, v in parse_qs(u.query).items()}
    if "manifest" in q:
        topo = load_manifest(q["manifest"])
    else:
        topo = synthetic_topology(int(q.get("prefill", 1)), int(q.get("decode", 1)), int(q.get("ptp", 1)), int(q.get("dtp", 1)),
                                  q.get("backend", "vllm"), model)
    layout = {"nhd": 0, "hnd": 1}.get(q.get("layout", "").lower())
    requests = int(q.get("requests", 16))
    isls = [int(x) for x in q.get("isl", "8192").split(",")]
    seed = int(q.get("seed", 5))
    isl = [isls[0]] * requests if len(isls) == 1 else draw_isls(requests, isls, seed)
    ready = None
    if q.get("arrival"):
        if q["arrival"] != "bursty":
            raise ValueError(f"unknown arrival process '{q['arrival']}': expected bursty")
        from .router import skewed_arrivals

        ready = skewed_arrivals(requests, seed + 1, float(q.get("gap_ms", 1.0)), int(q.get("burst", 4)))
    n_dec = sum(w.replicas for w in topo.role("decode"))
    per_rep = None
    if q.get("per_pair_seed") == "1" and len(isls) > 1 and ready is None:  # config #5: every pair draws its own stream
        per_rep = [draw_isls(requests, isls, seed + 17 * d) for d in range(n_dec)]
    corrupt = tuple(int(x) for x in q["corrupt"].split(",")) if q.get("corrupt") else None
    opts = {k: int(q[k]) for k in ("engine", "stages", "stage_bytes", "ctas_per_sm", "store_mode", "vec_bytes", "consumer_warps", "l2_hint", "pdl",
                                   "param_tables", "max_ctas") if k in q}
    cfg = StreamConfig(model=model, isl=isl, concurrency=int(q.get("concurrency", 1)), verify=q.get("verify", "1") != "0",
                       decode_layout=layout, options=opts, route=q.get("route", "pair"), batch=int(q.get("batch", 1)),
                       ready_ms=ready, per_replica_isl=per_rep, layerwise=int(q.get("layerwise", 1)),
                       prefill_ms=float(q.get("prefill_ms", 0.0)), corrupt=corrupt, queue=q.get("queue", "0") == "1")
    return topo, cfg


def main(argv: Optional[List[str]] = None) -> int:
    parsed, rc = parse_args(list(sys.argv[1:] if argv is None else argv))
    if parsed is None:
        return rc
    url, model, out_dir, name, plots = parsed
    if urlparse(url).scheme in ("http", "https"):
        # the reference's own usage: an OpenAI-compatible endpoint (the Frontend pod, or `python -m kvx.frontend`)
        from .httpload import run_http_benchmark

        os.makedirs(out_dir, exist_ok=True)
        summary = run_http_benchmark(url, model, os.path.join(out_dir, name))
        print(json.dumps(summary))
        if plots:
            from .plot import plot_dir

            for p in plot_dir(out_dir):
                print(f"plot: {p}")
        return 1 if (summary["failed_conversations"] or not summary["requests"]) else 0
    try:
        topo, cfg = build_job(url, model)
    except (ValueError, OSError) as e:
        sys.stderr.write(f"{e}\n")
        return 1
    os.makedirs(out_dir, exist_ok=True)
    run_dir = os.path.join(out_dir, name)
    from .harness import run_stream

    try:
        summary = run_stream(topo, cfg, run_dir)
    except Exception as e:  # any failed transfer = failed benchmark, like a non-zero exit of the reference's runner
        sys.stderr.write(f"benchmark failed: {e}\n")
        return 1
    print(json.dumps(summary))
    if plots:
        from .plot import plot_dir

        for p in plot_dir(out_dir):
            print(f"plot: {p}")
    if summary.get("verified_bit_exact") is None:  # only ever the caller's own verify=0: say so, never call it verified
        sys.stderr.write("warning: verify=0 — the bytes that landed were NOT checked\n")
        return 0 if (not cfg.verify and summary.get("prefill_blocks_released")) else 1
    ok = (summary["verified_bit_exact"] is True and summary.get("requests_verified") == summary.get("requests")
          and summary.get("prefill_blocks_released"))
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
