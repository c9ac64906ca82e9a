#!/usr/bin/env bash
# Same interface as the reference's run-benchmarks.sh (-u/-m/-o/-b[/-p/-h]); the benchmark behind it
# is the synthetic prefill->decode KV handoff stream (dynamo-k8s-llm-inference_b200/kvx/cli.py).
set -euo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
export PYTHONPATH="${HERE}/dynamo-k8s-llm-inference_b200${PYTHONPATH:+:${PYTHONPATH}}"
exec python3 -m kvx.cli "$@"
