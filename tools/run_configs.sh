#!/usr/bin/env bash
# BASELINE.json configs #3-#5 (plus incast / fan-out) through the runner, as many as the visible GPUs allow.
#   tools/run_configs.sh <output-dir>
# Each run leaves <output-dir>/<name>/results.json; a one-line summary per run goes to <output-dir>/summaries.jsonl.
set -uo pipefail
OUT="${1:-gpurun_out/configs}"
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")/.." && pwd)"
NGPU="$(nvidia-smi -L | wc -l)"
mkdir -p "$OUT"
: > "$OUT/summaries.jsonl"
run() {  # name model url
  echo "== $1" >&2
  if "$HERE/run-benchmarks.sh" -u "$3" -m "$2" -o "$OUT" -b "$1" > "$OUT/$1.summary.json" 2> "$OUT/$1.err"; then
    python3 - "$1" "$OUT/$1.summary.json" >> "$OUT/summaries.jsonl" <<'PY'
import json, sys
s = json.load(open(sys.argv[2]))
keep = ("requests", "concurrency", "aggregate_throughput_GBps", "per_decode_rank_ingress_GBps", "per_request_throughput_GBps_median",
        "pct_of_nvlink_900", "xfer_time_seconds", "time_to_first_token_added_ms", "verified_bit_exact", "prefill_blocks_released")
print(json.dumps({"name": sys.argv[1], **{k: s[k] for k in keep if k in s}}))
PY
  else
    echo "{\"name\": \"$1\", \"failed\": true}" >> "$OUT/summaries.jsonl"; tail -5 "$OUT/$1.err" >&2
  fi
}
L8=meta-llama/Llama-3.1-8B
L70=meta-llama/Llama-3.1-70B
if [ "$NGPU" -ge 2 ] && [ -z "${SKIP_CFG3:-}" ]; then
  run cfg3_1p1d_c1 $L8 'kvx://local?prefill=1&decode=1&isl=8192&requests=64&concurrency=1&layout=nhd'
  run cfg3_1p1d_c4 $L8 'kvx://local?prefill=1&decode=1&isl=8192&requests=64&concurrency=4&layout=nhd'
  run cfg3_1p1d_isl1k_c4 $L8 'kvx://local?prefill=1&decode=1&isl=1024&requests=128&concurrency=4&layout=nhd'
fi
if [ "$NGPU" -ge 6 ]; then
  run cfg4_tp4_tp2_hnd $L70 'kvx://local?prefill=1&decode=1&ptp=4&dtp=2&isl=8192&requests=16&concurrency=1'
  run cfg4_tp4_tp2_hnd_c2 $L70 'kvx://local?prefill=1&decode=1&ptp=4&dtp=2&isl=8192&requests=16&concurrency=2'
  run cfg4_tp2_tp4_hnd $L70 'kvx://local?prefill=1&decode=1&ptp=2&dtp=4&isl=8192&requests=16&concurrency=1'
  run cfg4_tp4_tp2_nhd $L70 'kvx://local?prefill=1&decode=1&ptp=4&dtp=2&isl=8192&requests=16&concurrency=1&layout=nhd'
fi
if [ "$NGPU" -ge 8 ]; then
  run cfg5_4p4d_isl8k $L8 'kvx://local?prefill=4&decode=4&isl=8192&requests=32&concurrency=1&layout=nhd'
  run cfg5_4p4d_mixed $L8 'kvx://local?prefill=4&decode=4&isl=1024,2048,4096,8192,16384,32768&requests=32&concurrency=1&layout=nhd'
fi
if [ "$NGPU" -ge 5 ]; then
  run incast_4p_to_1d $L8 'kvx://local?prefill=4&decode=1&isl=8192&requests=16&concurrency=4&route=spread'
  run fanout_1p_to_4d $L8 'kvx://local?prefill=1&decode=4&isl=8192&requests=32&concurrency=1'
fi
cat "$OUT/summaries.jsonl"
