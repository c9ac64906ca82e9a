#!/bin/bash
# Round-2 evidence on an 8-GPU box: N=8 bench line, router policies on a bursty 4P4D mixed-ISL stream, the 4P->1D incast,
# layer-wise handoff behind a synthetic prefill.  tools/run_round2.sh <out-dir>
OUT=${1:-gpurun_out/r02_results}
M8=meta-llama/Llama-3.1-8B
mkdir -p $OUT
export KVX_BENCH_TIMEOUT=200
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29540 bench.py --gpus 8 > $OUT/bench_n8.json 2> $OUT/bench_n8.err
echo "bench8 rc=$?"
U='kvx://local?prefill=4&decode=4&isl=1024,2048,4096,8192,16384,32768&requests=64&arrival=bursty&gap_ms=1.5&burst=6&seed=5'
for spec in "pair&concurrency=4" "spread&concurrency=4" "pair&concurrency=1" "sched&concurrency=1"; do
  name=route_$(echo $spec | sed 's/&concurrency=/_c/')
  timeout 240 ./run-benchmarks.sh -u "$U&route=$spec" -m $M8 -o $OUT -b $name > $OUT/$name.summary.json 2> $OUT/$name.err; echo "$name rc=$?"
done
I='kvx://local?prefill=4&decode=1&isl=8192&requests=32&arrival=bursty&gap_ms=2.0&burst=4&seed=7'
for spec in "spread&concurrency=4" "sched&concurrency=1"; do
  name=incast_4p1d_$(echo $spec | sed 's/&concurrency=/_c/')
  timeout 240 ./run-benchmarks.sh -u "$I&route=$spec" -m $M8 -o $OUT -b $name > $OUT/$name.summary.json 2> $OUT/$name.err; echo "$name rc=$?"
done
for g in 1 4 8 32; do
  name=layerwise_isl8k_g$g
  timeout 200 ./run-benchmarks.sh -u "kvx://local?prefill=1&decode=1&isl=8192&requests=8&layerwise=$g&prefill_ms=20" -m $M8 -o $OUT -b $name > $OUT/$name.summary.json 2> $OUT/$name.err; echo "$name rc=$?"
done
