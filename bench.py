#!/usr/bin/env python
"""bench.py — prefill→decode paged-KV handoff throughput (BASELINE.json's metric) on N GPUs of one node.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

A step = one request's handoff (all layers, K and V) from a prefill pool to a decode pool.
  N = 1   config #2: Qwen3-0.6B KV shape, ISL 8192 (896 MiB payload), both pools on the one GPU
          (intra-device gather→transform→scatter, fused kernel), HND → NHD in flight.
  N = 2,4,8  configs #3/#5: Llama-3.1-8B KV shape, ISL 8192 (1 GiB payload), N/2 independent
          P→D pairs; rank i (< N/2) is prefill worker i, rank N/2+i its decode worker, which PULLS
          the blocks over NVLink (--mode push: the prefill GPU stores them instead).  No collective
          on the data path ("scaling": "weak").
value = payload GB/s of the whole job (all pairs), inputs resident in HBM; e2e = the same through
kvx_handoff_host with host block-id lists (the ids' trip to the GPU and the completion word's trip
back to pinned host memory inside the timed region).  --impl reference times the CPU memcpy handoff (oracle/kvx_oracle.c) instead.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "dynamo-k8s-llm-inference_b200")
for _p in (ROOT, PKG):
    if _p not in sys.path:
        sys.path.insert(0, _p)

# BASELINE.json's metric is per P->D pair; the bench contract wants the whole job in `value`, so the line carries the
# sum over the concurrent pairs under this name and the per-pair figure next to it (`per_pair_GBps`).
METRIC = "kv_transfer_GBps_sum_over_PD_pairs"
UNIT = "GB/s"


def load_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs, burst copy)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def ncu_traffic():
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the fused kernel, from the committed
    ncu --set full capture of this same workload (profiles/); None when no capture is committed."""
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_fused_latest.json")) as f:
            d = json.load(f)
        tot = []
        for l in d["launches"]:
            v = 0.0
            for k in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
                num, unit = l[k].split()
                v += float(num) * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[unit]
            tot.append(v)
        return round(sum(tot) / len(tot)), d.get("report")
    except Exception:
        return None, None


NVLINK_PEAK = 770.0  # measured peer copy per direction on this pool (B200_PROFILING.md); nominal 900


class ClockSampler:
    """nvidia-smi clocks and throttle reasons DURING the timed region."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "50"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except ValueError:
                continue
            for name, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# --------------------------------------------------------------------------------------------------
# CPU arm (reference's own kind of handoff: memcpy per row on host cores)
# --------------------------------------------------------------------------------------------------
def cpu_handoff_rate(shape, isl, min_seconds, max_reps, threads):
    """GB/s of payload for the C restatement moving one request HND→NHD between host pools."""
    import numpy as np

    from oracle import kvx_oracle as o
    from oracle import kvx_oracle_c as oc

    n = shape.blocks_for(isl)
    nb = n + 1
    src = o.make_pool(shape.layers, shape.kv_heads, shape.head_dim, nb, o.HND, seed=1)
    dst = o.make_pool(shape.layers, shape.kv_heads, shape.head_dim, nb, o.NHD, fill="canary")
    ts, td = o.make_table(nb, n, 7), o.make_table(nb, n, 13)
    payload = n * shape.regions * shape.chunk_bytes()
    oc.handoff(src, ts, dst, td, threads=threads)  # warm (page faults)
    reps, t0 = 0, time.perf_counter()
    times = []
    while reps < max_reps:
        t1 = time.perf_counter()
        rc = oc.handoff(src, ts, dst, td, threads=threads)
        times.append(time.perf_counter() - t1)
        assert rc == 0
        reps += 1
        if time.perf_counter() - t0 > min_seconds:
            break
    return payload / (sum(times) / len(times)) / 1e9, reps, payload, times


def run_reference(args):
    from kvx.shapes import KV_SHAPES

    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    shape = KV_SHAPES["Qwen/Qwen3-0.6B"] if args.gpus == 1 else KV_SHAPES["meta-llama/Llama-3.1-8B"]
    isl = args.isl
    threads = len(os.sched_getaffinity(0))
    # each step = one request's handoff on host memory (bounded sample: `steps` requests)
    gbps, reps, payload, times = cpu_handoff_rate(shape, isl, min_seconds=1e9, max_reps=args.warmup + args.steps, threads=threads)
    timed = times[args.warmup:] or times
    ms = 1e3 * sum(timed) / len(timed)
    val = payload / (ms / 1e3) / 1e9
    line = {
        "impl": "reference", "metric": METRIC, "value": round(val, 3), "unit": UNIT, "n_gpus": args.gpus,
        "steps": len(timed), "warmup": args.warmup, "ms_per_step": round(ms, 3), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u8 (opaque bf16 bits)", "data": "synthetic",
        "config": workload_config(args.gpus, shape, isl),
        "cpu_baseline": {"value": round(val, 3), "unit": UNIT, "cores": threads, "kind": "port",
                         "sample": f"{len(timed)} requests of ISL {isl} ({payload / 2**20:.0f} MiB each), host pools, "
                                   "oracle/kvx_oracle.c memcpy-per-row HND->NHD"},
        "e2e": {"value": round(val, 3), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "the reference repo has no code for this path (bash+YAML); this arm is the CPU memcpy handoff "
                "BASELINE.md section 4 names, on all host threads",
    }
    print(json.dumps(line))
    return 0


def workload_config(n_gpus, shape, isl):
    if n_gpus == 1:
        wl = f"config#2 {shape.name} KV shape, single-GPU intra-device handoff HND->NHD, ISL {isl}"
    else:
        wl = (f"{shape.name} KV shape, {n_gpus // 2}P{n_gpus // 2}D over NVLink, HND->NHD, fixed ISL {isl} "
              f"({'config#3 topology' if n_gpus == 2 else 'config#5 topology; its mixed-ISL stream is the config5_mixed_isl block'})")
    return {"workload": wl, "layers": shape.layers, "kv_heads": shape.kv_heads, "head_dim": shape.head_dim,
            "block_tokens": shape.block_tokens, "isl": isl, "pairs": max(1, n_gpus // 2),
            "l2": "each request's payload (>= 896 MiB) exceeds the 126 MB L2 and steps rotate over 4 disjoint block sets"}


# --------------------------------------------------------------------------------------------------
# GPU arm
# --------------------------------------------------------------------------------------------------
def timed_loop(torch, stream, fn, steps):
    start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    start.record(stream)
    for k in range(steps):
        fn(k)
    end.record(stream)
    end.synchronize()
    return start.elapsed_time(end)  # ms


def library_baselines(torch, np, ctx, shape, P, D, perm_p, perm_d, n, nb, payload, stream, threads):
    """Baselines timed in the same run: (i) what the reference's images do today
    when layouts differ — land the raw blocks, then torch index_select -> permute -> index_copy_ per
    layer tensor (vLLM kv_connector/utils.py:255-277); (ii) a CPU-staged handoff: gather -> D2H into
    pinned memory -> host memcpy-per-row transform -> H2D -> scatter (the kv_buffer_device=cpu shape)."""
    from oracle import kvx_oracle as o
    from oracle import kvx_oracle_c as oc
    import ctypes

    N, H, Dh = shape.block_tokens, shape.kv_heads, shape.head_dim
    words = N * H * Dh
    src = P.tensor().view(torch.int16).view(shape.regions, nb, words)
    dst = D.tensor().view(torch.int16).view(shape.regions, nb, words)
    si = torch.from_numpy(perm_p[:n].astype(np.int64)).cuda()
    di = torch.from_numpy(perm_d[:n].astype(np.int64)).cuda()

    def torch_path(_k):
        for r in range(shape.regions):
            cache = dst[r].view(nb, N, H, Dh)
            cache.view(nb, -1).index_copy_(0, di, src[r].index_select(0, si))       # raw landing of HND bytes
            blocks = cache.index_select(0, di)                                      # receive-side permute
            cache.index_copy_(0, di, blocks.reshape(n, H, N, Dh).permute(0, 2, 1, 3))

    for k in range(2):
        torch_path(k)
    torch.cuda.synchronize()
    t_torch = timed_loop(torch, stream, torch_path, 3) / 3

    # CPU-staged: one request, three repetitions
    stage_dev = torch.empty(shape.regions, n, words, dtype=torch.int16, device="cuda")
    host_in = torch.empty(shape.regions, n, words, dtype=torch.int16).pin_memory()
    host_out = torch.empty(shape.regions, n, words, dtype=torch.int16).pin_memory()
    lib = oc.lib()
    chunk = words * 2

    def side(t, layout):
        return oc._Side(None, None, t.data_ptr(), 0, 0, layout, H, 0, 0)

    times = []
    for _ in range(3):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        torch.index_select(src, 1, si, out=stage_dev)
        host_in.copy_(stage_dev, non_blocking=True)
        torch.cuda.synchronize()
        s_side, d_side = side(host_in, o.HND), side(host_out, o.NHD)
        rc = lib.kvxo_xfer(ctypes.byref(s_side), ctypes.byref(d_side), shape.regions, n, N, H, Dh * 2, threads)
        assert rc == 0
        stage_dev.copy_(host_out, non_blocking=True)
        dst.index_copy_(1, di, stage_dev)
        torch.cuda.synchronize()
        times.append(time.perf_counter() - t0)
    t_cpu = min(times)
    del stage_dev
    return {
        "torch_land_then_index_select_permute_index_copy": {"ms": round(t_torch, 3), "payload_GBps": round(payload / t_torch / 1e6, 1),
                                                            "launches_per_request": 5 * shape.regions},
        "cpu_staged_d2h_hostmemcpy_h2d": {"ms": round(1e3 * t_cpu, 2), "payload_GBps": round(payload / t_cpu / 1e9, 2),
                                          "host_threads": threads, "sample": "1 request x 3 repetitions, best"},
    }


def run_single(args):
    import numpy as np
    import torch

    import kvx
    from kvx import synth

    hbm_peak, peak_src = load_peaks()
    torch.cuda.set_device(0)
    shape = kvx.KV_SHAPES["Qwen/Qwen3-0.6B"]
    isl = args.isl
    n = shape.blocks_for(isl)
    sets = 4
    nb = sets * n + 1
    ctx = kvx.Context(0)
    P = synth.alloc_pool(ctx, shape, nb, kvx.HND)
    D = synth.alloc_pool(ctx, shape, nb, kvx.NHD)
    synth.fill_random(P, synth.SEED_BASE + 2000)
    synth.fill_canary(D)
    perm_p = synth.make_table(nb, sets * n, synth.SEED_BASE + 2007)
    perm_d = synth.make_table(nb, sets * n, synth.SEED_BASE + 2013)
    h_src = [np.ascontiguousarray(perm_p[s * n:(s + 1) * n]) for s in range(sets)]
    h_dst = [np.ascontiguousarray(perm_d[s * n:(s + 1) * n]) for s in range(sets)]
    d_src = [torch.from_numpy(t).cuda() for t in h_src]
    d_dst = [torch.from_numpy(t).cuda() for t in h_dst]
    pin_src = [torch.from_numpy(t).pin_memory() for t in h_src]
    pin_dst = [torch.from_numpy(t).pin_memory() for t in h_dst]
    payload = n * shape.regions * shape.chunk_bytes()
    stream = torch.cuda.current_stream()
    flags, _ = ctx.flags()

    def step(k):
        s = k % sets
        ctx.handoff_fused(P.pool, d_src[s], D.pool, d_dst[s], n, None, flags, k + 1, stream)

    for k in range(args.warmup):
        step(k)
    torch.cuda.synchronize()
    launches0 = ctx.launch_count
    sampler = ClockSampler(0)
    sampler.start()
    time.sleep(0.15)
    torch.cuda.synchronize()
    ms = timed_loop(torch, stream, step, args.steps)
    torch.cuda.synchronize()
    launches = ctx.launch_count - launches0

    # e2e: host block-id lists in, completion word out, every step
    done = torch.zeros(args.steps + args.warmup, dtype=torch.int64).pin_memory()

    def e2e_step(k):
        s = k % sets
        ctx.handoff_host(P.pool, pin_src[s], D.pool, pin_dst[s], n, None, flags, 1000 + k, done[k:k + 1], stream)

    for k in range(args.warmup):
        e2e_step(k)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for k in range(args.steps):
        e2e_step(args.warmup + k)
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    assert int(done[args.warmup + args.steps - 1]) == 1000 + args.warmup + args.steps - 1, "completion word missing"
    clocks = sampler.stop()
    ctx.check()

    if args.no_extras:
        ms_step = ms / args.steps
        print(json.dumps({"metric": METRIC, "value": round(payload / (ms_step / 1e3) / 1e9, 2), "unit": UNIT, "n_gpus": 1,
                          "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(ms_step, 4),
                          "gpu_launches": int(launches), "clocks": clocks, "note": "--no-extras profiling run"}))
        P.free(); D.free(); ctx.close()
        return 0

    # the staged variant's kernels, each against the HBM roofline (2B bytes per launch)
    stage = torch.empty(payload, dtype=torch.uint8, device="cuda")
    stage2 = torch.empty(payload, dtype=torch.uint8, device="cuda")
    xf = kvx.Xform(0, 0, 0, kvx.NHD)
    D2 = synth.alloc_pool(ctx, shape, nb, kvx.HND)   # identity control: HND prefill pool -> HND decode pool
    asc = [torch.arange(1 + j * n, 1 + (j + 1) * n, dtype=torch.int32, device="cuda") for j in range(sets)]
    parts = {}
    reps = max(3, min(args.steps, 10))
    for name, fn in (
        ("gather_hnd_to_nhd", lambda k: ctx.gather(P.pool, d_src[k % sets], n, stage, xf, stream)),
        ("copy_local", lambda k: ctx.copy(stage, stage2, payload, None, 0, stream)),
        ("scatter_nhd", lambda k: ctx.scatter(stage2, D.pool, d_dst[k % sets], n, xf, stream)),
        ("fused_identity_hnd", lambda k: ctx.handoff_fused(P.pool, d_src[k % sets], D2.pool, d_dst[k % sets], n, None, None, 0, stream)),
        # best-case control: contiguous ascending block ids on both sides
        ("fused_hnd_to_nhd_contiguous_tables", lambda k: ctx.handoff_fused(P.pool, asc[k % sets], D.pool, asc[(k + 1) % sets], n, None, None, 0, stream)),
    ):
        for k in range(3):
            fn(k)
        torch.cuda.synchronize()
        t = timed_loop(torch, stream, fn, reps) / reps
        gbs = 2 * payload / (t / 1e3) / 1e9
        parts[name] = {"ms": round(t, 4), "hbm_GBps": round(gbs, 1), "frac_measured": round(gbs / hbm_peak, 4),
                       "frac_nominal_8TBps": round(gbs / 8000.0, 4)}
    # torch's plain copy of the same bytes in the same run (the roofline's own yardstick)
    for _ in range(3):
        stage2.copy_(stage)
    torch.cuda.synchronize()
    t = timed_loop(torch, stream, lambda k: stage2.copy_(stage), reps) / reps
    parts["torch_copy_same_bytes"] = {"ms": round(t, 4), "hbm_GBps": round(2 * payload / (t / 1e3) / 1e9, 1)}
    del stage, stage2
    D2.free()

    # ISL sweep of config #2 (fused, HND→NHD)
    sweep = {}
    for s_isl in (1024, 2048, 4096, 8192, 16384, 32768):
        m = shape.blocks_for(s_isl)
        if m > sets * n:
            continue
        tabs = [(torch.from_numpy(np.ascontiguousarray(perm_p[j * m:(j + 1) * m])).cuda(),
                 torch.from_numpy(np.ascontiguousarray(perm_d[j * m:(j + 1) * m])).cuda()) for j in range(max(1, (sets * n) // m))][:8]
        f = lambda k: ctx.handoff_fused(P.pool, tabs[k % len(tabs)][0], D.pool, tabs[k % len(tabs)][1], m, None, None, 0, stream)
        for k in range(3):
            f(k)
        torch.cuda.synchronize()
        r = 10
        t = timed_loop(torch, stream, f, r) / r
        b = m * shape.regions * shape.chunk_bytes()
        sweep[str(s_isl)] = {"ms": round(t, 4), "payload_GBps": round(b / (t / 1e3) / 1e9, 1),
                             "hbm_frac_measured": round(2 * b / (t / 1e3) / 1e9 / hbm_peak, 4)}

    # the host-table entry at the short end (launch-bound): per request, and 16 requests in one batch
    short = {}
    for s_isl in (128, 1024):
        m = shape.blocks_for(s_isl)
        R = 16
        ids = [(np.ascontiguousarray(perm_p[j * m:(j + 1) * m]), np.ascontiguousarray(perm_d[j * m:(j + 1) * m])) for j in range(R)]
        words = torch.zeros(R, dtype=torch.int64).pin_memory()
        reqs = [(a, b, None, 1, words[j:j + 1]) for j, (a, b) in enumerate(ids)]

        def one_by_one(_k):
            for j, (a, b) in enumerate(ids):
                ctx.handoff_host(P.pool, a, D.pool, b, m, None, None, 1, words[j:j + 1], stream)

        def batched(_k):
            ctx.handoff_batch(P.pool, D.pool, reqs, None, stream)

        rec = {"requests": R, "payload_MiB": round(R * m * shape.regions * shape.chunk_bytes() / 2**20, 1)}
        for name, fn in (("kvx_handoff_host_each_us", one_by_one), ("kvx_handoff_batch_us", batched)):
            for k in range(3):
                fn(k)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for k in range(10):
                fn(k)
            torch.cuda.synchronize()
            rec[name] = round(1e5 * (time.perf_counter() - t0), 1)
        short[str(s_isl)] = rec

    # the same requests through the resident receiver (kvx_queue_*): no launch per request, the CTAs run from one
    # request into the next.  Host wall clock from the first submit to the last request's completion word, per request
    # ("stream": requests submitted back to back) and submit -> completion of ONE request at a time ("single"); next to
    # them the per-request entry point measured the same way.
    queue = {}
    torch.cuda.synchronize()
    with ctx.queue(P.pool, D.pool) as q:
        for s_isl in (128, 1024, 2048, 4096, 8192):
            m = shape.blocks_for(s_isl)
            R = 16
            ids = [(np.ascontiguousarray(perm_p[(j * m) % (sets * n - m):][:m]), np.ascontiguousarray(perm_d[(j * m) % (sets * n - m):][:m])) for j in range(R)]
            word = torch.zeros(1, dtype=torch.int64).pin_memory()
            b = m * shape.regions * shape.chunk_bytes()

            def stream_q():
                for a_, b_ in ids:
                    t_ = q.submit(a_, b_, m)
                q.wait(t_, 10000)

            packed = q.pack([(a_, b_, None, 1, None) for a_, b_ in ids])

            def step_q():   # the 16 requests as one scheduler step: one call into the library, 16 separate completions
                q.wait(q.submit_many(packed), 10000)

            def single_q():
                for a_, b_ in ids:
                    q.wait(q.submit(a_, b_, m), 10000)

            def stream_l():
                for a_, b_ in ids:
                    ctx.handoff_host(P.pool, a_, D.pool, b_, m, None, None, 1, None, stream)
                stream.synchronize()

            def single_l():
                for a_, b_ in ids:
                    ctx.handoff_host(P.pool, a_, D.pool, b_, m, None, None, 1, word, stream)
                    stream.synchronize()

            rec = {"requests": R, "payload_MiB_each": round(b / 2**20, 1)}
            for name, fn in (("queue_stream_us", stream_q), ("queue_step_us", step_q), ("queue_single_us", single_q), ("launch_stream_us", stream_l), ("launch_single_us", single_l)):
                if name == "launch_stream_us":
                    q.park()   # the launched kernels get the GPU to themselves (the idle receiver holds registers on every SM)
                fn(); fn()
                best = 1e9
                for _ in range(5):
                    t0 = time.perf_counter()
                    fn()
                    best = min(best, time.perf_counter() - t0)
                rec[name] = round(1e6 * best / R, 2)
            rec["queue_stream_hbm_frac_measured"] = round(2 * b / (rec["queue_stream_us"] * 1e-6) / 1e9 / hbm_peak, 4)
            rec["queue_step_hbm_frac_measured"] = round(2 * b / (rec["queue_step_us"] * 1e-6) / 1e9 / hbm_peak, 4)
            rec["launch_stream_hbm_frac_measured"] = round(2 * b / (rec["launch_stream_us"] * 1e-6) / 1e9 / hbm_peak, 4)
            queue[str(s_isl)] = rec
        queue["ctas"], queue["kernel_launches_for_all_of_the_above"] = q.ctas, q.launches
        for k_isl, rec_ in queue.items():   # next to the launched path in the ISL sweep: the same request through the receiver
            if isinstance(rec_, dict) and k_isl in sweep:
                sweep[k_isl]["through_resident_receiver"] = {"us_per_request_in_a_stream": rec_["queue_stream_us"],
                                                             "hbm_frac_measured": rec_["queue_stream_hbm_frac_measured"]}
        q.park()
    torch.cuda.synchronize()

    # chunks of TP shards (Llama-3.1-70B, 80 layers, ISL 8192): 4, 2 and 1 KV heads per rank = 16, 8 and 4 KiB
    # chunks; small pieces switch the launch geometry (few consumer warps, more CTAs)
    shards = {}
    s70 = kvx.KV_SHAPES["meta-llama/Llama-3.1-70B"]
    for heads in (4, 2, 1):
        m = s70.blocks_for(8192)
        nbs = 2 * m + 1
        Ps = synth.alloc_pool(ctx, s70, nbs, kvx.HND, heads=heads)
        Ds = synth.alloc_pool(ctx, s70, nbs, kvx.NHD, heads=heads)
        tp_, td_ = synth.make_table(nbs, 2 * m, 31), synth.make_table(nbs, 2 * m, 32)
        tabs = [(torch.from_numpy(np.ascontiguousarray(tp_[j * m:(j + 1) * m])).cuda(),
                 torch.from_numpy(np.ascontiguousarray(td_[j * m:(j + 1) * m])).cuda()) for j in range(2)]
        f = lambda k: ctx.handoff_fused(Ps.pool, tabs[k % 2][0], Ds.pool, tabs[k % 2][1], m, None, None, 0, stream)
        for k in range(3):
            f(k)
        torch.cuda.synchronize()
        t = timed_loop(torch, stream, f, 10) / 10
        b = m * s70.regions * s70.block_tokens * heads * s70.head_dim * s70.elem_bytes
        shards[f"{heads}_heads_{b // (m * s70.regions) // 1024}KiB_chunks"] = {
            "ms": round(t, 4), "hbm_GBps": round(2 * b / (t / 1e3) / 1e9, 1), "frac_measured": round(2 * b / (t / 1e3) / 1e9 / hbm_peak, 4)}
        Ps.free(); Ds.free()

    threads = len(os.sched_getaffinity(0))
    baselines = library_baselines(torch, np, ctx, shape, P, D, perm_p, perm_d, n, nb, payload, stream, threads)
    cpu_gbps, cpu_reps, _, _ = cpu_handoff_rate(shape, isl, min_seconds=10.0, max_reps=200, threads=threads)

    ms_step = ms / args.steps
    value = payload / (ms_step / 1e3) / 1e9
    achieved = 2 * payload / (ms_step / 1e3) / 1e9
    traffic, traffic_src = ncu_traffic() if isl == 8192 else (None, None)
    line = {
        "metric": METRIC, "value": round(value, 2), "unit": UNIT, "n_gpus": 1, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": round(ms_step, 4), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u8 (opaque bf16 bits)", "data": "synthetic", "config": workload_config(1, shape, isl),
        "e2e": {"value": round(args.steps * payload / e2e_s / 1e9, 2), "unit": UNIT, "h2d_bytes_per_step": 8 * n,
                "d2h_bytes_per_step": 8, "ms_per_step": round(1e3 * e2e_s / args.steps, 4),
                "through": "kvx_handoff_host (C ABI): host int32 block-id lists in (carried in the kernel's parameter "
                           "block), completion word out (written by the kernel into pinned host memory); the KV "
                           "payload is produced by prefill in HBM and never crosses PCIe on this path"},
        "gpu_launches": int(launches),
        "clocks": clocks,
        "roofline": {"bound": "hbm", "achieved": round(achieved, 1), "peak": hbm_peak, "unit": "GB/s",
                     "frac": round(achieved / hbm_peak, 4), "traffic": traffic, "traffic_source": traffic_src,
                     "peak_source": peak_src, "kernel": "kvx_xfer_tma_kernel (fused gather+transform+scatter)",
                     "algorithmic_bytes_per_launch": 2 * payload, "frac_of_nominal_8TBps": round(achieved / 8000.0, 4)},
        "cpu_baseline": {"value": round(cpu_gbps, 3), "unit": UNIT, "cores": threads, "kind": "port",
                         "sample": f"{cpu_reps} requests of ISL {isl} ({payload / 2**20:.0f} MiB each) between host pools, "
                                   "oracle/kvx_oracle.c memcpy-per-row HND->NHD"},
        "added_ttft_ms_isl8k": round(ms_step, 4) if isl == 8192 else None,
        # what the fused kernel replaces, timed in this run on the same request: the post-process the reference's images run
        # on the decode GPU today (land raw blocks, index_select -> permute -> index_copy_ per layer tensor), the CPU-staged
        # handoff, and the pure-host handoff.  The e2e / cpu_baseline ratio compares HBM-resident data with host-resident
        # data; the first of these three is the like-for-like one.
        "same_run_speedups": {"vs_torch_land_then_postprocess": round(baselines["torch_land_then_index_select_permute_index_copy"]["ms"] / ms_step, 1),
                              "vs_cpu_staged_gpu_handoff": round(baselines["cpu_staged_d2h_hostmemcpy_h2d"]["ms"] / ms_step, 1),
                              "vs_host_memcpy_handoff": round(value / cpu_gbps, 1)},
        "kernels": parts, "isl_sweep": sweep, "short_requests_host_entry": short, "resident_queue": queue, "tp_shard_chunks": shards, "baselines": baselines,
        "options": {k: ctx.get_option(k) for k in kvx.OPT},
    }
    print(json.dumps(line))
    P.free(); D.free(); ctx.close()
    return 0


def pull_ncu_traffic():
    """Local DRAM bytes of one 1 GiB pull from the committed ncu capture (the payload arrives over NVLink and is
    written once to local HBM; nvlrx/nvltx counters read 0 on this driver)."""
    try:
        with open(os.path.join(ROOT, "profiles", "r01s2_ncu_pull_summary.json")) as f:
            d = json.load(f)
        return int(d["launches"][0]["dram_traffic_bytes"]), d.get("report")
    except Exception:
        return None, None


def harness_leg(dist, rank, world, topo, cfg, n_devices=None):
    """One verified request stream through kvx.harness.run_rank on the ranks torchrun started (the same code
    `run-benchmarks.sh -u kvx://...` runs in processes of its own).  Ranks the topology does not need only take part
    in the collectives."""
    from kvx import harness

    plans = harness.rank_plans(topo, n_devices or world)
    res = harness.run_rank(rank, plans, cfg, world_size=world)
    summary = None
    if rank == 0:
        summary = harness.summarize(res["all"], topo.describe(), cfg, len(plans), n_devices or world)
    dist.barrier()
    return summary


def brief(summary, extra=()):
    """The part of a harness summary that belongs in the bench line."""
    if summary is None:
        return None
    keys = ("requests", "concurrency", "route", "aggregate_throughput_GBps", "per_request_throughput_GBps_median",
            "per_edge_throughput_GBps_median", "per_decode_rank_ingress_GBps", "pct_of_nvlink_900", "per_isl",
            "time_to_first_token_added_ms", "xfer_time_seconds", "verified_bit_exact", "requests_verified",
            "blocks_checked", "canary_blocks_checked", "verification_failures", "prefill_blocks_released", "releases_observed", "wall_seconds") + tuple(extra)
    out = {k: summary[k] for k in keys if k in summary}
    for k, v in list(out.items()):
        if isinstance(v, float):
            out[k] = round(v, 3)
    out["topology"] = summary["topology"]["name"]
    return out


def phase(name):
    """Progress marker on stderr (every rank): which part of a multi-GPU run a hang or a failure belongs to."""
    sys.stderr.write(f"[bench rank {os.environ.get('RANK', '0')} +{time.perf_counter() - _T0:7.2f}s] {name}\n")
    sys.stderr.flush()


_T0 = time.perf_counter()


def run_pairs(args):
    import faulthandler

    import numpy as np
    import torch
    import torch.distributed as dist

    # a multi-GPU run that stops making progress dumps every thread's stack and exits instead of hanging the box
    faulthandler.dump_traceback_later(int(os.environ.get("KVX_BENCH_TIMEOUT", "900")), exit=True)

    import kvx
    from kvx import harness, synth
    from kvx.manifest import synthetic_topology
    from kvx.workers import DecodeWorker, PrefillWorker

    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    assert world == args.gpus and world % 2 == 0, "pairs need an even number of ranks"
    one_gpu = args.one_gpu   # debugging aid: every rank on cuda:0 (a one-GPU box can walk through the N-rank logic)
    if one_gpu:
        local = 0
    torch.cuda.set_device(local)
    # stdout carries exactly one JSON line: everything native libraries print there (NCCL's version
    # banner) is sent to stderr by pointing fd 1 at fd 2 for the whole run; the line goes to the saved fd.
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)
    if one_gpu:
        dist.init_process_group("gloo", rank=rank, world_size=world)
    else:
        dist.init_process_group("cpu:gloo,cuda:nccl", rank=rank, world_size=world, device_id=torch.device(f"cuda:{local}"))

    def all_reduce(t, op):
        if one_gpu:   # gloo reduces host tensors
            h = t.cpu()
            dist.all_reduce(h, op=op)
            t.copy_(h)
        else:
            dist.all_reduce(t, op=op)

    hbm_peak, peak_src = load_peaks()
    shape = kvx.KV_SHAPES["meta-llama/Llama-3.1-8B"]
    isl = args.isl
    n = shape.blocks_for(isl)
    sets = 4
    nb = sets * n + 1
    pairs = world // 2
    is_p = rank < pairs
    peer = rank + pairs if is_p else rank - pairs
    payload = n * shape.regions * shape.chunk_bytes()
    stream = torch.cuda.current_stream()

    if is_p:
        w = PrefillWorker(local, shape, nb, kvx.HND, seed=rank)
        synth.fill_random(w.kv, synth.SEED_BASE + 3000 + rank)
    else:
        w = DecodeWorker(local, shape, nb, kvx.NHD, seed=rank)
        synth.fill_canary(w.kv)
    for slot in range(4):
        w.expect(slot)
    torch.cuda.synchronize()
    metas = [None] * world
    dist.all_gather_object(metas, w.metadata())
    w.add_remote(0, metas[peer])
    perm_p = synth.make_table(nb, sets * n, synth.SEED_BASE + 3007 + (rank if is_p else peer))
    perm_d = synth.make_table(nb, sets * n, synth.SEED_BASE + 3013 + (peer if is_p else rank))
    total_steps = args.warmup + args.steps
    h_src = [np.ascontiguousarray(perm_p[s * n:(s + 1) * n]) for s in range(sets)]
    h_dst = [np.ascontiguousarray(perm_d[s * n:(s + 1) * n]) for s in range(sets)]
    d_src = [torch.from_numpy(t).cuda() for t in h_src]
    d_dst = [torch.from_numpy(t).cuda() for t in h_dst]
    pin_src = [torch.from_numpy(t).pin_memory() for t in h_src]
    pin_dst = [torch.from_numpy(t).pin_memory() for t in h_dst]
    pull = args.mode == "pull"
    clocks = None
    mover = (not is_p) if pull else is_p  # the rank whose GPU runs the transfer kernel
    alone_steps = min(args.steps, 20)
    # block sets this rank's pool takes part in: warm-up + timed steps, the e2e steps, the one-pair run (pair 0), and —
    # with the extras — the staged path's five scatters and the CPU-staged request (set 0)
    used_sets = {k % sets for k in range(total_steps)} | {k % sets for k in range(args.steps)}
    if rank in (0, pairs):
        used_sets |= {k % sets for k in range(alone_steps)}
    if not args.no_extras:
        used_sets |= {r % sets for r in range(5)} | {0}

    def move(k, slot, value, host, h_done=None):
        s = k % sets
        src, dst = (pin_src[s], pin_dst[s]) if host else (d_src[s], d_dst[s])
        if pull:
            w.pull(0, src, dst, notify_slot=slot, value=value, host_tables=host, h_done=h_done, stream=stream)
        else:
            w.push(0, src, dst, slot=slot, value=value, host_tables=host, h_done=h_done, stream=stream)

    def step(k, slot=0, base=0, active=True):
        if not active:
            return
        if mover:
            move(k, slot, base + k + 1, False)
        else:  # the other side only learns about completion through its flag
            w.wait(slot, base + k + 1, 10000, stream)

    torch.cuda.synchronize()
    dist.barrier()
    phase("headline: warm-up + timed steps")
    for k in range(args.warmup):
        step(k)
    torch.cuda.synchronize()
    dist.barrier()
    l0 = w.ctx.launch_count
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        time.sleep(0.15)
    torch.cuda.synchronize()
    dist.barrier()
    ms = timed_loop(torch, stream, lambda k: step(args.warmup + k), args.steps)
    torch.cuda.synchronize()
    launches = w.ctx.launch_count - l0
    dist.barrier()
    w.ctx.check()

    # e2e through the host-table entry on the moving side; the other side waits on its flag
    phase("e2e steps")
    base = 100000
    done = torch.zeros(total_steps, dtype=torch.int64).pin_memory()
    torch.cuda.synchronize()
    dist.barrier()
    t0 = time.perf_counter()
    for k in range(args.steps):
        if mover:
            move(k, 1, base + k + 1, True, done[k:k + 1])
        else:
            w.wait(1, base + k + 1, 10000, stream)
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    dist.barrier()
    w.ctx.check()
    if rank == 0:
        clocks = sampler.stop()

    # Short requests (ISL 1k, 128 MiB): per-request cost over NVLink is the bytes plus ~20 us of launch and link round
    # trips.  Per request through the per-request entry (launch) and through the resident receiver (kvx_queue_*), one at
    # a time ("single": submit -> completion observed, then the next) and back to back ("stream"); host wall clock on
    # the decode rank, all pairs at once.  Same (source, destination) block pairs as the headline's sets, so the
    # whole-pool verification below covers what these moved.
    short = None
    phase("short requests (queue / launch)")
    if not args.no_extras and pull:
        m_short = shape.blocks_for(1024)
        R = 16
        rec = {}
        if not is_p:
            ids = [(np.ascontiguousarray(perm_p[j * m_short:(j + 1) * m_short]), np.ascontiguousarray(perm_d[j * m_short:(j + 1) * m_short])) for j in range(R)]
            word = torch.zeros(1, dtype=torch.int64).pin_memory()
            q = w.pull_queue(0)

            def q_single():
                for a_, b_ in ids:
                    q.wait(q.submit(a_, b_, m_short), 10000)

            def q_stream():
                for a_, b_ in ids:
                    t_ = q.submit(a_, b_, m_short)
                q.wait(t_, 10000)

            def l_single():
                for a_, b_ in ids:
                    w.pull(0, a_, b_, host_tables=True, h_done=word, stream=stream)
                    stream.synchronize()

            def l_stream():
                for a_, b_ in ids:
                    w.pull(0, a_, b_, host_tables=True, stream=stream)
                stream.synchronize()

            b_short = m_short * shape.regions * shape.chunk_bytes()
            for name, fn in (("queue_single", q_single), ("queue_stream", q_stream), ("launch_single", l_single), ("launch_stream", l_stream)):
                if name == "launch_single":
                    q.park()
                fn(); fn()
                best = 1e9
                for _ in range(5):
                    t0 = time.perf_counter()
                    fn()
                    best = min(best, time.perf_counter() - t0)
                rec[name + "_us"] = round(1e6 * best / R, 2)
                rec[name + "_GBps"] = round(b_short / (best / R) / 1e9, 1)
            rec["queue_ctas"] = q.ctas
            q.park()
            torch.cuda.synchronize()
        # the prefill ranks wait here on the CPU (gloo): an NCCL collective would park a spinning kernel on the very GPUs
        # whose memory is being read
        dist.all_reduce(torch.zeros(1))
        allr = [None] * world
        dist.all_gather_object(allr, rec)
        if rank == 0:
            dec = [r for r in allr if r]
            short = {"isl": 1024, "payload_MiB_each": round(m_short * shape.regions * shape.chunk_bytes() / 2**20, 1), "requests": R,
                     "per_request_min_over_pairs": {k: (min if k.endswith("GBps") else max)(r[k] for r in dec) for k in dec[0]},
                     "note": "worst pair; queue_single = one request at a time through the resident receiver (concurrency 1)"}
        dist.barrier()

    # pair 0 alone (the other pairs idle): what the concurrent pairs cost each other
    active = rank in (0, pairs)
    phase("pair 0 alone")
    torch.cuda.synchronize()
    dist.barrier()
    ms_alone = timed_loop(torch, stream, lambda k: step(k, slot=3, base=0, active=active), alone_steps) if active else 0.0
    torch.cuda.synchronize()
    dist.barrier()

    # the unfused path of the same request: gather (P GPU) -> copy over NVLink (D GPU pulls the staging
    # image) -> scatter (D GPU), each kernel timed on its own GPU; next to it the library yardstick for the copy
    # (the copy engines moving the same image: cudaMemcpyAsync between the two GPUs' memories = cudaMemcpyPeerAsync)
    # and the CPU-staged handoff (D2H of the request's blocks into pinned host memory -> host
    # memcpy-per-row transform -> H2D -> scatter)
    staged, baselines = None, None
    phase("staged path + library / CPU-staged baselines")
    threads = len(os.sched_getaffinity(0))
    if not args.no_extras:
        reps = 5
        st_ptr = w.ctx.mem_alloc(payload)
        blobs = [None] * world
        dist.all_gather_object(blobs, w.ctx.mem_export(st_ptr, payload) if is_p else None)
        tiny = torch.zeros(64, dtype=torch.uint8, device="cuda")
        xf = kvx.Xform(0, 0, 0, kvx.NHD)
        xf_raw = kvx.Xform(0, 0, 0, kvx.HND)   # CPU-staged leg: the image leaves the prefill GPU untransformed
        torch.cuda.synchronize()
        dist.barrier()
        evs = [[torch.cuda.Event(enable_timing=True) for _ in range(6)] for _ in range(reps)]
        peer_stage = None if is_p else w.ctx.mem_import(blobs[peer])
        mine_t = synth.as_torch(st_ptr, payload, local)
        peer_t = None if is_p else synth.as_torch(peer_stage, payload, local)
        for r in range(reps):
            if is_p:
                evs[r][0].record(stream)
                w.ctx.gather(w.pool, d_src[r % sets], n, st_ptr, xf, stream)
                evs[r][1].record(stream)
                w.ctx.copy(tiny, tiny[32:], 16, w.remotes[0].flag_addr(2), r + 1, stream)   # "image ready" to D
                w.wait(2, r + 1, 10000, stream)                                             # D has read it
            else:
                w.wait(2, r + 1, 10000, stream)
                evs[r][0].record(stream)
                w.ctx.copy(peer_stage, st_ptr, payload, None, 0, stream)                    # NVLink read (kvx_copy)
                evs[r][1].record(stream)
                mine_t.copy_(peer_t)                                                        # same bytes, copy engines
                evs[r][4].record(stream)
                w.ctx.copy(tiny, tiny[32:], 16, w.remotes[0].flag_addr(2), r + 1, stream)   # P may reuse its image
                evs[r][2].record(stream)
                w.ctx.scatter(st_ptr, w.pool, d_dst[r % sets], n, xf, stream)
                evs[r][3].record(stream)
        torch.cuda.synchronize()
        dist.barrier()
        w.ctx.check()

        # CPU-staged handoff of one request (set 0), twice, best: P gathers the raw blocks into its image (timed
        # on P), D copies that image to pinned host memory (a D2H read of the PREFILL GPU's memory), transforms it
        # with the C memcpy-per-row loop on all host threads, uploads and scatters
        cpu_ms, g_raw_ms = None, None
        if is_p:
            g_raw = []
            for r in range(2):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(stream)
                w.ctx.gather(w.pool, d_src[0], n, st_ptr, xf_raw, stream)
                e1.record(stream)
                w.ctx.copy(tiny, tiny[32:], 16, w.remotes[0].flag_addr(2), 100 + r, stream)
                w.wait(2, 100 + r, 60000, stream)
                torch.cuda.synchronize()
                g_raw.append(e0.elapsed_time(e1))
            g_raw_ms = min(g_raw)
        else:
            import ctypes

            from oracle import kvx_oracle as o
            from oracle import kvx_oracle_c as oc

            words = shape.chunk_bytes() // 2
            host_in = torch.empty(shape.regions, n, words, dtype=torch.int16).pin_memory()
            host_out = torch.empty(shape.regions, n, words, dtype=torch.int16).pin_memory()
            lib = oc.lib()
            H, N, Dh = shape.kv_heads, shape.block_tokens, shape.head_dim
            side = lambda t, layout: oc._Side(None, None, t.data_ptr(), 0, 0, layout, H, 0, 0)  # noqa: E731
            t_cpu = []
            for r in range(2):
                w.wait(2, 100 + r, 60000, stream)
                torch.cuda.synchronize()
                t1 = time.perf_counter()
                host_in.view(-1).copy_(peer_t.view(torch.int16), non_blocking=True)
                torch.cuda.synchronize()
                s_side, d_side = side(host_in, o.HND), side(host_out, o.NHD)
                rc = lib.kvxo_xfer(ctypes.byref(s_side), ctypes.byref(d_side), shape.regions, n, N, H, Dh * 2, threads)
                assert rc == 0
                mine_t.view(torch.int16).copy_(host_out.view(-1), non_blocking=True)
                w.ctx.scatter(st_ptr, w.pool, d_dst[0], n, xf, stream)
                torch.cuda.synchronize()
                t_cpu.append(time.perf_counter() - t1)
                w.ctx.copy(tiny, tiny[32:], 16, w.remotes[0].flag_addr(2), 100 + r, stream)
                torch.cuda.synchronize()
            cpu_ms = 1e3 * min(t_cpu)
            del host_in, host_out
        dist.barrier()
        if is_p:
            mine = {"gather_ms": min(e[0].elapsed_time(e[1]) for e in evs[1:]), "gather_raw_ms": g_raw_ms}
        else:
            mine = {"copy_ms": min(e[0].elapsed_time(e[1]) for e in evs[1:]),
                    "memcpy_ms": min(e[1].elapsed_time(e[4]) for e in evs[1:]),
                    "scatter_ms": min(e[2].elapsed_time(e[3]) for e in evs[1:]), "cpu_staged_ms": cpu_ms}
            w.ctx.mem_unmap(peer_stage)
        allm = [None] * world
        dist.all_gather_object(allm, mine)
        if rank == 0:
            g = max(m["gather_ms"] for m in allm if "gather_ms" in m)
            c = max(m["copy_ms"] for m in allm if "copy_ms" in m)
            mc = max(m["memcpy_ms"] for m in allm if "memcpy_ms" in m)
            sc = max(m["scatter_ms"] for m in allm if "scatter_ms" in m)
            cs = max(m["cpu_staged_ms"] for m in allm if m.get("cpu_staged_ms")) + max(m["gather_raw_ms"] for m in allm if m.get("gather_raw_ms"))
            staged = {"gather_ms": round(g, 4), "gather_hbm_GBps": round(2 * payload / g / 1e6, 1),
                      "copy_ms": round(c, 4), "copy_nvlink_GBps": round(payload / c / 1e6, 1),
                      "copy_pct_of_nvlink_900": round(100 * payload / c / 1e6 / 900.0, 2),
                      "scatter_ms": round(sc, 4), "scatter_hbm_GBps": round(2 * payload / sc / 1e6, 1),
                      "note": "max over pairs of the best of 4 repetitions; copy = kvx_copy run by the decode GPU on the "
                              "prefill GPU's staging image"}
            baselines = {"cudaMemcpyPeerAsync_GBps": round(payload / mc / 1e6, 1), "cudaMemcpyPeerAsync_ms": round(mc, 4),
                         "cudaMemcpyPeerAsync_note": "the copy engines moving the same contiguous 1 GiB image between the same two GPUs "
                                                     "(torch copy_ = cudaMemcpyAsync over UVA, the cudaMemcpyPeerAsync path), all pairs at once, max over pairs",
                         "cpu_staged_GBps": round(payload / cs / 1e6, 2), "cpu_staged_ms": round(cs, 2), "cpu_staged_host_threads": threads,
                         "cpu_staged_note": "one request: gather on the prefill GPU (timed there) + D2H of that image into pinned host memory, "
                                            "memcpy-per-row HND->NHD on the host (oracle/kvx_oracle.c), H2D, scatter on the decode GPU; best of 2"}
        dist.barrier()
        w.ctx.mem_free(st_ptr)

    # Bit-exactness of EVERYTHING that landed: the decode rank regenerates its prefill peer's pool from the seed
    # (same generator, on this GPU — not through the mapping the transfer used), builds the expected decode pool
    # for every block set the run touched, and compares the ENTIRE pool: touched blocks, canary blocks, block 0.
    phase("whole-pool verification")
    vstat = torch.zeros(3, dtype=torch.int64, device="cuda")  # [ok, touched blocks, canary blocks]
    if not is_p:
        canary = np.array(synth.CANARY16, dtype=np.uint16).view(np.int16).item()
        words = shape.chunk_bytes() // 2
        dv = w.kv.tensor().view(torch.int16).view(shape.regions, nb, words)
        src = harness.regen_random(torch, shape.regions * nb * words * 2, synth.SEED_BASE + 3000 + peer, local).view(shape.regions, nb, words)
        exp = torch.full_like(dv, canary)
        H, N, Dh = shape.kv_heads, shape.block_tokens, shape.head_dim
        for s in sorted(used_sets):
            si = torch.from_numpy(h_src[s].astype(np.int64)).cuda()
            di = torch.from_numpy(h_dst[s].astype(np.int64)).cuda()
            exp[:, di] = src[:, si].view(shape.regions, n, H, N, Dh).transpose(2, 3).reshape(shape.regions, n, words)
        vstat[0] = int(torch.equal(exp, dv))
        vstat[1] = len(used_sets) * n
        vstat[2] = nb - len(used_sets) * n
        del src, exp
    ok = vstat[:1].clone() if not is_p else torch.ones(1, dtype=torch.int64, device="cuda")
    all_reduce(ok, dist.ReduceOp.MIN)
    vsum = vstat[1:].clone()
    all_reduce(vsum, dist.ReduceOp.SUM)
    vstat[1:] = vsum

    t = torch.tensor([ms, e2e_s * 1e3, ms_alone], device="cuda", dtype=torch.float64)
    all_reduce(t, dist.ReduceOp.MAX)
    lt = torch.tensor([float(launches)], device="cuda", dtype=torch.float64)
    all_reduce(lt, dist.ReduceOp.SUM)
    ms_max, e2e_ms, alone_ms = float(t[0]), float(t[1]), float(t[2])
    dist.barrier()
    w.close()
    torch.cuda.empty_cache()

    # CPU baseline of the same request (host pools, the C restatement on every host thread), rank 0, bounded
    cpu = None
    phase("cpu baseline (rank 0)")
    if rank == 0 and not args.no_extras:
        cpu_gbps, cpu_reps, _, _ = cpu_handoff_rate(shape, isl, min_seconds=6.0, max_reps=100, threads=threads)
        cpu = {"value": round(cpu_gbps, 3), "unit": UNIT, "cores": threads, "kind": "port",
               "sample": f"{cpu_reps} requests of ISL {isl} ({payload / 2**20:.0f} MiB each) between host pools, "
                         "oracle/kvx_oracle.c memcpy-per-row HND->NHD, the other ranks idle"}
    dist.barrier()

    # BASELINE.json configs #3, #4 and #5 as verified request streams (kvx.harness; not the headline)
    legs = {}
    if not args.no_extras:
        M8, M70 = "meta-llama/Llama-3.1-8B", "meta-llama/Llama-3.1-70B"
        ndev = 1 if one_gpu else world
        topo = synthetic_topology(pairs, pairs, 1, 1, "vllm", M8, decode_layout=kvx.NHD)
        want = [x for x in args.legs.split(",") if x]
        for c in ((1, 4) if ("all" in want or "config3" in want) else ()):  # config #3's load shape: 64-request closed loop at ISL 8192 per pair, concurrency 1 and 4
            phase(f"leg config3 c={c}")
            cfg = harness.StreamConfig(model=M8, isl=[8192] * args.stream_requests, concurrency=c)
            legs[f"config3_stream_isl8k_c{c}"] = brief(harness_leg(dist, rank, world, topo, cfg, n_devices=ndev))
        # config #5's: every pair its own stream of 32 requests, ISL log-uniform over 1k..32k, all pairs at once
        choices = [1024, 2048, 4096, 8192] if one_gpu else [1024, 2048, 4096, 8192, 16384, 32768]
        per_rep = [harness.draw_isls(args.mixed_requests, choices, 5 + 17 * d) for d in range(pairs)]
        if "all" in want or "config5" in want:
            phase("leg config5 mixed ISL")
            cfg = harness.StreamConfig(model=M8, isl=per_rep[0], per_replica_isl=per_rep, concurrency=1)
            legs["config5_mixed_isl"] = brief(harness_leg(dist, rank, world, topo, cfg, n_devices=ndev))
            phase("leg config5 mixed ISL through the resident receiver")
            cfg = harness.StreamConfig(model=M8, isl=per_rep[0], per_replica_isl=per_rep, concurrency=1, queue=True)
            legs["config5_mixed_isl_resident_receiver"] = brief(harness_leg(dist, rank, world, topo, cfg, n_devices=ndev), extra=("resident_receiver",))
        # config #4's: TP reshard on transfer, prefill TP = 2 x decode TP, every edge concurrent.  6 of 8 GPUs run
        # TP4 -> TP2 (Llama-3.1-70B); with 4 GPUs the same code runs TP2 -> TP1 on 3 of them.
        if world >= 4 and ("all" in want or "config4" in want):
            ptp, dtp = (4, 2) if world >= 6 else (2, 1)
            phase(f"leg config4 TP{ptp}->TP{dtp}")
            topo4 = synthetic_topology(1, 1, ptp, dtp, "vllm", M70)
            cfg = harness.StreamConfig(model=M70, isl=[8192] * 8, concurrency=1)
            s4 = harness_leg(dist, rank, world, topo4, cfg, n_devices=ndev)
            legs[f"config4_tp{ptp}_to_tp{dtp}"] = brief(s4)

    phase("done")
    faulthandler.cancel_dump_traceback_later()
    if rank == 0:
        ms_step = ms_max / args.steps
        per_pair = payload / (ms_step / 1e3) / 1e9
        alone = payload / (alone_ms / alone_steps / 1e3) / 1e9 if alone_ms > 0 else None
        value = pairs * per_pair
        peak = baselines["cudaMemcpyPeerAsync_GBps"] if baselines else NVLINK_PEAK
        traffic, traffic_src = pull_ncu_traffic() if (pull and isl == 8192) else (None, None)
        verified = bool(ok[0] > 0)
        line = {
            "metric": METRIC, "value": round(value, 2), "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": round(ms_step, 4), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u8 (opaque bf16 bits)", "data": "synthetic",
            "config": workload_config(world, shape, isl),
            "value_is": f"sum over the {pairs} concurrent P->D pair(s); one pair = per_pair_GBps",
            "per_pair_GBps": round(per_pair, 2), "pct_of_nvlink_900": round(100 * per_pair / 900.0, 2),
            "one_pair_alone_GBps": round(alone, 2) if alone else None,
            "pairs_efficiency_vs_one_pair_alone": round(per_pair / alone, 4) if alone else None,
            "e2e": {"value": round(pairs * args.steps * payload / (e2e_ms / 1e3) / 1e9, 2), "unit": UNIT,
                    "h2d_bytes_per_step": 8 * n, "d2h_bytes_per_step": 8, "ms_per_step": round(e2e_ms / args.steps, 4),
                    "through": f"{'DecodeWorker.pull' if pull else 'PrefillWorker.push'} -> kvx_handoff_host (host block-id "
                               "lists in, completion word back; release flag stored in the peer's memory)"},
            "gpu_launches": int(lt[0]), "clocks": clocks,
            "roofline": {"bound": "nvlink", "achieved": round(per_pair, 1), "peak": peak, "unit": "GB/s",
                         "frac": round(per_pair / peak, 4), "traffic": traffic, "traffic_source": traffic_src,
                         "peak_source": ("measured in this run: the copy engines moving the same 1 GiB between the same two GPUs "
                                         "(baselines.cudaMemcpyPeerAsync_GBps); nominal 900 per direction") if baselines else
                                        "770 GB/s per direction (B200_PROFILING.md peer copy; --no-extras run); nominal 900",
                         "kernel": "kvx_xfer_tma_kernel (fused: " + ("peer gather over NVLink + transform + local scatter, run by the decode GPU)" if pull else "local gather + transform + peer store, run by the prefill GPU)"),
                         "algorithmic_bytes_per_launch": payload, "frac_of_nominal_900": round(per_pair / 900.0, 4)},
            "cpu_baseline": cpu,
            "added_ttft_ms_isl8k": round(ms_step, 4) if isl == 8192 else None,
            "direction": args.mode,
            "verified_bit_exact": verified, "blocks_checked": int(vstat[1]), "canary_blocks_checked": int(vstat[2]),
            "verification": "every decode rank: ENTIRE pool (all regions of every block, block 0 included) against the expected pool "
                            "rebuilt from the prefill rank's seed, after the timed, e2e, one-pair and staged runs",
            "short_requests_isl1k": short, "staged_path": staged, "baselines": baselines, **legs,
        }
        os.write(json_fd, (json.dumps(line) + "\n").encode())
    os.close(json_fd)
    bad = int(ok[0] <= 0)
    leg_bad = [k for k, v in legs.items() if rank == 0 and v is not None and v.get("verified_bit_exact") is not True]
    dist.barrier()
    dist.destroy_process_group()
    if bad or leg_bad:
        sys.stderr.write(f"bench: verification FAILED (headline {'bad' if bad else 'ok'}, legs {leg_bad})\n")
        return 1
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="kvx", choices=["kvx", "reference"])
    ap.add_argument("--isl", type=int, default=8192)
    ap.add_argument("--no-extras", action="store_true", help="N=1: skip the per-kernel table, ISL sweep and CPU baseline (profiling runs)")
    ap.add_argument("--stream-requests", type=int, default=64, help="N>1: requests per pair in the config #3 closed-loop streams")
    ap.add_argument("--mixed-requests", type=int, default=32, help="N>1: requests per pair in the config #5 mixed-ISL stream")
    ap.add_argument("--one-gpu", action="store_true", help="N>1 debugging aid: all ranks share cuda:0 (gloo only; numbers meaningless)")
    ap.add_argument("--legs", default="all", help="N>1: which verified request-stream legs run after the headline: all | none | "
                                                  "comma list of config3,config4,config5")
    ap.add_argument("--mode", default="pull", choices=["pull", "push"],
                    help="N>1: decode GPU reads the prefill pool (pull, default) or prefill GPU stores into the decode pool")
    args = ap.parse_args()
    args.warmup = max(3, args.warmup) if args.impl == "kvx" else args.warmup
    if args.impl == "reference":
        return run_reference(args)
    if args.gpus == 1 and int(os.environ.get("WORLD_SIZE", "1")) == 1:
        return run_single(args)
    return run_pairs(args)


if __name__ == "__main__":
    sys.exit(main())
