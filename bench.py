#!/usr/bin/env python
"""bench.py -- Prometheus samples/s through the aggregate path (BASELINE.json metric).

    python bench.py --gpus N --steps K --warmup W            # this engine (libpmx.so on B200)
    python bench.py --impl reference --gpus N ...            # the pandas path on the host cores

A step is one pass of the hot path over one batch: fused align + rolling min/max/mean + rate over
every series of the shard, per-metric group reductions and exact p50/p95/p99 of the aligned values
(all-reduced across ranks when N > 1).  Workload at every N: config 2 of BASELINE.json per GPU
(64 nodes x 8 cards x 12 metrics, 24 h @ 15 s = 35 389 440 samples; weak scaling), synthetic,
seeded, generated on the device.  Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

N_SERIES, N_STEPS, N_METRICS = 6144, 5760, 12
WINDOW_STEPS, STEP_MS, START_MS, LOOKBACK_MS = 60, 15_000, 1_700_000_000_000, 300_000
QS = (0.5, 0.95, 0.99)
FUSED_BYTES_PER_SAMPLE = 56          # 8 ts + 8 val in, 5 x 8 out
METRIC = "Prometheus samples/sec through aggregate path"
WORKLOAD = ("config2: 64 nodes x 8 GPUs x 12 metrics, 24h @15s (35,389,440 samples) per GPU; "
            "window 15min, lookback 5min, p50/p95/p99 per metric")


def measured_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:  # noqa: BLE001
        return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled while the timed regions run."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "50", "-i", str(index)], stdout=subprocess.PIPE, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:  # noqa: BLE001
            self.proc = None
        self.marks = []

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), line.strip()))

    def mark(self, t0, t1):
        self.marks.append((t0, t1))

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        time.sleep(0.12)
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        inside = [r for r in self.rows if any(a - 0.05 <= r[0] <= b + 0.05 for a, b in self.marks)] or self.rows
        for _, line in inside:
            f = [x.strip() for x in line.split(",")]
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except Exception:  # noqa: BLE001
                continue
            for n, v in zip(names, f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# --------------------------------------------------------------------------- CPU legs
_CPU = {}


def _cpu_chunk(rng):
    """Worker: the pandas whole-step on series [a, b) of the forked parent's arrays."""
    from oracle import prom_oracle as O
    a, b = rng
    d = _CPU
    off = d["off"][a:b + 1] - d["off"][a]
    lo, hi = int(d["off"][a]), int(d["off"][b])
    O.aggregate_step_pandas(d["ts"][lo:hi], d["val"][lo:hi], off, START_MS, STEP_MS, N_STEPS,
                            WINDOW_STEPS * STEP_MS, d["gid"][a:b], N_METRICS, QS, LOOKBACK_MS)
    return int(hi - lo)


def cpu_time_pandas(ts, val, off, gid, procs):
    """Seconds for one pandas pass over the given series with ``procs`` worker processes."""
    import multiprocessing as mp
    S = len(off) - 1
    _CPU.update(ts=ts, val=val, off=off, gid=gid)
    if procs <= 1:
        t0 = time.perf_counter()
        _cpu_chunk((0, S))
        return time.perf_counter() - t0
    per = max(1, (S + procs - 1) // procs)
    chunks = [(a, min(S, a + per)) for a in range(0, S, per)]
    with mp.get_context("fork").Pool(procs) as pool:
        pool.map(_cpu_chunk, chunks[:1])                     # spin the workers up outside the timing
        t0 = time.perf_counter()
        pool.map(_cpu_chunk, chunks, chunksize=1)
        return time.perf_counter() - t0


def cpu_other_configs():
    """The pandas whole-step on a stated slice of configs 3, 4 and 5: one process, a few
    seconds each.  Returns a list of {config, value, unit, cores, kind, sample}."""
    import numpy as np
    from oracle import prom_oracle as O
    from k8s_rocm_metrics_dashboard_b200 import synth
    out = []

    def run(name, d, n_steps, step_ms, window_ms, what):
        S = len(d["off"]) - 1
        gid = (np.arange(S) % N_METRICS).astype(np.int32)
        t0 = time.perf_counter()
        O.aggregate_step_pandas(d["ts"], d["val"], d["off"], d["start_ms"], step_ms, n_steps, window_ms, gid, N_METRICS, QS,
                                LOOKBACK_MS)
        sec = time.perf_counter() - t0
        n = int(d["off"][-1])
        out.append({"config": name, "value": n / sec, "unit": "samples/s", "cores": 1, "kind": "port",
                    "sample": f"{what} ({n} samples, {sec:.1f} s), pandas {_pandas_version()} single process"})

    d = synth.dense(1, 8, 16, 40320, seed=3, nan_frac=1e-3, kinds="ttk")
    run("config3", d, 40320, d["step_ms"], WINDOW_STEPS * d["step_ms"], "1 of 1024 nodes: 128 series x 40320 steps")
    d = synth.dense(8, 8, 8, 5760, seed=4, all_counters=True, reset_prob=1e-3)
    run("config4", d, 5760, d["step_ms"], WINDOW_STEPS * d["step_ms"], "8 of 64 nodes (8 metrics each): 512 counter series with resets x 5760 steps")
    d = synth.ragged(4000, seed=5)
    d.setdefault("start_ms", int(d["ts"].min()))
    run("config5", d, 480, 15000, WINDOW_STEPS * 15000, "4000 of 1 000 000 ragged series resampled to a 15 s grid of 480 steps")
    return out


def host_sample(n_series, seed=0):
    """Config-2-shaped series on the host (same generator family as the device data)."""
    import numpy as np
    from k8s_rocm_metrics_dashboard_b200 import synth
    nodes = max(1, n_series // (8 * N_METRICS))
    d = synth.dense(nodes, 8, N_METRICS, N_STEPS, seed=seed, nan_frac=1e-3, kinds="ttk")
    gid = (np.arange(len(d["off"]) - 1) % N_METRICS).astype(np.int32)
    return d["ts"], d["val"], d["off"], gid


def run_reference(args):
    """The reference arm: the pandas restatement of the path (oracle port; the reference itself
    implements only the instant path) on all host cores, on a bounded sample per step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    n_series = min(N_SERIES, max(96, cores * 48))
    n_series -= n_series % (8 * N_METRICS)
    ts, val, off, gid = host_sample(n_series)
    # as much of the workload per step as a ~2.5 minute run allows, up to all of it
    t_cal = cpu_time_pandas(ts, val, off, gid, cores)
    want = int(150.0 / max(1, args.steps + min(args.warmup, 1)) / max(t_cal, 1e-3) * n_series)
    want = min(N_SERIES, want) - min(N_SERIES, want) % (8 * N_METRICS)
    if want > n_series:
        n_series = want
        ts, val, off, gid = host_sample(n_series)
    n = int(off[-1])
    for _ in range(max(0, min(args.warmup, 1))):
        cpu_time_pandas(ts, val, off, gid, cores)
    times = [cpu_time_pandas(ts, val, off, gid, cores) for _ in range(args.steps)]
    total = sum(times)
    v = n * args.steps / total
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": "samples/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": WORKLOAD, "series_per_step": n_series, "steps_per_series": N_STEPS,
                       "note": f"CPU arm: each step runs {n_series} of the workload's {N_SERIES} series (all of them when "
                               "steps x 1.5 s fits ~2.5 minutes; a per-sample rate otherwise); it runs on rank 0's host "
                               "cores only, whatever N is"},
            "cpu_baseline": {"value": v, "unit": "samples/s", "cores": cores, "kind": "port",
                             "sample": f"{n_series} of {N_SERIES} series x {N_STEPS} steps ({n} samples) per step, "
                                       f"pandas {_pandas_version()} fan-out over {cores} processes"},
            "e2e": {"value": v, "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def instant_path_times(eng):
    """The reference's OWN path (app.py:153-227, one node, 8 cards x 5 series = 40 samples per call):
    unmodified fetch_gpu_metrics under import stubs (baseline/_ref/app.py, copied there by build())
    against this engine's aggregate_instant on the same body.  A 40-sample call is latency-bound on
    a GPU; reported for completeness, not as the optimisation target."""
    try:
        from k8s_rocm_metrics_dashboard_b200 import adapter, synth
        from oracle import ref_app
        _, _, body = synth.instant_snapshot(0, 8, seed=0)
        out = {"samples_per_call": 40}
        for _ in range(3):
            adapter.aggregate_instant(body, eng)
        t0 = time.perf_counter()
        for _ in range(20):
            adapter.aggregate_instant(body, eng)
        out["engine_ms_per_call"] = (time.perf_counter() - t0) / 20 * 1e3
        if ref_app.reference_path():
            mod, _ = ref_app.load_reference()
            text = body.decode()
            for _ in range(3):
                ref_app.run_fetch(mod, text)
            t0 = time.perf_counter()
            for _ in range(20):
                ref_app.run_fetch(mod, text)
            out["reference_ms_per_call"] = (time.perf_counter() - t0) / 20 * 1e3
        else:
            out["reference_ms_per_call"] = None
        return out
    except Exception as e:  # noqa: BLE001
        return {"error": str(e)[:200]}


def config3_block(eng, sh, world, rank, barrier, steps=5):
    """One config-3 shard per GPU (16 384 series x 40 320 steps, 16 metrics): ms per step of the whole
    aggregate step, device-timed, max over ranks; and the same step without the exchange between
    ranks (what one GPU alone does on its shard), so that the cost of the collectives is visible in
    one run.  Data is generated on the device."""
    import numpy as np
    import torch
    import torch.distributed as dist
    from k8s_rocm_metrics_dashboard_b200 import group_index, synth
    from k8s_rocm_metrics_dashboard_b200.sharding import ShardedEngine
    S3, T3, M3 = 16384, 40320, 16
    dev = eng.device
    ts, val, off = synth.dense_device(S3, T3, M3, seed=3000 + rank, nan_frac=1e-3, device=dev)
    gid = (np.arange(S3) % M3).astype(np.int32)
    seg = torch.as_tensor(gid, device=dev)
    perm, goff = group_index(gid, M3)
    pg = (torch.as_tensor(perm, device=dev), torch.as_tensor(goff, device=dev))
    out = {k: torch.empty((S3, T3), dtype=torch.float64, device=dev)
           for k in ("aligned", "roll_min", "roll_max", "roll_mean", "rate")}
    alone = ShardedEngine(eng, local_only=True)

    def timed(engine_like):
        def one():
            engine_like.aggregate_step(ts, val, off, START_MS, STEP_MS, T3, WINDOW_STEPS, seg, M3, pg, QS, out=out,
                                       lookback_ms=LOOKBACK_MS, defer_check=True, graph=True)
        for _ in range(4):                                       # the graphs of the step are captured on the third call
            one()
        barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(steps):
            one()
        b.record()
        barrier()
        tt = torch.tensor([a.elapsed_time(b) / steps], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        return float(tt.item())

    ms = timed(sh)
    ms_alone = timed(alone) if world > 1 else ms
    ok = eng.deferred_ok(wait=True)
    n = S3 * T3
    return {"workload": f"config3 shard per GPU: {S3} series x {T3} steps x {M3} metrics ({n} samples); at 8 GPUs this is "
                        "config 3 (5,284,823,040 samples)",
            "ms_per_step": ms, "samples_per_s": world * n / (ms * 1e-3), "steps": steps,
            "ms_per_step_without_exchange": ms_alone, "exchange_efficiency": ms_alone / ms,
            "deferred_checks_ok": ok}


def e2e_text_block(eng, steps=5):
    """End to end from the RESPONSE TEXT: a query_range body in host memory -> H2D ->
    pmx_tokenize_device -> pmx_parse_samples -> fused align+rolling+rate -> per-metric summary and
    p50/p95/p99 -> summary back on the host.  Workload: 768 series x 5760 steps of config 2's mix
    (4.4 M samples; formatting all 35 M samples as text in Python would take minutes), the same
    calls ``adapter.aggregate_range`` makes, without building the pandas frame."""
    import numpy as np
    import torch
    from k8s_rocm_metrics_dashboard_b200 import adapter, group_index, synth
    d = synth.dense(8, 8, N_METRICS, N_STEPS, seed=7, nan_frac=1e-3, kinds="ttk")
    body = synth.to_matrix_json(d["labels"], d["ts"], d["val"], d["off"])
    n = len(d["ts"])
    out = None
    # the body in PINNED host memory (what response.raw.readinto(buffer) leaves behind)
    pinned = torch.frombuffer(bytearray(body), dtype=torch.uint8).pin_memory()

    def one():
        nonlocal out
        tok = eng.tokenize_device(pinned)                                 # H2D of the body + the cut
        ts, val, decoded = eng.parse_async(tok)                            # the decode kernel runs while the host interns the labels,
        labels = tok.labels()                                             # as aggregate_range does
        names, gid = adapter._group_ids(labels, ("__name__",))             # integer work on the interned table
        decoded()                                                         # the kernel's verdict on malformed tokens
        out = eng.fused(ts, val, tok.series_off, START_MS, STEP_MS, N_STEPS, WINDOW_STEPS, LOOKBACK_MS, out=out)
        pg = group_index(gid, len(names))
        sm, cnt, mn, mx, mean = eng.group_reduce(out["aligned"], None, len(names), perm_goff=pg)
        q = eng.quantiles(out["aligned"], gid, len(names), list(QS), row_order=pg[0], summary=(cnt, mn, mx))
        return torch.stack([sm, mean, q["value"][:, 0], q["value"][:, 1], q["value"][:, 2]]).cpu()

    for _ in range(2):
        one()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(steps):
        one()
    torch.cuda.synchronize()
    sec = (time.perf_counter() - t0) / steps
    return {"value": n / sec, "unit": "samples/s", "ms_per_step": sec * 1e3, "samples": n, "body_bytes": len(body),
            "h2d_bytes_per_step": len(body), "d2h_bytes_per_step": 5 * N_METRICS * 8 + 8 * (len(d["off"]) + 3 * (len(d["off"]) - 1)),
            "path": "body (pinned host bytes) -> H2D -> pmx_tokenize_device -> pmx_parse_samples -> fused -> group reduce + quantiles"}


def _pandas_version():
    import pandas
    return pandas.__version__


# --------------------------------------------------------------------------- GPU arm
def run_engine(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    from k8s_rocm_metrics_dashboard_b200 import Engine, group_index, synth
    from k8s_rocm_metrics_dashboard_b200.sharding import ShardedEngine

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; this engine has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    eng = Engine(local)
    sh = ShardedEngine(eng, peer_exchange=not args.no_peer_exchange)
    dev = eng.device

    S, T = N_SERIES, N_STEPS
    ts, val, off = synth.dense_device(S, T, N_METRICS, seed=1000 + rank, nan_frac=1e-3, device=dev)
    gid = (np.arange(S) % N_METRICS).astype(np.int32)
    seg = torch.as_tensor(gid, device=dev)
    perm, goff = group_index(gid, N_METRICS)
    perm_goff = (torch.as_tensor(perm, device=dev), torch.as_tensor(goff, device=dev))
    out = {k: torch.empty((S, T), dtype=torch.float64, device=dev)
           for k in ("aligned", "roll_min", "roll_max", "roll_mean", "rate")}
    n_samples = S * T

    checks = {"ok": True}

    def step():
        # no host synchronisation inside the step: whether every segment fitted the key exchange of
        # the cluster-wide quantiles is a device counter, looked at one step later (Engine.deferred_ok)
        r = sh.aggregate_step(ts, val, off, START_MS, STEP_MS, T, WINDOW_STEPS, seg, N_METRICS, perm_goff,
                              QS, out=out, lookback_ms=LOOKBACK_MS, defer_check=True, graph=True)
        checks["ok"] = eng.deferred_ok(wait=False) and checks["ok"]
        return r

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- parity before timing: every rank's shard of a small config against the oracle on the union
    parity = None
    if world > 1 and not args.no_parity_check:
        from tests.multi_gpu_check import parity_check
        parity = parity_check(eng, sh, rank, world)
        barrier()

    sampler = ClockSampler(local) if rank == 0 else None
    for _ in range(max(args.warmup, 3)):
        step()
    barrier()
    # ---- device-timed region: K steps, CUDA events, fused kernel timed on its own inside
    l0 = eng.launches
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    fe = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    orig_fused = eng.fused
    state = {"i": 0}

    def timed_fused(*a, **k):
        i = state["i"]
        if i < len(fe):
            fe[i][0].record()
        r = orig_fused(*a, **k)
        if i < len(fe):
            fe[i][1].record()
        state["i"] = i + 1
        return r

    eng.fused = timed_fused
    w0 = time.time()
    ev0.record()
    for _ in range(args.steps):
        step()
    ev1.record()
    barrier()
    w1 = time.time()
    checks["ok"] = eng.deferred_ok(wait=True) and checks["ok"]
    eng.fused = orig_fused
    launches = eng.launches - l0
    ms = ev0.elapsed_time(ev1)
    fused_ms = sum(a.elapsed_time(b) for a, b in fe) / len(fe)
    # the same kernel with the GPU to itself (inside the step, on one rank, the bracket kernels of the
    # summary run beside it on a side stream): reported next to the in-step figure, not instead of it
    fa = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(20)]
    for e0, e1 in fa:
        e0.record()
        eng.fused(ts, val, off, START_MS, STEP_MS, T, WINDOW_STEPS, LOOKBACK_MS, out=out)
        e1.record()
    torch.cuda.synchronize()
    fused_alone_ms = sum(a.elapsed_time(b) for a, b in fa[2:]) / len(fa[2:])
    tt = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    ms = float(tt.item())
    if sampler:
        sampler.mark(w0, w1)

    # ---- end to end through the public API with HOST buffers (pinned), copies inside the timing
    h_ts, h_val, h_off = ts.cpu().pin_memory(), val.cpu().pin_memory(), off.cpu().pin_memory()
    h_out = {k: torch.empty((S, T), dtype=torch.float64).pin_memory() for k in out}
    h_sum = torch.empty((5 + 3, N_METRICS), dtype=torch.float64).pin_memory()
    d_ts, d_val, d_off = torch.empty_like(ts), torch.empty_like(val), torch.empty_like(off)

    host = {"dev": {"ts": d_ts, "val": d_val, "off": d_off, **out}, "out": h_out}

    def e2e_step():
        # public API with HOST buffers: pipelined H2D -> fused kernel -> D2H over series slices,
        # then the cluster-wide summary, whose result is read back too
        o, summ, q = sh.aggregate_step(h_ts, h_val, h_off, START_MS, STEP_MS, T, WINDOW_STEPS, seg, N_METRICS,
                                       perm_goff, QS, lookback_ms=LOOKBACK_MS, host=host)
        packed = torch.stack([summ[0], summ[1].to(torch.float64), summ[2], summ[3], summ[4],
                              q["value"][:, 0], q["value"][:, 1], q["value"][:, 2]])
        h_sum.copy_(packed, non_blocking=True)
        torch.cuda.synchronize()

    e2e_steps = max(1, min(args.steps, 10))
    for _ in range(2):
        e2e_step()
    barrier()
    w2 = time.time()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_step()
    barrier()
    e2e_s = time.perf_counter() - t0
    w3 = time.time()
    te = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_s = float(te.item())
    if sampler:
        sampler.mark(w2, w3)
    clocks = sampler.stop() if sampler else None
    h2d = (h_ts.numel() + h_val.numel() + h_off.numel()) * 8
    d2h = sum(v.numel() for v in h_out.values()) * 8 + h_sum.numel() * 8

    # ---- config 3 (1024 nodes x 8 GPUs x 16 metrics, 7 d @ 15 s) = 16 384 series x 40 320 steps per GPU on
    # 8 GPUs: the same step on that shard, with and without the exchange between ranks
    cfg3 = None
    ns_cpu = 768                                                 # the CPU baseline's sample: keep it, free the rest
    c_ts, c_val = h_ts[:ns_cpu * T].numpy().copy(), h_val[:ns_cpu * T].numpy().copy()
    c_off = h_off[:ns_cpu + 1].numpy().copy()
    if not args.no_config3:
        del ts, val, off, out, h_ts, h_val, h_off, h_out, d_ts, d_val, d_off, host
        torch.cuda.empty_cache()
        cfg3 = config3_block(eng, sh, world, rank, barrier)

    if rank == 0:
        peak, which = measured_peak()
        achieved = n_samples * FUSED_BYTES_PER_SAMPLE / (fused_ms * 1e-3) / 1e9
        traffic = None
        try:
            with open(os.path.join(ROOT, "profiles", "fused_traffic.json")) as f:
                traffic = json.load(f).get("dram_bytes_per_launch")
        except Exception:  # noqa: BLE001
            pass
        cpu = None
        if world == 1 and not args.no_cpu_baseline:
            from oracle import prom_oracle  # noqa: F401  (checker timed as the CPU baseline)
            ns = ns_cpu
            sec = cpu_time_pandas(c_ts, c_val, c_off, gid[:ns], 1)
            cpu = {"value": ns * T / sec, "unit": "samples/s", "cores": 1, "kind": "port",
                   "sample": f"first {ns} of {S} series x {T} steps of the same device data ({ns * T} samples), "
                             f"pandas {_pandas_version()} single process, host has {os.cpu_count()} cores"}
        instant = None
        e2e_text = None
        cpu_other = None
        if world == 1 and not args.no_cpu_baseline:
            try:
                cpu_other = cpu_other_configs()
            except Exception as e:  # noqa: BLE001
                cpu_other = [{"error": str(e)[:200]}]
            instant = instant_path_times(eng)
            try:
                e2e_text = e2e_text_block(eng)
            except Exception as e:  # noqa: BLE001
                e2e_text = {"error": str(e)[:200]}
        line = {"metric": METRIC, "value": world * n_samples * args.steps / (ms * 1e-3), "unit": "samples/s",
                "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps,
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": {"workload": WORKLOAD, "series_per_gpu": S, "steps_per_series": T,
                           "l2": "inputs+outputs per step (2.0 GB) exceed the 126 MB L2; no explicit flush",
                           "parallelism": f"series sharded over {world} GPU(s); histogram/partial all-reduce only"},
                "roofline": {"bound": "hbm", "kernel": "window_fast_kernel<128,true> (fused align+rolling+rate)",
                             "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                             "traffic": traffic, "peak_source": which, "avg_launch_ms": fused_ms,
                             "algorithmic_bytes_per_launch": n_samples * FUSED_BYTES_PER_SAMPLE,
                             "alone": {"avg_launch_ms": fused_alone_ms,
                                       "achieved": n_samples * FUSED_BYTES_PER_SAMPLE / (fused_alone_ms * 1e-3) / 1e9,
                                       "frac": n_samples * FUSED_BYTES_PER_SAMPLE / (fused_alone_ms * 1e-3) / 1e9 / peak,
                                       "note": "same launches outside the timed region, nothing else on the GPU; inside the step "
                                               "(one rank) the bracket kernels of the summary share the GPU with it"}},
                "cpu_baseline": cpu, "cpu_baseline_other": cpu_other,
                "e2e": {"value": world * n_samples * e2e_steps / e2e_s, "unit": "samples/s",
                        "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "steps": e2e_steps},
                "gpu_launches": launches, "clocks": clocks, "instant_path": instant,
                "deferred_checks_ok": checks["ok"], "parity_check": parity, "config3": cfg3, "e2e_text": e2e_text}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    # stdout carries exactly one JSON line: anything a library prints meanwhile (NCCL's version
    # banner goes to stdout) is sent to stderr instead
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    sys.stdout = os.fdopen(real_stdout, "w")
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="engine", choices=["engine", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-parity-check", action="store_true")
    ap.add_argument("--no-config3", action="store_true")
    ap.add_argument("--no-peer-exchange", action="store_true", help="several ranks: NCCL calls instead of stores into peer memory")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_engine(args)


if __name__ == "__main__":
    main()
