#!/usr/bin/env python
"""Per-kernel throughput on one B200 (not the driver's bench.py contract; feeds profiles/).

    python bench_kernels.py [--out profiles/kernels_rNN.json] [--big]

Times every C-ABI kernel with CUDA events on torch's current stream (the stream the kernels are
launched on), inputs resident in HBM and larger than L2, 3 warm-ups + 10 timed launches, and
reports samples/s and algorithmic GB/s (bytes per sample as in SURVEY.md 8(d) / DESIGN.md 4)
against the measured HBM peak.  --big adds the config-3 per-GPU shard (16 384 series x 40 320).
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np   # noqa: E402
import torch         # noqa: E402

from k8s_rocm_metrics_dashboard_b200 import Engine, group_index, synth, tokenize  # noqa: E402


def timeit(fn, iters=10, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters * 1e-3


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=None)
    ap.add_argument("--big", action="store_true")
    args = ap.parse_args()
    peak = 6555.2
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:  # noqa: BLE001
        pass
    eng = Engine(0)
    dev = eng.device
    rows = []

    def rec(name, n, bytes_per_sample, sec, note=""):
        gbs = n * bytes_per_sample / sec / 1e9
        rows.append({"kernel": name, "samples": n, "ms": sec * 1e3, "samples_per_s": n / sec,
                     "bytes_per_sample": bytes_per_sample, "GBps": gbs, "frac_of_measured_peak": gbs / peak, "note": note})
        print(f"{name:34s} {n / 1e6:9.1f} M samples  {sec * 1e3:8.3f} ms  {n / sec / 1e9:7.2f} G/s  "
              f"{gbs:7.0f} GB/s  {100 * gbs / peak:5.1f}%  {note}", flush=True)

    # ---- config 2 shape
    S, T, W = 6144, 5760, 60
    ts, val, off = synth.dense_device(S, T, 12, seed=1, device=dev)
    N = S * T
    out = {k: torch.empty((S, T), dtype=torch.float64, device=dev) for k in ("aligned", "roll_min", "roll_max", "roll_mean", "rate")}
    rec("fused align+rolling+rate (cfg2)", N, 56, timeit(lambda: eng.fused(ts, val, off, synth.START_MS, synth.STEP_MS, T, W, out=out)))
    al = out["aligned"]
    rec("K2 align (values+idx)", N, 28, timeit(lambda: eng.align(ts, val, off, synth.START_MS, synth.STEP_MS, T)))
    rec("K3 rolling min/max/mean", N, 32, timeit(lambda: eng.rolling(al, W, want_count=False)))
    rec("K3 rolling + count", N, 36, timeit(lambda: eng.rolling(al, W)))
    rec("K4 rate+increase", N, 24, timeit(lambda: eng.rate(al, W, synth.STEP_MS, want_resets=False)))
    rec("K4 rate+increase+n_resets", N, 28, timeit(lambda: eng.rate(al, W, synth.STEP_MS)))
    gid = (np.arange(S) % 12).astype(np.int32)
    pg = group_index(gid, 12)
    pg = (torch.as_tensor(pg[0], device=dev), torch.as_tensor(pg[1], device=dev))
    rec("K5 group reduce per metric", N, 8, timeit(lambda: eng.group_reduce(al, None, 12, perm_goff=pg)))
    rec("K5 group reduce per step [12,T]", N, 8, timeit(lambda: eng.group_reduce(al, None, 12, per_step=True, perm_goff=pg)))
    seg = torch.as_tensor(gid, device=dev)
    rec("K6 quantiles p50/p95/p99 x12", N, 24, timeit(lambda: eng.quantiles(al, seg, 12, [0.5, 0.95, 0.99], row_order=pg[0])),
        "2 full histogram passes + the list pass (no summary given: pass 0 included)")
    rec("K5+K6 summary + p50/p95/p99 x12, one sweep", N, 8,
        timeit(lambda: eng.summary_quantiles(al, seg, 12, (0.5, 0.95, 0.99), pg)),
        "pmx_summary_quantiles: sample, brackets, ONE read of the matrix, select (all serial here; in the step the first two run beside the fused kernel)")
    # one refresh of the series cache (one new grid point per series) against the full pass
    from k8s_rocm_metrics_dashboard_b200.cache import SeriesCache
    labs = synth.series_labels(64, 8, 12)
    keep = 200
    cache = SeriesCache(eng, synth.START_MS, synth.STEP_MS, W, capacity=256, max_series=S)
    cache.append(labs, ts.view(S, T)[:, T - keep:].reshape(-1).contiguous(), val.view(S, T)[:, T - keep:].reshape(-1).contiguous(),
                 np.arange(S + 1, dtype=np.int64) * keep)
    upto = synth.START_MS + (T - 1) * synth.STEP_MS + synth.STEP_MS - 1

    def one_refresh():
        cache.next_step = T - 1
        cache.refresh(upto)
    sec = timeit(one_refresh)
    rec("series cache: refresh of 1 new step", S, 56 * (W + 20), sec, f"{S} series, window {W}; the full-range fused pass above does the same in {rows[0]['ms']:.3f} ms")
    del cache
    # config 4: all counters with resets
    ts4, val4, off4 = synth.dense_device(S, T, 12, seed=2, all_counters=True, reset_prob=1e-3, nan_frac=0.0, device=dev)
    rec("fused (cfg4 counters + resets)", N, 56, timeit(lambda: eng.fused(ts4, val4, off4, synth.START_MS, synth.STEP_MS, T, W, out=out)))
    del ts4, val4, off4

    # ---- K1 parse: config-1-shaped text, tiled to ~40 M samples
    d1 = synth.dense(16, 8, 12, 5760, seed=3, nan_frac=1e-3, kinds="ttk")      # 8.8 M samples, ~240 MB of text (> L2)
    t0 = time.perf_counter()
    body = synth.to_matrix_json(d1["labels"], d1["ts"], d1["val"], d1["off"])
    t_fmt = time.perf_counter() - t0
    tokenize(body, threads=1)                                  # touch the pages once
    t0 = time.perf_counter()
    tok = tokenize(body, threads=1)
    t_tok = time.perf_counter() - t0
    n1 = tok.n_samples
    ncpu = os.cpu_count() or 1
    tokenize(body, threads=ncpu)
    t0 = time.perf_counter()
    tokenize(body, threads=ncpu)
    t_mt = time.perf_counter() - t0
    print(f"tokenizer (host, {ncpu} threads): {n1 / t_mt / 1e6:.1f} M samples/s, {len(body) / t_mt / 1e9:.2f} GB/s")
    rows.append({"kernel": f"X1 tokenizer (host C++, {ncpu} threads)", "samples": n1, "ms": t_mt * 1e3,
                 "samples_per_s": n1 / t_mt, "bytes_per_sample": len(body) / n1, "GBps": len(body) / t_mt / 1e9})
    print(f"tokenizer (host, 1 core): {len(body) / 1e6:.1f} MB, {n1 / 1e6:.2f} M samples in {t_tok * 1e3:.1f} ms = "
          f"{n1 / t_tok / 1e6:.1f} M samples/s, {len(body) / t_tok / 1e9:.2f} GB/s (text formatting took {t_fmt:.1f} s)")
    rows.append({"kernel": "X1 tokenizer (host C++, 1 core)", "samples": n1, "ms": t_tok * 1e3, "samples_per_s": n1 / t_tok,
                 "bytes_per_sample": len(body) / n1, "GBps": len(body) / t_tok / 1e9})
    # X1 on the device: the body is copied once (pageable host memory here) and cut by pmx_tokenize_device
    dtok = eng.tokenize_device(body)
    assert dtok.n_samples == n1
    t0 = time.perf_counter()
    dtok = eng.tokenize_device(body)
    t_dev = time.perf_counter() - t0
    import ctypes as C
    from k8s_rocm_metrics_dashboard_b200 import _abi
    cap_s, cap_r = len(body) // 8 + 1, len(body) // 24 + 1
    b_ts, b_val = torch.empty(cap_s, dtype=torch.int32, device=dev), torch.empty(cap_s, dtype=torch.int32, device=dev)
    b_so = torch.empty(cap_r + 1, dtype=torch.int64, device=dev)
    b_mo, b_ml = torch.empty(cap_r, dtype=torch.int32, device=dev), torch.empty(cap_r, dtype=torch.int32, device=dev)
    b_st = torch.empty(4, dtype=torch.int64, device=dev)
    pp = lambda x: C.c_void_p(x.data_ptr())
    sec = timeit(lambda: _abi.check(eng.lib.pmx_tokenize_device(
        eng.ctx, pp(dtok.d_text), len(body), pp(b_ts), pp(b_val), cap_s, pp(b_so), pp(b_mo), pp(b_ml), cap_r, pp(b_st),
        C.c_void_p(torch.cuda.current_stream().cuda_stream))))
    rec("X1 tokenizer (device kernels)", n1, len(body) / n1 + 8, sec,
        f"{len(body) / sec / 1e9:.0f} GB/s of text; with H2D of the body and the label spans back: {n1 / t_dev / 1e6:.0f} M samples/s")
    del b_ts, b_val, b_so, b_mo, b_ml, dtok
    text = torch.from_numpy(tok.flat.copy()).to(dev)
    voff = torch.from_numpy(tok.val_off.view(np.int32).copy()).to(dev)
    toff = torch.from_numpy(tok.ts_off.view(np.int32).copy()).to(dev)
    import ctypes as C
    from k8s_rocm_metrics_dashboard_b200 import _abi
    pv, pt = torch.empty(n1, dtype=torch.float64, device=dev), torch.empty(n1, dtype=torch.int64, device=dev)
    bad = torch.empty(2, dtype=torch.int64, device=dev)
    p = lambda x: C.c_void_p(x.data_ptr())
    st = lambda: C.c_void_p(torch.cuda.current_stream().cuda_stream)

    def parse():
        _abi.check(eng.lib.pmx_parse_samples(eng.ctx, p(text), text.numel(), p(toff), p(voff), n1, p(pt), p(pv), p(bad), st()))
    lv = float(np.mean([len(synth.format_value(v)) for v in d1["val"][:20000]]))
    lt = float(np.mean([len(synth.format_ts(v)) for v in d1["ts"][:20000]]))
    rec("K1 parse value+timestamp", n1, lv + lt + 24, timeit(parse), f"mean token bytes {lv:.1f}+{lt:.1f}; working set {len(body) / 1e6:.0f} MB")
    del text, voff, toff, pv, pt

    # ---- config 5 at its stated size: 1 000 000 ragged series (~400 M samples), generated on the device
    del ts, val, off, out, al
    torch.cuda.empty_cache()
    r_ts, r_val, r_off = synth.ragged_device(1_000_000, seed=5, device=dev)
    n5 = int(r_off[-1])
    rec("ragged rolling (cfg5, 1M series)", n5, 44, timeit(lambda: eng.rolling_ragged(r_ts, r_val, r_off, 900_000), iters=3, warm=1))
    rec("ragged rate (cfg5, 1M series)", n5, 36, timeit(lambda: eng.rate_ragged(r_ts, r_val, r_off, 900_000, want_resets=False), iters=3, warm=1))
    t0i = int(r_ts.min())
    rec("fused on ragged cfg5 -> 15 s grid", 1_000_000 * 480, 56,
        timeit(lambda: eng.fused(r_ts, r_val, r_off, t0i, 15_000, 480, 60), iters=3, warm=1), "per GRID point; 480 steps (2 h)")
    rec("bucket 15 s -> 1 min (cfg5)", n5, 16 + 20 * 480 * 1_000_000 / 4 / n5, timeit(lambda: eng.bucket(r_ts, r_val, r_off, t0i, 60_000, 120), iters=3, warm=1),
        "resample().last/mean/count; 120 one-minute bins per series")
    del r_ts, r_val, r_off

    if args.big:
        torch.cuda.empty_cache()
        S3, T3 = 16384, 40320
        ts, val, off = synth.dense_device(S3, T3, 16, seed=4, device=dev)
        out = {k: torch.empty((S3, T3), dtype=torch.float64, device=dev) for k in ("aligned", "roll_min", "roll_max", "roll_mean", "rate")}
        rec("fused (cfg3 per-GPU shard)", S3 * T3, 56, timeit(lambda: eng.fused(ts, val, off, synth.START_MS, synth.STEP_MS, T3, W, out=out), iters=3, warm=1),
            "16384 series x 40320 steps, 37 GB of traffic per launch")
    res = {"peak_GBps": peak, "gpu": torch.cuda.get_device_name(0), "rows": rows}
    if args.out:
        with open(args.out, "w") as f:
            json.dump(res, f, indent=1)


if __name__ == "__main__":
    main()
