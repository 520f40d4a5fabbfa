"""GPU: K2 align, K3 rolling, K4 rate and the fused kernel against the oracle (C ABI path)."""
import numpy as np
import pytest

from k8s_rocm_metrics_dashboard_b200 import synth
from oracle import prom_oracle as O
from tests.util import assert_bits_equal, assert_close, assert_exact_or_nan

pytestmark = pytest.mark.gpu


def _np(t):
    return t.cpu().numpy()


def _check_roll(eng, x, w, min_periods=1):
    mn, mx, mean, cnt = eng.rolling(x, w, min_periods)
    omn, omx, omean, ocnt = O.rolling(x, w, min_periods)
    assert_exact_or_nan(_np(mn), omn, f"min w={w}")
    assert_exact_or_nan(_np(mx), omx, f"max w={w}")
    assert np.array_equal(_np(cnt), ocnt), f"count w={w}"
    assert_close(_np(mean), omean, scale=O.rolling_scale(x, w), rtol=1e-12, what=f"mean w={w}")


@pytest.mark.parametrize("w", [1, 2, 7, 8, 9, 16, 17, 60, 61, 64, 240, 1000, 1025, 3000])
def test_rolling_windows(eng, w):
    rng = np.random.default_rng(w)
    T = 5000 if w <= 1025 else 3500
    x = np.round(rng.uniform(-50, 100, (5, T)), 3)
    x[rng.random(x.shape) < 0.03] = np.nan
    x[0, 100] = np.inf; x[1, 200] = -np.inf; x[2, :2100] = np.nan; x[3, 7] = -0.0
    _check_roll(eng, x, w)


@pytest.mark.parametrize("shape", [(1, 1), (3, 5), (1, 2047), (2, 2048), (2, 2049), (7, 1989), (1, 1990), (300, 17)])
def test_rolling_shapes(eng, shape):
    rng = np.random.default_rng(shape[1])
    x = rng.normal(size=shape)
    x[rng.random(shape) < 0.1] = np.nan
    for w in (1, 5, 60):
        _check_roll(eng, x, w)


def test_rolling_min_periods_and_all_nan(eng):
    rng = np.random.default_rng(0)
    x = rng.normal(size=(3, 700)); x[1, 300] = np.nan; x[2] = np.nan
    _check_roll(eng, x, 60, min_periods=60)
    _check_roll(eng, x, 60, min_periods=3)


def test_rolling_mean_is_fresh_per_window(eng):
    """A huge value leaving the window must not poison later means (no prefix differences)."""
    x = np.ones((1, 4000)); x[0, 10] = 1e300; x[0, 2500] = -1e250
    _check_roll(eng, x, 60)
    mn, mx, mean, cnt = eng.rolling(x, 60)
    assert _np(mean)[0, 100] == 1.0 and _np(mean)[0, 3000] == 1.0


def _check_rate(eng, x, w, step_ms=15000):
    inc, rt, nr = eng.rate(x, w, step_ms)
    oinc, ort, onr = O.rate(x, w, step_ms)
    assert np.array_equal(_np(nr), onr), f"n_resets w={w}"
    # relative to the RESULT (north_star: 1e-12 relative for rates); the oracle's increase is the
    # formula's exact value rounded once
    assert_close(_np(inc), oinc, rtol=1e-12, what=f"increase w={w}")
    with np.errstate(divide="ignore", invalid="ignore"):
        assert_close(_np(rt), ort, rtol=1e-12, what=f"rate w={w}")


@pytest.mark.parametrize("w", [1, 2, 3, 8, 9, 60, 100, 1030])
def test_rate_with_resets(eng, w):
    d = synth.dense(1, 2, 3, 2600, seed=w, all_counters=True, reset_prob=0.02, nan_frac=0.02)
    x = d["val"].reshape(6, 2600)
    x[5, :2200] = np.nan
    _check_rate(eng, x, w)


def test_rate_small_cases(eng):
    x = np.array([[10., 20., 5., 7., np.nan, 9., 2., 2.], [np.nan] * 8, [1., np.nan, np.nan, np.nan, np.nan, np.nan, np.nan, 2.]])
    _check_rate(eng, x, 4, 1000)
    _check_rate(eng, x, 8, 1000)


def _check_align(eng, d, n_steps, start=None, step=None, lookback=300_000):
    start = d.get("start_ms", synth.START_MS) if start is None else start
    step = d.get("step_ms", synth.STEP_MS) if step is None else step
    al, idx = eng.align(d["ts"], d["val"], d["off"], start, step, n_steps, lookback)
    oal, oidx = O.align(d["ts"], d["val"], d["off"], start, step, n_steps, lookback)
    assert np.array_equal(_np(idx), oidx), "alignment indices"
    assert_bits_equal(_np(al), oal, "aligned values")


def test_align_regular_and_empty(eng):
    d = synth.dense(2, 2, 3, 3000, seed=1, nan_frac=0.01)
    _check_align(eng, d, 3000)
    _check_align(eng, d, 7000, start=d["start_ms"] - 500 * 15000)       # grid wider than the data
    _check_align(eng, d, 100, step=60_000)                               # coarser grid (down-sampling)
    _check_align(eng, d, 4500, step=1000)                                # finer grid than the scrape
    d["off"] = np.concatenate([d["off"][:3], d["off"][2:]])              # an empty series in the middle
    d["off"][3] = d["off"][2]
    _check_align(eng, d, 3000)


def test_align_ragged_staleness_and_markers(eng):
    d = synth.ragged(300, seed=3, mean_len=300, nan_frac=0.02, stale_frac=0.02, gap_frac=0.03)
    t0 = int(d["ts"].min())
    _check_align(eng, d, 2500, start=t0, step=15000)
    _check_align(eng, d, 2500, start=t0, step=15000, lookback=20_000)
    _check_align(eng, d, 300, start=t0 + 7, step=200_000)                # many samples per grid step


def test_fused_matches_oracle_config1(eng):
    d = synth.config("config1", seed=0)
    w = 60
    out = eng.fused(d["ts"], d["val"], d["off"], d["start_ms"], d["step_ms"], d["n_steps"], w)
    ref = O.aggregate_range(d["ts"], d["val"], d["off"], d["start_ms"], d["step_ms"], d["n_steps"], w * d["step_ms"])
    assert_bits_equal(_np(out["aligned"]), ref["aligned"], "aligned")
    assert_exact_or_nan(_np(out["roll_min"]), ref["roll_min"], "roll_min")
    assert_exact_or_nan(_np(out["roll_max"]), ref["roll_max"], "roll_max")
    assert_close(_np(out["roll_mean"]), ref["roll_mean"], scale=O.rolling_scale(ref["aligned"], w), what="roll_mean")
    assert_close(_np(out["rate"]), ref["rate"], rtol=1e-12, what="rate")


@pytest.mark.parametrize("name,scale", [("config2", 1 / 32), ("config4", 1 / 32)])
def test_fused_matches_oracle_scaled_configs(eng, name, scale):
    d = synth.config(name, seed=1, scale=scale)
    w = 60
    out = eng.fused(d["ts"], d["val"], d["off"], d["start_ms"], d["step_ms"], d["n_steps"], w)
    ref = O.aggregate_range(d["ts"], d["val"], d["off"], d["start_ms"], d["step_ms"], d["n_steps"], w * d["step_ms"])
    assert_bits_equal(_np(out["aligned"]), ref["aligned"], "aligned")
    assert_exact_or_nan(_np(out["roll_min"]), ref["roll_min"], "roll_min")
    assert_exact_or_nan(_np(out["roll_max"]), ref["roll_max"], "roll_max")
    assert_close(_np(out["roll_mean"]), ref["roll_mean"], scale=O.rolling_scale(ref["aligned"], w), what="roll_mean")
    assert_close(_np(out["rate"]), ref["rate"], rtol=1e-12, what="rate")


def test_fused_equals_unfused_at_full_config2(eng):
    """Size-independent property at BASELINE.json's full size: the fused kernel and the three
    separate kernels (align -> rolling, rate) agree bit for bit on 35.4 M samples."""
    t = eng.torch
    ts, val, off = synth.dense_device(6144, 5760, 12, seed=0, device=eng.device)
    out = eng.fused(ts, val, off, synth.START_MS, synth.STEP_MS, 5760, 60)
    al, _ = eng.align(ts, val, off, synth.START_MS, synth.STEP_MS, 5760, want_idx=False)
    assert t.equal(out["aligned"].view(t.int64), al.view(t.int64))
    mn, mx, mean, _ = eng.rolling(al, 60, want_count=False)
    for a, b in ((out["roll_min"], mn), (out["roll_max"], mx), (out["roll_mean"], mean)):
        assert t.equal(a.view(t.int64), b.view(t.int64))
    # pmx_rate with n_resets runs the general kernel (true division, ordered fold): 1e-12, not bits
    _, rt, _ = eng.rate(al, 60, synth.STEP_MS)
    both_nan = t.isnan(rt) & t.isnan(out["rate"])
    err = (out["rate"] - rt).abs()
    assert bool((both_nan | (err <= 1e-12 * rt.abs())).all())
    # rolling min <= mean <= max wherever defined; first grid point of every series has no sample
    ok = ~t.isnan(mean)
    assert bool(((mn <= mean + 1e-9 * mean.abs()) & (mean <= mx + 1e-9 * mx.abs()))[ok].all())
    assert bool(t.isnan(al[:, 0]).float().mean() > 0.99)


def test_fused_window_limit_is_reported_and_engine_falls_back(eng):
    """The C entry refuses windows beyond its tile (513 steps) with a message; Engine.fused then runs
    align + rolling + rate instead (a 3 h window at a 15 s step), same results as the oracle."""
    import ctypes as C
    from k8s_rocm_metrics_dashboard_b200 import _abi
    t = eng.torch
    d = synth.dense(1, 1, 3, 1500, seed=0, nan_frac=0.01, kinds="ttk")
    ts, val, off = (t.as_tensor(d[k], device=eng.device) for k in ("ts", "val", "off"))
    outs = [t.empty((3, 1500), dtype=t.float64, device=eng.device) for _ in range(5)]
    p = lambda x: C.c_void_p(x.data_ptr())
    rc = eng.lib.pmx_fused_align_rolling_rate(eng.ctx, p(ts), p(val), p(off), 3, d["start_ms"], d["step_ms"], 1500, 300000, 720, 1,
                                              *(p(o) for o in outs), C.c_void_p(t.cuda.current_stream().cuda_stream))
    assert rc == _abi.E_LIMIT and b"exceeds the fused limit" in eng.lib.pmx_last_error()
    _check_fused(eng, d, d["start_ms"], d["step_ms"], 1500, 720)
    with pytest.raises(ValueError, match="window_steps must be >= 1"):
        eng.fused(d["ts"], d["val"], d["off"], d["start_ms"], d["step_ms"], 100, 0)


def test_ragged_rolling_and_rate(eng):
    d = synth.ragged(400, seed=5, mean_len=150, nan_frac=0.02, stale_frac=0.01, gap_frac=0.02)
    mn, mx, mean, cnt = eng.rolling_ragged(d["ts"], d["val"], d["off"], 900_000)
    omn, omx, omean, ocnt = O.rolling_ragged(d["ts"], d["val"], d["off"], 900_000)
    assert_exact_or_nan(_np(mn), omn); assert_exact_or_nan(_np(mx), omx)
    assert np.array_equal(_np(cnt), ocnt)
    assert_close(_np(mean), omean, rtol=1e-12)
    inc, rt, nr = eng.rate_ragged(d["ts"], d["val"], d["off"], 900_000)
    oinc, ort, onr = O.rate_ragged(d["ts"], d["val"], d["off"], 900_000)
    assert np.array_equal(_np(nr), onr)
    assert_close(_np(inc), oinc, rtol=1e-12, what="ragged increase")
    assert_close(_np(rt), ort, rtol=1e-12, what="ragged rate")


def _check_fused(eng, d, start, step, n_steps, w, lookback=300_000):
    out = eng.fused(d["ts"], d["val"], d["off"], start, step, n_steps, w, lookback)
    ref = O.aggregate_range(d["ts"], d["val"], d["off"], start, step, n_steps, w * step, lookback)
    assert_bits_equal(_np(out["aligned"]), ref["aligned"], "aligned")
    assert_exact_or_nan(_np(out["roll_min"]), ref["roll_min"], "roll_min")
    assert_exact_or_nan(_np(out["roll_max"]), ref["roll_max"], "roll_max")
    assert_close(_np(out["roll_mean"]), ref["roll_mean"], scale=O.rolling_scale(ref["aligned"], w), what="roll_mean")
    assert_close(_np(out["rate"]), ref["rate"], rtol=1e-12, what="rate")


@pytest.mark.parametrize("w", [9, 16, 60, 61, 63, 64, 250])
def test_fused_gaps_nan_runs_and_stale_markers(eng, w):
    """Ragged series with 6-30 min scrape gaps (> lookback): long NaN runs in the aligned grid,
    windows whose first chunk part is empty, stale markers, series that start late / end early."""
    d = synth.ragged(120, seed=w, mean_len=500, nan_frac=0.03, stale_frac=0.02, gap_frac=0.04)
    t0 = int(d["ts"].min())
    _check_fused(eng, d, t0 - 45_000, 15_000, 2600 + w, w)


def test_fused_many_samples_per_step_uses_global_search(eng):
    d = synth.ragged(40, seed=2, mean_len=3000, nan_frac=0.01, stale_frac=0.0, gap_frac=0.0)
    t0 = int(d["ts"].min())
    _check_fused(eng, d, t0, 200_000, 1500, 12)            # ~13 samples per grid step
    _check_fused(eng, d, t0, 1_000, 5000, 30, lookback=20_000)   # 15 grid steps per sample


@pytest.mark.parametrize("T", [1021, 1022, 1023, 4097])
def test_fused_row_alignment_variants(eng, T):
    """n_steps not a multiple of 4: every row starts at a different 32-byte phase."""
    d = synth.dense(1, 3, 3, T, seed=T, nan_frac=0.02, kinds="ttk")
    _check_fused(eng, d, d["start_ms"], d["step_ms"], T, 60)


def test_rate_long_nan_runs(eng):
    rng = np.random.default_rng(8)
    x = np.cumsum(rng.uniform(0, 100, (6, 3000)), axis=1)
    for s in range(6):
        pos = 0
        while pos < 3000:
            run = int(rng.integers(1, 40))
            if rng.random() < 0.5:
                x[s, pos:pos + run] = np.nan
            elif rng.random() < 0.2:
                x[s, pos:pos + run] -= x[s, pos] * 0.999       # a reset
            pos += run
    for w in (9, 12, 60, 200):
        _check_rate(eng, x, w)
        _check_roll(eng, x, w)


def test_general_kernel_path_for_unaligned_outputs(eng):
    """Output rows that are not 32-byte aligned fall back to the general kernel; same results."""
    t = eng.torch
    import ctypes as C
    from k8s_rocm_metrics_dashboard_b200 import _abi
    rng = np.random.default_rng(3)
    x = rng.normal(size=(3, 1000)); x[rng.random(x.shape) < 0.05] = np.nan
    dx = t.as_tensor(x, device=eng.device)
    buf = t.empty(3 * 1000 + 1, dtype=t.float64, device=eng.device)
    mean_unaligned = buf[1:]                                    # 8-byte aligned only
    _abi.check(eng.lib.pmx_rolling(eng.ctx, C.c_void_p(dx.data_ptr()), 3, 1000, 60, 1, None, None,
                                   C.c_void_p(mean_unaligned.data_ptr()), None,
                                   C.c_void_p(t.cuda.current_stream().cuda_stream)))
    _, _, mean, _ = eng.rolling(x, 60)
    assert t.equal(mean.view(t.int64).ravel(), mean_unaligned.contiguous().view(t.int64)) or \
        np.allclose(_np(mean).ravel(), _np(mean_unaligned), rtol=1e-14, equal_nan=True)


def test_fused_from_host_pipeline_equals_fused(eng):
    """Host-buffer entry (three streams, series slices) returns the same bits as the one-shot call."""
    t = eng.torch
    d = synth.dense(3, 8, 6, 700, seed=31, nan_frac=0.01, kinds="ttk")
    S, T = len(d["off"]) - 1, 700
    ref = eng.fused(d["ts"], d["val"], d["off"], d["start_ms"], d["step_ms"], T, 60)
    h_ts, h_val, h_off = (t.as_tensor(d[k]).pin_memory() for k in ("ts", "val", "off"))
    keys = ("aligned", "roll_min", "roll_max", "roll_mean", "rate")
    dev = {"ts": t.empty(len(d["ts"]), dtype=t.int64, device=eng.device),
           "val": t.empty(len(d["val"]), dtype=t.float64, device=eng.device),
           "off": t.empty(S + 1, dtype=t.int64, device=eng.device),
           **{k: t.full((S, T), -1.0, dtype=t.float64, device=eng.device) for k in keys}}
    h_out = {k: t.empty((S, T), dtype=t.float64).pin_memory() for k in keys}
    for n_chunks in (1, 5, 8):
        out = eng.fused_from_host(h_ts, h_val, h_off, d["start_ms"], d["step_ms"], T, 60, dev, h_out, n_chunks=n_chunks)
        t.cuda.synchronize()
        for k in keys:
            assert t.equal(out[k].view(t.int64), ref[k].view(t.int64)), k
            assert t.equal(h_out[k].view(t.int64), ref[k].cpu().view(t.int64)), k


def _check_ragged(eng, d, window_ms):
    mn, mx, mean, cnt = eng.rolling_ragged(d["ts"], d["val"], d["off"], window_ms)
    omn, omx, omean, ocnt = O.rolling_ragged(d["ts"], d["val"], d["off"], window_ms)
    assert_exact_or_nan(_np(mn), omn, "ragged min"); assert_exact_or_nan(_np(mx), omx, "ragged max")
    assert np.array_equal(_np(cnt), ocnt), "ragged count"
    assert_close(_np(mean), omean, rtol=1e-12, what="ragged mean")
    inc, rt, _ = eng.rate_ragged(d["ts"], d["val"], d["off"], window_ms, want_resets=False)
    oinc, ort, _ = O.rate_ragged(d["ts"], d["val"], d["off"], window_ms)
    assert_close(_np(inc), oinc, rtol=1e-12, what="ragged increase")
    assert_close(_np(rt), ort, rtol=1e-12, what="ragged rate")


@pytest.mark.parametrize("window_ms", [20_000, 120_000, 900_000, 3_600_000])
def test_ragged_tiled_kernel_windows(eng, window_ms):
    """Config-5 shaped data through the tiled kernel: windows of 1-2 samples (inside one chunk),
    ~8, ~60 and ~240 samples (the last one longer than the halo: direct folds)."""
    d = synth.ragged(300, seed=window_ms % 97, mean_len=250, nan_frac=0.03, stale_frac=0.01, gap_frac=0.03)
    _check_ragged(eng, d, window_ms)


def test_ragged_tiled_kernel_tiny_and_empty_series(eng):
    """Hundreds of series starts per tile, single-sample and empty series, long NaN runs."""
    rng = np.random.default_rng(4)
    lens = rng.integers(0, 5, 3000)
    lens[::50] = rng.integers(200, 1500, len(lens[::50]))
    off = np.zeros(len(lens) + 1, dtype=np.int64); np.cumsum(lens, out=off[1:])
    N = int(off[-1])
    gaps = 15_000 + rng.integers(-2000, 2001, N)
    ts = np.cumsum(gaps).astype(np.int64)
    val = np.round(rng.uniform(0, 1000, N), 3)
    pos = 0
    while pos < N:                                              # NaN runs of up to 30 samples
        run = int(rng.integers(1, 60))
        if rng.random() < 0.3:
            val[pos:pos + min(run, 30)] = np.nan
        pos += run
    d = {"ts": ts, "val": val, "off": off}
    _check_ragged(eng, d, 900_000)
    _check_ragged(eng, d, 60_000)


def test_ragged_tiled_kernel_more_series_starts_than_fit(eng):
    """> 1152 (empty) series starting inside one tile: that tile's outputs are folded directly."""
    rng = np.random.default_rng(6)
    lens = np.concatenate([[700], np.zeros(2000, dtype=np.int64), [900], np.zeros(10, dtype=np.int64), [1500]])
    off = np.zeros(len(lens) + 1, dtype=np.int64); np.cumsum(lens, out=off[1:])
    N = int(off[-1])
    ts = np.cumsum(15_000 + rng.integers(-2000, 2001, N)).astype(np.int64)
    val = np.round(rng.uniform(0, 100, N), 3); val[rng.random(N) < 0.02] = np.nan
    _check_ragged(eng, {"ts": ts, "val": val, "off": off}, 900_000)


@pytest.mark.parametrize("w", [9, 60, 1030])
def test_extrapolated_rate_mode(eng, w):
    d = synth.dense(1, 2, 3, 2600, seed=w, all_counters=True, reset_prob=0.02, nan_frac=0.05)
    x = d["val"].reshape(6, 2600); x[5, :2200] = np.nan; x[4, 100:400] = np.nan
    inc, rt, nr = eng.rate(x, w, 15000, mode=1)
    oinc, ort, onr = O.rate(x, w, 15000, mode=1)
    assert np.array_equal(_np(nr), onr)
    assert_close(_np(inc), oinc, rtol=1e-12, what="extrapolated increase")
    assert_close(_np(rt), ort, rtol=1e-12, what="extrapolated rate")


def test_extrapolated_rate_mode_ragged(eng):
    d = synth.ragged(150, seed=11, mean_len=200, nan_frac=0.03, stale_frac=0.01, gap_frac=0.03)
    d["val"] = np.abs(np.cumsum(np.where(np.isnan(d["val"]), 0.0, d["val"])) % 5000.0)   # counter-like with resets
    inc, rt, nr = eng.rate_ragged(d["ts"], d["val"], d["off"], 900_000, mode=1)
    oinc, ort, onr = O.rate_ragged(d["ts"], d["val"], d["off"], 900_000, mode=1)
    assert np.array_equal(_np(nr), onr)
    assert_close(_np(inc), oinc, rtol=1e-12, what="ragged extrapolated increase")
    assert_close(_np(rt), ort, rtol=1e-12, what="ragged extrapolated rate")
