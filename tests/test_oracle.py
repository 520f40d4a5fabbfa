"""CPU: the oracle against the golden vectors and against pandas/numpy themselves."""
import json
import os
import struct

import numpy as np
import pandas as pd
import pytest

from k8s_rocm_metrics_dashboard_b200 import synth
from oracle import prom_oracle as O
from tests.util import assert_bits_equal, assert_close, assert_exact_or_nan


def fbits(x):
    return "%016x" % struct.unpack("<Q", struct.pack("<d", float(x)))[0]


def test_float_golden_matches_python_float(golden_dir):
    g = json.load(open(os.path.join(golden_dir, "float_golden.json")))
    assert len(g) > 1000
    for tok, exp in g:
        try:
            got = fbits(O.parse_value(tok))
        except ValueError:
            got = "ValueError"
        assert got == exp, tok


def test_instant_restatement_matches_reference_golden(golden_dir):
    """oracle.aggregate_instant == what the UNMODIFIED reference returned (tests/golden/make_golden.py)."""
    g = json.load(open(os.path.join(golden_dir, "instant_golden.json")))
    assert g["versions"]["pandas"] == pd.__version__, "regenerate goldens for this pandas"
    for case in g["cases"]:
        result = json.loads(case["body"])["data"]["result"]
        if "error" in case:
            with pytest.raises(Exception) as ei:
                O.aggregate_instant(result)
            assert str(ei.value) in case["error"], case["name"]
            continue
        piv, stats = O.aggregate_instant(result)
        assert list(piv.index) == case["pivot"]["index"], case["name"]
        assert list(piv.columns) == case["pivot"]["columns"], case["name"]
        for gid, rowd in case["pivot"]["cells"].items():
            for c, v in rowd.items():
                cell = piv.loc[gid, c]
                assert (fbits(cell) if isinstance(cell, float) else cell) == v, (case["name"], gid, c)
        for k in ("mean", "max", "min"):
            for c, v in case["stats"][k].items():
                assert fbits(stats[k][c]) == v, (case["name"], k, c)
        avg = O.selected_averages(piv, case["selected"])
        for c, v in case["averages"].items():
            assert fbits(avg[c]) == v, (case["name"], "avg", c)


def test_reference_runs_here_and_matches_golden(golden_dir):
    """When the reference is present (build container), re-run it and compare with the fixture."""
    from oracle import ref_app
    if ref_app.reference_path() is None:
        pytest.skip("reference app.py not on this machine")
    mod, errors = ref_app.load_reference()
    g = json.load(open(os.path.join(golden_dir, "instant_golden.json")))
    case = g["cases"][0]
    piv, stats = ref_app.run_fetch(mod, case["body"])
    assert list(piv.index) == case["pivot"]["index"]
    for c, v in case["stats"]["mean"].items():
        assert fbits(stats["mean"][c]) == v


def _pandas_rolling(x, w, step_ms):
    idx = pd.to_datetime(synth.START_MS + step_ms * np.arange(x.shape[1], dtype=np.int64), unit="ms")
    df = pd.DataFrame(x.T, index=idx)
    r = df.rolling(f"{w * step_ms}ms")
    return r.min().values.T, r.max().values.T, r.mean().values.T, r.count().values.T


@pytest.mark.parametrize("w", [1, 4, 60])
def test_rolling_matches_pandas(w):
    rng = np.random.default_rng(w)
    x = np.round(rng.uniform(0, 100, (6, 300)), 3)
    x[rng.random(x.shape) < 0.05] = np.nan
    x[0, 10] = np.inf; x[1, 20] = -np.inf; x[2, :80] = np.nan; x[3, 5] = -0.0
    mn, mx, mean, cnt = O.rolling(x, w)
    pmn, pmx, pmean, pcnt = _pandas_rolling(x, w, 15000)
    assert_exact_or_nan(mn, pmn, "min"); assert_exact_or_nan(mx, pmx, "max")
    assert np.array_equal(cnt, pcnt.astype(np.int32))
    assert_close(mean, pmean, scale=O.rolling_scale(x, w), rtol=1e-12, what="mean")


def test_rolling_fixed_count_min_periods():
    rng = np.random.default_rng(0)
    x = rng.normal(size=(3, 200)); x[1, 50] = np.nan
    mn, mx, mean, cnt = O.rolling(x, 60, min_periods=60)
    r = pd.DataFrame(x.T).rolling(60)
    assert_exact_or_nan(mn, r.min().values.T); assert_exact_or_nan(mx, r.max().values.T)
    assert_close(mean, r.mean().values.T, scale=O.rolling_scale(x, 60), rtol=1e-12)
    assert np.isnan(mean[:, :59]).all()


def test_align_matches_searchsorted_and_reindex():
    d = synth.dense(1, 2, 3, 50, seed=1)
    al, idx = O.align(d["ts"], d["val"], d["off"], d["start_ms"], d["step_ms"], 50)
    # phase > 0: grid point j sees sample j-1; first grid point has none
    for s in range(6):
        a, b = d["off"][s], d["off"][s + 1]
        ser = pd.Series(d["val"][a:b], index=d["ts"][a:b])
        grid = d["start_ms"] + d["step_ms"] * np.arange(50)
        exp = ser.reindex(grid, method="ffill", tolerance=300_000 - 1)
        assert_exact_or_nan(al[s], exp.values)
    # lookback is exclusive, stale marker hides the series, plain NaN is a value
    ts = np.array([0, 1000, 2000], dtype=np.int64)
    val = np.array([1.0, np.nan, 3.0]); val.view(np.uint64)[2] = O.STALE_NAN_BITS
    al, idx = O.align(ts, val, np.array([0, 3]), 0, 500, 12, lookback_ms=1500)
    assert idx[0].tolist() == [0, 0, 1, 1, -1, -1, -1, -1, -1, -1, -1, -1]
    assert al[0, 0] == 1.0 and np.isnan(al[0, 2]) and idx[0, 2] == 1


def test_rate_definition_small():
    x = np.array([[10., 20., 5., 7., np.nan, 9., 2., 2.]])
    inc, rt, nr = O.rate(x, 4, 1000)
    # window of 4 ending at j=3: 10,20,5,7 -> (7-10)+20 = 17 over 3 s
    assert inc[0, 3] == 17.0 and rt[0, 3] == 17.0 / 3 and nr[0, 3] == 1
    assert np.isnan(inc[0, 0]) and inc[0, 1] == 10.0
    # j=5: window 5,7,nan,9 -> 4 over 3 s ; j=6: 7,nan,9,2 -> (2-7)+9 = 4
    assert inc[0, 5] == 4.0 and inc[0, 6] == 4.0 and nr[0, 6] == 1


def test_pandas_idiom_pipeline_matches_oracle():
    d = synth.dense(1, 2, 6, 400, seed=5, nan_frac=0.02, all_counters=False, reset_prob=0.0)
    d4 = synth.dense(1, 2, 3, 400, seed=6, all_counters=True, reset_prob=0.02)
    for dd in (d, d4):
        a = O.aggregate_range(dd["ts"], dd["val"], dd["off"], dd["start_ms"], dd["step_ms"], 400, 900_000)
        b = O.aggregate_range_pandas(dd["ts"], dd["val"], dd["off"], dd["start_ms"], dd["step_ms"], 400, 900_000)
        assert_exact_or_nan(b["roll_min"], a["roll_min"]); assert_exact_or_nan(b["roll_max"], a["roll_max"])
        assert_close(b["roll_mean"], a["roll_mean"], scale=O.rolling_scale(a["aligned"], 60), rtol=1e-12)
        sc = O.rate_scale(a["aligned"], 60) / 15.0
        assert_close(b["rate"], a["rate"], scale=sc, rtol=1e-11, what="rate")


def test_group_reduce_matches_pandas_groupby():
    rng = np.random.default_rng(2)
    x = rng.uniform(0, 10, (12, 40)); x[rng.random(x.shape) < 0.1] = np.nan
    gid = np.array([0, 1, 2] * 4)
    sm, cnt, mn, mx, mean = O.group_reduce(x, gid, 3)
    df = pd.DataFrame({"g": np.repeat(gid, 40), "v": x.ravel()})
    agg = df.groupby("g")["v"].agg(["sum", "count", "min", "max", "mean"])
    assert_close(sm, agg["sum"].values, rtol=1e-12); assert np.array_equal(cnt, agg["count"].values)
    assert_exact_or_nan(mn, agg["min"].values); assert_exact_or_nan(mx, agg["max"].values)
    assert_close(mean, agg["mean"].values, rtol=1e-12)
    sm2, cnt2, mn2, mx2, _ = O.group_reduce(x, gid, 3, per_step=True)
    assert_close(sm2[1], np.nansum(x[gid == 1], axis=0), rtol=1e-12)
    assert np.array_equal(cnt2[2], (~np.isnan(x[gid == 2])).sum(axis=0))


def test_quantiles_match_pandas_and_numpy():
    rng = np.random.default_rng(3)
    for n in (1, 2, 3, 10, 1001):
        x = rng.normal(size=n); x[rng.random(n) < 0.1] = np.nan
        if np.isnan(x).all():
            x[0] = 1.0
        q = O.quantiles(x, [0.0, 0.5, 0.95, 0.99, 1.0])
        exp = pd.Series(x).quantile([0.0, 0.5, 0.95, 0.99, 1.0]).values
        assert_bits_equal(q["value"], exp, f"n={n}")
        assert_bits_equal(q["value"], np.percentile(x[~np.isnan(x)], [0, 50, 95, 99, 100]))
    assert O.quantiles([np.nan], [0.5])["n"] == 0
    k = O.f64_to_key(np.array([-np.inf, -1.0, -0.0, 0.0, 1e-300, 1.0, np.inf]))
    assert np.all(np.diff(k.astype(object)) > 0)
    assert_bits_equal(O.key_to_f64(k), np.array([-np.inf, -1.0, -0.0, 0.0, 1e-300, 1.0, np.inf]))


def test_ragged_rolling_matches_pandas_time_window():
    d = synth.ragged(5, seed=1, mean_len=60, nan_frac=0.05, stale_frac=0.02)
    mn, mx, mean, cnt = O.rolling_ragged(d["ts"], d["val"], d["off"], 900_000)
    for s in range(5):
        a, b = d["off"][s], d["off"][s + 1]
        ser = pd.Series(d["val"][a:b], index=pd.to_datetime(d["ts"][a:b], unit="ms"))
        r = ser.rolling("900s")
        assert_exact_or_nan(mn[a:b], r.min().values); assert_exact_or_nan(mx[a:b], r.max().values)
        assert np.array_equal(cnt[a:b], r.count().values.astype(np.int32))
        assert_close(mean[a:b], r.mean().values, rtol=1e-12)


def test_extrapolated_rate_matches_prometheus_examples():
    """Hand-checked extrapolatedRate cases (promql/functions.go): dense counter, counter that starts
    mid-window (zero-crossing cap), and a window edge further than 1.1 spacings away."""
    # dense: 60 points 15 s apart over a 900 s window: first is 15 s after the window start -> full stretch
    x = np.arange(100, dtype=np.float64)[None, :] * 30.0 + 1000.0
    inc, rt, _ = O.rate(x, 60, 15000, mode=1)
    j = 80
    raw = 59 * 30.0
    assert np.isclose(inc[0, j], raw * 900.0 / 885.0, rtol=1e-15) and np.isclose(rt[0, j], inc[0, j] / 900.0, rtol=1e-15)
    # counter born inside the window at a small value: start extrapolation capped at the zero crossing
    y = np.full((1, 100), np.nan); y[0, 50:] = 3.0 + 30.0 * np.arange(50)
    inc, rt, _ = O.rate(y, 60, 15000, mode=1)
    j = 60                                              # samples at 50..60, first value 3, raw = 300 over 150 s
    to_zero = 150.0 * (3.0 / 300.0)
    assert np.isclose(inc[0, j], 300.0 * (150.0 + to_zero) / 150.0, rtol=1e-15)
    # mode 0 unchanged
    inc0, rt0, _ = O.rate(y, 60, 15000)
    assert inc0[0, j] == 300.0 and rt0[0, j] == 2.0
