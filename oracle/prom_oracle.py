"""CPU oracle for the Prometheus aggregate hot path.  TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this module.  The product package
(``k8s_rocm_metrics_dashboard_b200``) never does: it fails loudly when ``libpmx.so``
is missing instead of falling back to anything in here.

What it restates, and from where
--------------------------------
* H1-H10 (the stages that exist in the reference): ``aggregate_instant`` follows
  ``/root/reference/app.py:180-221`` (records -> ``pd.DataFrame`` -> ``pivot`` -> VRAM
  ratio -> ``mean/max/min``) and ``selected_averages`` follows ``app.py:335-345``.
  Pinned: ``tests/golden/instant_*.json`` were produced by running the UNMODIFIED
  reference under import stubs (``oracle/ref_app.py`` + ``tests/golden/make_golden.py``).
* X1-X7 (stages BASELINE.json's north_star names that the reference does not
  implement): pandas/numpy expressions fixed in SURVEY.md section 8(a2).  **Parity
  unpinned** for these rows: the reference holds no implementation, test or golden
  vector for them; the restatement is validated against pandas' own ``rolling`` /
  ``quantile`` / ``groupby`` in ``tests/test_oracle.py``.

Conventions: ``ts`` is int64 milliseconds, values are float64, ragged series are CSR
(``off`` int64[S+1] into flat ``ts``/``val``).  ``STALE_NAN_BITS`` is Prometheus' staleness
marker; an ordinary NaN sample is a valid sample whose value is NaN.
"""
from __future__ import annotations

import json
import math
from decimal import Decimal

import numpy as np
import pandas as pd

STALE_NAN_BITS = 0x7FF0000000000002
CANON_NAN_BITS = 0x7FF8000000000000
DEFAULT_LOOKBACK_MS = 300_000


# --------------------------------------------------------------------------- X1/X2
def parse_value(token: str) -> float:
    """app.py:188 -- ``float(metric["value"][1])``; raises ValueError like the reference."""
    return float(token)


def parse_ts_ms(token: str) -> int:
    """Timestamp token (JSON number, seconds, <=3 decimals) -> exact int64 ms."""
    d = Decimal(token) * 1000
    if d != d.to_integral_value():
        raise ValueError(f"timestamp {token!r} has more than 3 decimals")
    return int(d)


def tokenize_matrix(body: str | bytes):
    """json.loads + per-sample float(), the way app.py:178/183-189 treats a vector.

    Returns (labels list[dict], ts int64[N], val float64[N], off int64[S+1]).
    Accepts both resultType "matrix" ("values") and "vector" ("value").
    """
    doc = json.loads(body)
    res = doc["data"]["result"]
    labels, ts, val, off = [], [], [], [0]
    for series in res:
        labels.append(series["metric"])
        pairs = series["values"] if "values" in series else [series["value"]]
        for t, v in pairs:
            ts.append(int(round(t * 1000)))
            val.append(float(v))
        off.append(len(ts))
    return (labels, np.asarray(ts, dtype=np.int64), np.asarray(val, dtype=np.float64),
            np.asarray(off, dtype=np.int64))


def raw_tokens(body: str | bytes):
    """The raw timestamp / value token TEXT per sample (for tokenizer-offset parity)."""
    doc = json.loads(body, parse_float=lambda s: s, parse_int=lambda s: s)
    ts_tok, val_tok, off = [], [], [0]
    for series in doc["data"]["result"]:
        pairs = series["values"] if "values" in series else [series["value"]]
        for t, v in pairs:
            ts_tok.append(t)
            val_tok.append(v)
        off.append(len(ts_tok))
    return ts_tok, val_tok, off


# --------------------------------------------------------------------------- X3
def align(ts, val, off, start_ms, step_ms, n_steps, lookback_ms=DEFAULT_LOOKBACK_MS):
    """Step-aligned resample, SURVEY X3 (searchsorted form).

    idx_j = searchsorted(ts, g_j, 'right') - 1; valid iff idx_j >= 0,
    g_j - ts[idx_j] < lookback and val[idx_j] is not the stale marker.
    Returns (aligned float64[S, n_steps], src_idx int32[S, n_steps]); invalid -> NaN / -1.
    src_idx is relative to the series start.
    """
    S = len(off) - 1
    grid = start_ms + step_ms * np.arange(n_steps, dtype=np.int64)
    out = np.full((S, n_steps), np.nan)
    idx_out = np.full((S, n_steps), -1, dtype=np.int32)
    bits = val.view(np.uint64)
    for s in range(S):
        a, b = int(off[s]), int(off[s + 1])
        if a == b:
            continue
        t = ts[a:b]
        idx = np.searchsorted(t, grid, side="right") - 1
        ok = idx >= 0
        safe = np.where(ok, idx, 0)
        ok &= (grid - t[safe]) < lookback_ms
        ok &= bits[a:b][safe] != np.uint64(STALE_NAN_BITS)
        out[s] = np.where(ok, val[a:b][safe], np.nan)
        # pure gather: keep the payload of an ordinary NaN sample
        ob = out[s].view(np.uint64)
        ob[ok] = bits[a:b][safe][ok]
        idx_out[s] = np.where(ok, idx, -1)
    return out, idx_out


def bucket(ts, val, off, start_ms, step_ms, n_steps):
    """Bucket down-sampling, the second form of SURVEY X3: ``Series.resample('<step>')`` bins
    ``[start + j*step, start + (j+1)*step)`` labelled left; per bin the LAST non-NaN sample, the MEAN
    of the non-NaN samples (summed in time order) and their COUNT.  A stale marker is a NaN like any
    other here (pandas knows no markers).  Empty bin: NaN / NaN / 0.
    Returns (last float64[S, n_steps], mean float64[S, n_steps], count int32[S, n_steps]).
    """
    S = len(off) - 1
    last = np.full((S, n_steps), np.nan); mean = np.full((S, n_steps), np.nan)
    cnt = np.zeros((S, n_steps), dtype=np.int32)
    for s in range(S):
        a, b = int(off[s]), int(off[s + 1])
        t, v = ts[a:b], val[a:b]
        j = (t - start_ms) // step_ms
        ok = (j >= 0) & (j < n_steps) & ~np.isnan(v)
        sums = np.zeros(n_steps)
        for jj, vv in zip(j[ok].tolist(), v[ok].tolist()):
            sums[jj] += vv
            cnt[s, jj] += 1
            last[s, jj] = vv
        with np.errstate(invalid="ignore", divide="ignore"):
            mean[s] = np.where(cnt[s] > 0, sums / cnt[s], np.nan)
    return last, mean, cnt


# --------------------------------------------------------------------------- X4
def _windows(x, w):
    """[S, T] -> [S, T, w] trailing windows (t-w, t], left-padded with NaN."""
    S, T = x.shape
    pad = np.full((S, w - 1), np.nan)
    xp = np.concatenate([pad, x], axis=1)
    return np.lib.stride_tricks.sliding_window_view(xp, w, axis=1)


def rolling(x, w, min_periods=1):
    """Rolling min/max/mean/count over the trailing ``w`` grid points, SURVEY X4.

    Semantics are pandas' ``Series.rolling('<w*step>')`` on a regular DatetimeIndex
    (closed='right', min_periods=1) and ``rolling(w)`` (min_periods=w):
      * count = number of non-NaN observations (``+-inf`` counts);
      * min/max/mean use FINITE observations only -- pandas turns +-inf into NaN before
        the window kernels (pandas/core/window/rolling.py ``_prep_values``);
      * a stat is NaN when fewer than ``min_periods`` finite observations are present.
    The mean is computed freshly per window (pairwise sum), not with pandas' online
    add/remove accumulator.
    Returns (mn, mx, mean float64[S,T], count int32[S,T]).
    """
    x = np.asarray(x, dtype=np.float64)
    S, T = x.shape
    mn = np.empty((S, T)); mx = np.empty((S, T)); mean = np.empty((S, T))
    cnt = np.empty((S, T), dtype=np.int32)
    chunk = max(1, int(4e7 // max(1, T * w)))
    for s0 in range(0, S, chunk):
        win = _windows(x[s0:s0 + chunk], w)
        fin = np.isfinite(win)
        nfin = fin.sum(axis=2)
        cnt[s0:s0 + chunk] = (~np.isnan(win)).sum(axis=2)
        ok = nfin >= max(1, min_periods)
        mn[s0:s0 + chunk] = np.where(ok, np.where(fin, win, np.inf).min(axis=2), np.nan)
        mx[s0:s0 + chunk] = np.where(ok, np.where(fin, win, -np.inf).max(axis=2), np.nan)
        with np.errstate(invalid="ignore", divide="ignore"):
            mean[s0:s0 + chunk] = np.where(ok, np.where(fin, win, 0.0).sum(axis=2) / nfin, np.nan)
    return mn, mx, mean, cnt


def rolling_scale(x, w):
    """mean |x| over the finite observations of each window: the tolerance scale for means."""
    win = _windows(np.abs(np.asarray(x, dtype=np.float64)), w)
    fin = np.isfinite(win)
    n = fin.sum(axis=2)
    with np.errstate(invalid="ignore", divide="ignore"):
        return np.where(n > 0, np.where(fin, win, 0.0).sum(axis=2) / n, 0.0)


def rolling_ragged(ts, val, off, window_ms, min_periods=1):
    """Time-window rolling stats evaluated AT EACH SAMPLE of a ragged series (config 5).

    Window (t_i - W, t_i] over the samples of the same series; same NaN/inf rules as
    ``rolling``.  Stale markers are NaN bit patterns and are skipped like any NaN.
    Returns flat (mn, mx, mean, count).
    """
    N = len(ts)
    mn = np.full(N, np.nan); mx = np.full(N, np.nan); mean = np.full(N, np.nan)
    cnt = np.zeros(N, dtype=np.int32)
    for s in range(len(off) - 1):
        a, b = int(off[s]), int(off[s + 1])
        t = ts[a:b]; v = val[a:b]
        lo = np.searchsorted(t, t - window_ms, side="right")
        for i in range(b - a):
            wv = v[lo[i]:i + 1]
            cnt[a + i] = np.count_nonzero(~np.isnan(wv))
            f = wv[np.isfinite(wv)]
            if len(f) >= max(1, min_periods):
                mn[a + i] = f.min(); mx[a + i] = f.max(); mean[a + i] = math.fsum(f) / len(f)
    return mn, mx, mean, cnt


# --------------------------------------------------------------------------- X5
def extrapolate(inc, first_val, n, sampled_s, to_start_s, to_end_s, range_s):
    """Prometheus extrapolatedRate for counters (promql/functions.go), SURVEY X5 mode 1.
    Returns (extrapolated increase, rate)."""
    avg = sampled_s / (n - 1)
    if inc > 0 and first_val >= 0:
        to_zero = sampled_s * (first_val / inc)
        if to_zero < to_start_s:
            to_start_s = to_zero
    thr = avg * 1.1
    ext = sampled_s
    if to_start_s >= thr:
        to_start_s = avg / 2.0
    ext += to_start_s
    if to_end_s >= thr:
        to_end_s = avg / 2.0
    ext += to_end_s
    factor = ext / sampled_s
    return inc * factor, (inc * factor) / range_s


def _rate_series(k, x, t_of, lo_of, T_out, mode=0, win_end_ms=None, range_ms=None):
    """Shared body: k = positions of non-NaN samples, x their values.

    For output j the window holds the non-NaN samples with position in [lo_of(j), j].
    increase = (x_last - x_first) + sum_{resets m>first} x_{m-1}   (SURVEY X5), evaluated exactly and
    rounded once
    mode 0: rate = increase / ((t_last - t_first)/1000)
    mode 1: Prometheus extrapolation over the window (win_end_ms(j) - range_ms, win_end_ms(j)]
    """
    inc = np.full(T_out, np.nan); rate = np.full(T_out, np.nan)
    nres = np.zeros(T_out, dtype=np.int32)
    if len(k) == 0:
        return inc, rate, nres
    reset = np.zeros(len(k), dtype=bool)
    reset[1:] = x[1:] < x[:-1]
    rterm = np.zeros(len(k))
    rterm[1:] = np.where(reset[1:], x[:-1], 0.0)
    for j in range(T_out):
        p = np.searchsorted(k, lo_of(j), side="left")
        q = np.searchsorted(k, j, side="right") - 1
        if q - p + 1 < 2:
            continue
        r = math.fsum(rterm[p + 1:q + 1])
        nres[j] = int(reset[p + 1:q + 1].sum())
        with np.errstate(invalid="ignore", over="ignore", divide="ignore"):
            # the formula's value rounded ONCE (after a reset "x_last - x_first" alone is a huge
            # negative number whose rounding error would be the oracle's, not the kernel's)
            if math.isfinite(x[q]) and math.isfinite(x[p]) and math.isfinite(r):
                inc[j] = math.fsum((x[q], -x[p], *rterm[p + 1:q + 1]))
            else:
                inc[j] = (x[q] - x[p]) + r
            dt = (t_of(k[q]) - t_of(k[p])) / 1000.0
            rate[j] = inc[j] / dt if dt != 0 else np.nan
            if mode == 1 and dt != 0:
                end = win_end_ms(j)
                inc[j], rate[j] = extrapolate(inc[j], x[p], q - p + 1, dt, (t_of(k[p]) - (end - range_ms)) / 1000.0,
                                              (end - t_of(k[q])) / 1000.0, range_ms / 1000.0)
    return inc, rate, nres


def rate(x, w, step_ms, mode=0):
    """Counter increase / rate with reset handling over trailing ``w`` grid points.

    ``x`` is the aligned [S, T] matrix (NaN = absent).  NaN samples are skipped; +-inf are
    ordinary values here.  Fewer than 2 samples in the window -> NaN.  mode 1 = Prometheus
    extrapolation (``extrapolate``).  Returns (increase, rate float64[S,T], n_resets int32[S,T]).
    """
    x = np.asarray(x, dtype=np.float64)
    S, T = x.shape
    inc = np.empty((S, T)); rt = np.empty((S, T)); nr = np.empty((S, T), dtype=np.int32)
    for s in range(S):
        k = np.flatnonzero(~np.isnan(x[s]))
        inc[s], rt[s], nr[s] = _rate_series(k, x[s][k], lambda p: p * step_ms, lambda j: j - w + 1, T,
                                            mode, lambda j: j * step_ms, w * step_ms)
    return inc, rt, nr


def rate_ragged(ts, val, off, window_ms, mode=0):
    """``rate`` evaluated at each sample of a ragged series with window (t_i - W, t_i]."""
    N = len(ts)
    inc = np.full(N, np.nan); rt = np.full(N, np.nan); nr = np.zeros(N, dtype=np.int32)
    for s in range(len(off) - 1):
        a, b = int(off[s]), int(off[s + 1])
        t = ts[a:b]; v = val[a:b]
        lo = np.searchsorted(t, t - window_ms, side="right")
        k = np.flatnonzero(~np.isnan(v))
        i_, r_, n_ = _rate_series(k, v[k], lambda p: int(t[p]), lambda j: int(lo[j]), b - a,
                                  mode, lambda j: int(t[j]), window_ms)
        inc[a:b], rt[a:b], nr[a:b] = i_, r_, n_
    return inc, rt, nr


def rate_scale(x, w):
    """max |x| over each window (finite only): tolerance scale for increase."""
    win = _windows(np.abs(np.asarray(x, dtype=np.float64)), w)
    return np.where(np.isfinite(win), win, 0.0).max(axis=2)


# --------------------------------------------------------------------------- X6
def group_reduce(x, group_id, n_groups, per_step=False):
    """NaN-skipping sum/count/min/max/mean per group, SURVEY X6.

    Equivalent to ``df.groupby(keys)['value'].agg(['sum','count','min','max','mean'])``
    with groups numbered in sorted key order.  per_step=False reduces over (series in
    group, all steps) -> [G]; per_step=True reduces over series only -> [G, T].
    Series are visited in index order, steps in time order (pandas' row order).
    Empty group: sum 0, count 0, min/max/mean NaN.
    """
    x = np.asarray(x, dtype=np.float64)
    S, T = x.shape
    shape = (n_groups, T) if per_step else (n_groups,)
    sm = np.zeros(shape); cnt = np.zeros(shape, dtype=np.int64)
    mn = np.full(shape, np.inf); mx = np.full(shape, -np.inf)
    for s in range(S):
        g = int(group_id[s])
        if g < 0:
            continue
        ok = ~np.isnan(x[s])
        if per_step:
            sm[g] += np.where(ok, x[s], 0.0); cnt[g] += ok
            mn[g] = np.minimum(mn[g], np.where(ok, x[s], np.inf))
            mx[g] = np.maximum(mx[g], np.where(ok, x[s], -np.inf))
        else:
            v = x[s][ok]
            if len(v):
                sm[g] += math.fsum(v) if np.all(np.isfinite(v)) else v.sum()
                cnt[g] += len(v)
                mn[g] = min(mn[g], v.min()); mx[g] = max(mx[g], v.max())
    with np.errstate(invalid="ignore", divide="ignore"):
        mean = np.where(cnt > 0, sm / cnt, np.nan)
    mn = np.where(cnt > 0, mn, np.nan); mx = np.where(cnt > 0, mx, np.nan)
    return sm, cnt, mn, mx, mean


# --------------------------------------------------------------------------- X7
def lerp(a, b, t):
    """numpy's ``_lerp`` (numpy/lib/_function_base_impl.py): a+(b-a)t, b-(b-a)(1-t) for t>=.5."""
    d = b - a
    out = a + d * t
    if t >= 0.5:
        out = b - d * (1 - t)
    if d == 0:
        out = a
    return out


def quantiles(x, qs):
    """Series.quantile(qs) == np.percentile(x[~isnan], method='linear'), SURVEY X7.

    Returns dict with n, and per q: lo, hi (ranks), x_lo, x_hi (order statistics), value.
    n == 0 -> everything NaN / -1.
    """
    v = np.sort(np.asarray(x, dtype=np.float64).ravel())
    v = v[~np.isnan(v)]
    n = len(v)
    out = {"n": n, "lo": [], "hi": [], "x_lo": [], "x_hi": [], "value": []}
    for q in qs:
        if n == 0:
            out["lo"].append(-1); out["hi"].append(-1)
            out["x_lo"].append(np.nan); out["x_hi"].append(np.nan); out["value"].append(np.nan)
            continue
        pos = (n - 1) * q
        lo = int(math.floor(pos)); hi = min(lo + 1, n - 1); t = pos - lo
        with np.errstate(invalid="ignore"):
            val = lerp(v[lo], v[hi], t)
        out["lo"].append(lo); out["hi"].append(hi)
        out["x_lo"].append(float(v[lo])); out["x_hi"].append(float(v[hi])); out["value"].append(float(val))
    return out


def f64_to_key(x):
    """Order-preserving float64 -> uint64 key (sign flip trick); NaN keys are not used."""
    b = np.asarray(x, dtype=np.float64).view(np.uint64)
    neg = (b >> np.uint64(63)).astype(bool)
    return np.where(neg, ~b, b | np.uint64(1 << 63))


def key_to_f64(k):
    k = np.asarray(k, dtype=np.uint64)
    pos = (k >> np.uint64(63)).astype(bool)
    return np.where(pos, k & np.uint64((1 << 63) - 1), ~k).view(np.float64)


# --------------------------------------------------------------------------- H1-H10
def aggregate_instant(result):
    """app.py:180-221 restated on an instant-vector ``data.result`` list.

    Returns (df_pivot, stats) exactly as ``fetch_gpu_metrics`` does (object dtype cells,
    lexicographic gpu_id index, ``vram_usage_ratio`` column, mean/max/min Series).
    Exceptions propagate (the reference's caller turns them into ``(None, None)``).
    """
    models, records = {}, []
    for m in result:                                        # app.py:183-194
        lab = m["metric"]
        records.append({"gpu_id": lab["gpu_id"], "metric_name": lab["__name__"],
                        "value": float(m["value"][1])})
        if lab["gpu_id"] not in models:
            models[lab["gpu_id"]] = lab["card_model"]
    for gpu_id, model in models.items():                    # app.py:196-201
        records.append({"gpu_id": gpu_id, "metric_name": "card_model", "value": model})
    df = pd.DataFrame(records)                              # app.py:204
    piv = df.pivot(index="gpu_id", columns="metric_name", values="value")   # app.py:207
    piv["vram_usage_ratio"] = (piv["amd_gpu_used_vram"] / piv["amd_gpu_total_vram"]) * 100
    cols = [c for c in piv.columns if c != "card_model"]    # app.py:216
    stats = {"mean": piv[cols].mean(), "max": piv[cols].max(), "min": piv[cols].min()}
    return piv, stats


def selected_averages(df_pivot, selected):
    """app.py:335-345: means over the selected cards, power mean over strictly positive readings."""
    f = df_pivot.loc[selected]
    cols = [c for c in f.columns if c != "card_model"]
    avg = f[cols].mean()
    p = f["amd_gpu_average_package_power"]
    nz = p[p > 0]
    if not nz.empty:
        avg["amd_gpu_average_package_power"] = nz.mean()
    return avg


# --------------------------------------------------------------------------- pipeline
def aggregate_range(ts, val, off, start_ms, step_ms, n_steps, window_ms,
                    lookback_ms=DEFAULT_LOOKBACK_MS):
    """The range path end to end on decoded arrays (align -> rolling -> rate)."""
    w = int(window_ms // step_ms)
    al, idx = align(ts, val, off, start_ms, step_ms, n_steps, lookback_ms)
    mn, mx, mean, cnt = rolling(al, w)
    inc, rt, nres = rate(al, w, step_ms)
    return {"aligned": al, "src_idx": idx, "roll_min": mn, "roll_max": mx, "roll_mean": mean,
            "roll_count": cnt, "increase": inc, "rate": rt, "n_resets": nres}


def aggregate_range_pandas(ts, val, off, start_ms, step_ms, n_steps, window_ms,
                           lookback_ms=DEFAULT_LOOKBACK_MS):
    """Same path in the idiom the reference uses (pandas frames); this is what the CPU
    baseline times.  One wide frame [n_steps x S]; rolling() runs column-wise in C.
    Returns dict of [S, n_steps] arrays (aligned, roll_min, roll_max, roll_mean, rate).
    """
    w = int(window_ms // step_ms)
    al, _ = align(ts, val, off, start_ms, step_ms, n_steps, lookback_ms)
    index = pd.to_datetime(start_ms + step_ms * np.arange(n_steps, dtype=np.int64), unit="ms")
    df = pd.DataFrame(al.T, index=index)
    r = df.rolling(f"{window_ms}ms")
    mn, mx, mean = r.min(), r.max(), r.mean()
    # rate in pandas idiom: d = diff with reset correction, windowed sum of d
    # over the present samples minus the first present sample's own d (telescoping form of
    # (x_last - x_first) + sum of pre-reset values), divided by the elapsed seconds.
    present = df.notna()
    v = df.ffill()
    d = v.diff()
    d = d.where(d >= 0, v)            # a drop (or the very first sample) contributes v itself
    dm = d.where(present)
    tot = dm.fillna(0.0).rolling(w, min_periods=1).sum()
    dfb = dm.bfill()
    d_first = dfb.shift(w - 1).fillna(dfb.iloc[0])
    sec = pd.DataFrame(np.broadcast_to(np.arange(n_steps)[:, None] * (step_ms / 1000.0), df.shape),
                       index=index, columns=df.columns).where(present)
    t_first = sec.rolling(w, min_periods=1).min()
    t_last = sec.rolling(w, min_periods=1).max()
    n = present.astype(np.float64).rolling(w, min_periods=1).sum()
    rate_ = ((tot - d_first) / (t_last - t_first)).where(n >= 2)
    return {"aligned": al, "roll_min": mn.values.T, "roll_max": mx.values.T,
            "roll_mean": mean.values.T, "rate": rate_.values.T}


def aggregate_step_pandas(ts, val, off, start_ms, step_ms, n_steps, window_ms, group_id, n_groups,
                          qs=(0.5, 0.95, 0.99), lookback_ms=DEFAULT_LOOKBACK_MS):
    """One whole aggregate step the way the reference would write it with pandas: the range
    pipeline, then per-group sum/count/min/max/mean and quantiles of the aligned values.
    This is what bench.py's cpu_baseline / --impl reference legs time."""
    r = aggregate_range_pandas(ts, val, off, start_ms, step_ms, n_steps, window_ms, lookback_ms)
    al = r["aligned"]
    S = al.shape[0]
    long = pd.DataFrame({"g": np.repeat(np.asarray(group_id), n_steps), "v": al.ravel()})
    gb = long[long["g"] >= 0].groupby("g")["v"]
    r["summary"] = gb.agg(["sum", "count", "min", "max", "mean"]).reindex(range(n_groups))
    r["quantiles"] = gb.quantile(list(qs)).unstack().reindex(range(n_groups))
    del S
    return r
