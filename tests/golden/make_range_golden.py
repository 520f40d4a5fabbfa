"""Generate ``range_golden.npz``: pinned vectors for the range stages X3-X7.

    python tests/golden/make_range_golden.py

The reference holds no implementation, test or vector for these stages (its whole numeric surface
is app.py:216-221 and :335-345), so the only pin available is the library the reference is written
in: every expected array below is produced by pandas / numpy THEMSELVES --
``Series.reindex(method='ffill', tolerance=)``, ``Series.resample().last()/.mean()``,
``rolling('<W>ms')``, ``rolling(n)``, ``groupby().agg()``, ``Series.quantile`` /
``np.percentile`` / ``np.sort`` -- and nothing here imports ``oracle/prom_oracle.py``.  The one
stage pandas has no primitive for, the counter increase with resets (X5), is evaluated window by
window straight from SURVEY X5's sentence (last - first + the values before each drop) with plain
numpy indexing.  Both the oracle (tests/test_range_golden.py) and the CUDA path
(tests/test_gpu_range_golden.py) are checked against this file; the versions that produced it are
stored inside, and a pandas upgrade that changes a result shows up as a failing CPU test.

Inputs are seeded and shaped like BASELINE.json's configs, scaled to fixture size, and stored in
the file so that the tests do not depend on the generator's random streams:
  c1  config 1 (1 node x 8 cards x 6 metrics, 1 h @ 15 s)
  c2  config 2's series mix (gauge / gauge / counter, 0.1 % NaN), 12 series x 6 h
  c4  config 4 (all counters, resets), 12 series x 6 h
  c5  config 5 (ragged, irregular timestamps, gaps, NaN and stale markers), 40 series
"""
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

import numpy as np                                             # noqa: E402
import pandas as pd                                            # noqa: E402

from k8s_rocm_metrics_dashboard_b200 import synth              # noqa: E402

STALE = np.uint64(0x7FF0000000000002)
LOOKBACK = 300_000
QS = [0.5, 0.95, 0.99]


def align_pandas(ts, val, off, start, step, n_steps, lookback):
    """X3: last sample at or before each grid point, not older than lookback (exclusive), stale
    markers hide the series.  ``reindex(ffill, tolerance)`` is inclusive, hence lookback - 1 ms."""
    S = len(off) - 1
    grid = start + step * np.arange(n_steps, dtype=np.int64)
    out = np.full((S, n_steps), np.nan)
    idx_out = np.full((S, n_steps), -1, dtype=np.int32)
    bits = val.view(np.uint64)
    for s in range(S):
        a, b = int(off[s]), int(off[s + 1])
        if a == b:
            continue
        pos = pd.Series(np.arange(b - a, dtype=np.float64), index=ts[a:b]).reindex(grid, method="ffill",
                                                                                   tolerance=lookback - 1)
        idx = pos.fillna(-1).to_numpy().astype(np.int64)
        ok = idx >= 0
        ok &= bits[a:b][np.where(ok, idx, 0)] != STALE
        idx = np.where(ok, idx, -1)
        idx_out[s] = idx
        ob = out[s].view(np.uint64)
        ob[ok] = bits[a:b][idx[ok]]                             # a gather: NaN payloads survive
    return out, idx_out


def bucket_pandas(ts, val, off, start, step, n_steps):
    """X3, down-sampling variant: ``resample`` bins [t, t+step) labelled left; last / mean / count
    of the non-NaN samples of each bin (a stale marker is a NaN to pandas)."""
    S = len(off) - 1
    grid = pd.to_datetime(start + step * np.arange(n_steps, dtype=np.int64), unit="ms")
    last = np.full((S, n_steps), np.nan); mean = np.full((S, n_steps), np.nan)
    cnt = np.zeros((S, n_steps), dtype=np.int32)
    for s in range(S):
        a, b = int(off[s]), int(off[s + 1])
        keep = (ts[a:b] >= start) & (ts[a:b] < start + step * n_steps)
        if not keep.any():
            continue
        ser = pd.Series(val[a:b][keep], index=pd.to_datetime(ts[a:b][keep], unit="ms"))
        r = ser.resample(f"{step}ms", origin=pd.Timestamp(start, unit="ms"), closed="left", label="left")
        last[s] = r.last().reindex(grid).to_numpy()
        mean[s] = r.mean().reindex(grid).to_numpy()
        cnt[s] = r.count().reindex(grid).fillna(0).to_numpy().astype(np.int32)
    return last, mean, cnt


def rolling_pandas(x, w, step, start, fixed=False):
    """X4: ``rolling('<w*step>ms')`` on a DatetimeIndex (closed right, min_periods 1), or
    ``rolling(w)`` (min_periods w)."""
    idx = pd.to_datetime(start + step * np.arange(x.shape[1], dtype=np.int64), unit="ms")
    df = pd.DataFrame(x.T, index=idx)
    r = df.rolling(w) if fixed else df.rolling(f"{w * step}ms")
    return (r.min().to_numpy().T, r.max().to_numpy().T, r.mean().to_numpy().T,
            r.count().to_numpy().T.astype(np.int32))


def rolling_ragged_pandas(ts, val, off, window_ms):
    N = len(ts)
    mn = np.full(N, np.nan); mx = np.full(N, np.nan); mean = np.full(N, np.nan); cnt = np.zeros(N, dtype=np.int32)
    for s in range(len(off) - 1):
        a, b = int(off[s]), int(off[s + 1])
        if a == b:
            continue
        r = pd.Series(val[a:b], index=pd.to_datetime(ts[a:b], unit="ms")).rolling(f"{window_ms}ms")
        mn[a:b], mx[a:b], mean[a:b] = r.min().to_numpy(), r.max().to_numpy(), r.mean().to_numpy()
        cnt[a:b] = r.count().to_numpy().astype(np.int32)
    return mn, mx, mean, cnt


def rate_window(v, t_ms):
    """SURVEY X5 for one window: v = its non-NaN samples in time order, t_ms their times (int64 ms;
    the elapsed time is formed in integers, then turned into seconds)."""
    if len(v) < 2:
        return np.nan, np.nan, 0
    drop = v[1:] < v[:-1]
    inc = (v[-1] - v[0]) + float(np.sum(v[:-1][drop]))
    dt = int(t_ms[-1] - t_ms[0]) / 1000.0
    return inc, (inc / dt if dt != 0 else np.nan), int(drop.sum())


def rate_dense(x, w, step):
    S, T = x.shape
    inc = np.full((S, T), np.nan); rt = np.full((S, T), np.nan); nr = np.zeros((S, T), dtype=np.int32)
    ms = np.arange(T, dtype=np.int64) * int(step)
    for s in range(S):
        for j in range(T):
            lo = max(0, j - w + 1)
            ok = ~np.isnan(x[s, lo:j + 1])
            inc[s, j], rt[s, j], nr[s, j] = rate_window(x[s, lo:j + 1][ok], ms[lo:j + 1][ok])
    return inc, rt, nr


def rate_ragged(ts, val, off, window_ms):
    N = len(ts)
    inc = np.full(N, np.nan); rt = np.full(N, np.nan); nr = np.zeros(N, dtype=np.int32)
    for s in range(len(off) - 1):
        a, b = int(off[s]), int(off[s + 1])
        t, v = ts[a:b], val[a:b]
        for i in range(b - a):
            sel = (t > t[i] - window_ms) & (t <= t[i]) & ~np.isnan(v)
            inc[a + i], rt[a + i], nr[a + i] = rate_window(v[sel], t[sel])
    return inc, rt, nr


def groupby_pandas(x, labels, keys):
    """X6: ``df.groupby(keys)['value'].agg([...])``, groups in sorted key order."""
    S, T = x.shape
    cols = {k: np.repeat([l[k] for l in labels], T) for k in keys}
    df = pd.DataFrame({**cols, "value": x.ravel()})
    g = df.groupby(list(keys), sort=True)["value"].agg(["sum", "count", "min", "max", "mean"])
    names = ["\x1f".join(k) if isinstance(k, tuple) else k for k in g.index]
    return names, g


def quantiles_pandas(x, gid, n_groups):
    """X7: per group, n / lo / hi / x[lo] / x[hi] from np.sort, value from Series.quantile (checked
    here against np.percentile(method='linear'))."""
    out = {k: [] for k in ("n", "lo", "hi", "x_lo", "x_hi", "value")}
    for g in range(n_groups):
        v = x[gid == g].ravel()
        v = np.sort(v[~np.isnan(v)])
        n = len(v)
        val = pd.Series(v).quantile(QS).to_numpy() if n else np.full(len(QS), np.nan)
        if n:
            assert np.array_equal(val.view(np.uint64), np.percentile(v, [100 * q for q in QS]).view(np.uint64))
        pos = [(n - 1) * q for q in QS]
        lo = [int(np.floor(p)) if n else -1 for p in pos]
        hi = [min(l + 1, n - 1) if n else -1 for l in lo]
        out["n"].append(n); out["lo"].append(lo); out["hi"].append(hi)
        out["x_lo"].append([v[l] if n else np.nan for l in lo]); out["x_hi"].append([v[h] if n else np.nan for h in hi])
        out["value"].append(val)
    return {k: np.asarray(v) for k, v in out.items()}


def dense_case(G, name, d, w, group_keys):
    ts, val, off = d["ts"], d["val"], d["off"]
    start, step, T = d["start_ms"], d["step_ms"], d["n_steps"]
    S = len(off) - 1
    G[f"{name}/ts"], G[f"{name}/val"], G[f"{name}/off"] = ts, val, off
    G[f"{name}/params"] = np.array([start, step, T, w, LOOKBACK], dtype=np.int64)
    G[f"{name}/labels"] = np.array([json_labels(l) for l in d["labels"]])
    al, idx = align_pandas(ts, val, off, start, step, T, LOOKBACK)
    G[f"{name}/aligned"], G[f"{name}/src_idx"] = al, idx
    mn, mx, mean, cnt = rolling_pandas(al, w, step, start)
    G[f"{name}/roll_min"], G[f"{name}/roll_max"], G[f"{name}/roll_mean"], G[f"{name}/roll_count"] = mn, mx, mean, cnt
    mn, mx, mean, _ = rolling_pandas(al, w, step, start, fixed=True)
    G[f"{name}/fix_min"], G[f"{name}/fix_max"], G[f"{name}/fix_mean"] = mn, mx, mean
    inc, rt, nr = rate_dense(al, w, step)
    G[f"{name}/increase"], G[f"{name}/rate"], G[f"{name}/n_resets"] = inc, rt, nr
    for tag, keys in group_keys.items():
        names, g = groupby_pandas(al, d["labels"], keys)
        G[f"{name}/group_{tag}/names"] = np.array(names)
        for col in g.columns:
            G[f"{name}/group_{tag}/{col}"] = g[col].to_numpy()
    mnames = sorted({l["__name__"] for l in d["labels"]})
    gid = np.array([mnames.index(l["__name__"]) for l in d["labels"]], dtype=np.int32)
    G[f"{name}/metric_id"] = gid
    for k, v in quantiles_pandas(al, gid, len(mnames)).items():
        G[f"{name}/quantile/{k}"] = v
    assert S == al.shape[0]


def json_labels(l):
    import json
    return json.dumps(l, sort_keys=True, separators=(",", ":"))


def main():
    G = {}
    G["versions"] = np.array([pd.__version__, np.__version__, sys.version.split()[0]])
    keys = {"card": ("instance", "gpu_id", "__name__"), "metric": ("__name__",)}
    dense_case(G, "c1", synth.dense(1, 8, 6, 240, seed=11), 60, keys)
    dense_case(G, "c2", synth.dense(1, 1, 12, 1440, seed=12, nan_frac=1e-3, kinds="ttk"), 60, keys)
    d4 = synth.dense(1, 1, 12, 1440, seed=14, all_counters=True, reset_prob=5e-3)
    d4["val"][np.random.default_rng(4).random(len(d4["val"])) < 2e-3] = np.nan
    dense_case(G, "c4", d4, 60, keys)
    # config 5: ragged series, every stage evaluated at the samples themselves and on a 15 s grid
    d5 = synth.ragged(40, seed=15, mean_len=150, nan_frac=0.01, stale_frac=0.005, gap_frac=0.02)
    ts, val, off = d5["ts"], d5["val"], d5["off"]
    G["c5/ts"], G["c5/val"], G["c5/off"] = ts, val, off
    W = 900_000
    t0 = int(ts.min()) - 45_000
    n_steps = 800
    G["c5/params"] = np.array([t0, 15_000, n_steps, 60, LOOKBACK, W], dtype=np.int64)
    mn, mx, mean, cnt = rolling_ragged_pandas(ts, val, off, W)
    G["c5/roll_min"], G["c5/roll_max"], G["c5/roll_mean"], G["c5/roll_count"] = mn, mx, mean, cnt
    inc, rt, nr = rate_ragged(ts, val, off, W)
    G["c5/increase"], G["c5/rate"], G["c5/n_resets"] = inc, rt, nr
    al, idx = align_pandas(ts, val, off, t0, 15_000, n_steps, LOOKBACK)
    G["c5/aligned"], G["c5/src_idx"] = al, idx
    mn, mx, mean, cnt = rolling_pandas(al, 60, 15_000, t0)
    G["c5/grid_roll_min"], G["c5/grid_roll_max"], G["c5/grid_roll_mean"], G["c5/grid_roll_count"] = mn, mx, mean, cnt
    inc, rt, nr = rate_dense(al, 60, 15_000)
    G["c5/grid_increase"], G["c5/grid_rate"], G["c5/grid_n_resets"] = inc, rt, nr
    last, bmean, bcnt = bucket_pandas(ts, val, off, t0, 60_000, n_steps // 4)
    G["c5/bucket_params"] = np.array([t0, 60_000, n_steps // 4], dtype=np.int64)
    G["c5/bucket_last"], G["c5/bucket_mean"], G["c5/bucket_count"] = last, bmean, bcnt
    # config 2's 15 s scrape down-sampled to 1 min buckets as well
    d2 = {k[3:]: v for k, v in G.items() if k in ("c2/ts", "c2/val", "c2/off")}
    last, bmean, bcnt = bucket_pandas(d2["ts"], d2["val"], d2["off"], synth.START_MS, 60_000, 360)
    G["c2/bucket_params"] = np.array([synth.START_MS, 60_000, 360], dtype=np.int64)
    G["c2/bucket_last"], G["c2/bucket_mean"], G["c2/bucket_count"] = last, bmean, bcnt
    path = os.path.join(HERE, "range_golden.npz")
    np.savez_compressed(path, **G)
    print("wrote", path, os.path.getsize(path), "bytes,", len(G), "arrays")


if __name__ == "__main__":
    main()
