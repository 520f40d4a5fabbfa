"""GPU: BASELINE.json's large configs at their stated per-GPU size.  The kernels run on data generated
on the device; a deterministic subsample of the series goes to the host and is compared with the
oracle, and the whole output is checked through size-independent properties (the fused kernel
against the separate kernels bit for bit, order relations, counts)."""
import numpy as np
import pytest

from k8s_rocm_metrics_dashboard_b200 import synth
from oracle import prom_oracle as O
from tests.util import assert_bits_equal, assert_close, assert_exact_or_nan

pytestmark = pytest.mark.gpu


def test_config3_shard_16384_series_x_40320_steps(eng):
    """Config 3 (1024 nodes x 8 GPUs x 16 metrics, 7 d @ 15 s) is 16 384 series x 40 320 steps per GPU on
    eight GPUs: 660 M samples, 42 tiles per series."""
    t = eng.torch
    S, T, M, w = 16384, 40320, 16, 60
    ts, val, off = synth.dense_device(S, T, M, seed=33, device=eng.device)
    out = eng.fused(ts, val, off, synth.START_MS, synth.STEP_MS, T, w)
    # -- every 256th series against the oracle (rolling), every 1024th also for the rates
    rows = np.arange(0, S, 256)
    sel = t.as_tensor(rows, device=eng.device)
    h_ts = ts.view(S, T)[sel].cpu().numpy().ravel(); h_val = val.view(S, T)[sel].cpu().numpy().ravel()
    h_off = np.arange(len(rows) + 1, dtype=np.int64) * T
    got = {k: v[sel].cpu().numpy() for k, v in out.items()}
    al, _ = O.align(h_ts, h_val, h_off, synth.START_MS, synth.STEP_MS, T)
    assert_bits_equal(got["aligned"], al, "config3 aligned")
    mn, mx, mean, _ = O.rolling(al, w)
    assert_exact_or_nan(got["roll_min"], mn, "config3 roll_min"); assert_exact_or_nan(got["roll_max"], mx, "config3 roll_max")
    assert_close(got["roll_mean"], mean, scale=O.rolling_scale(al, w), rtol=1e-12, what="config3 roll_mean")
    _, rt, _ = O.rate(al[::4], w, synth.STEP_MS)
    assert_close(got["rate"][::4], rt, rtol=1e-12, what="config3 rate")
    del got
    # -- the whole output: the separate kernels agree bit for bit, min <= mean <= max, first point empty
    al_d, _ = eng.align(ts, val, off, synth.START_MS, synth.STEP_MS, T, want_idx=False)
    assert t.equal(out["aligned"].view(t.int64), al_d.view(t.int64))
    del ts, val
    mn_d, mx_d, mean_d, _ = eng.rolling(al_d, w, want_count=False)
    for a, b in ((out["roll_min"], mn_d), (out["roll_max"], mx_d), (out["roll_mean"], mean_d)):
        assert t.equal(a.view(t.int64), b.view(t.int64))
    del mn_d, mx_d
    ok = ~t.isnan(mean_d)
    assert bool(((out["roll_min"] <= mean_d + 1e-9 * mean_d.abs()) & (mean_d <= out["roll_max"] + 1e-9 * mean_d.abs()))[ok].all())
    assert float(t.isnan(out["aligned"][:, 0]).float().mean()) > 0.99
    assert int(t.isnan(out["aligned"]).sum()) == int(t.isnan(al_d).sum())
    # per-metric summary and exact quantiles on the shard, against numpy on the host for two metrics
    from k8s_rocm_metrics_dashboard_b200 import group_index
    gid = (np.arange(S) % M).astype(np.int32)
    pg = group_index(gid, M)
    sm, cnt, gmn, gmx, _ = eng.group_reduce(al_d, None, M, perm_goff=pg)
    q = eng.quantiles(al_d, gid, M, [0.5, 0.95, 0.99], row_order=pg[0], summary=(cnt, gmn, gmx))
    for g in (0, 2):                                              # a gauge and a counter
        x = al_d[t.as_tensor(np.flatnonzero(gid == g), device=eng.device)].cpu().numpy().ravel()
        o = O.quantiles(x, [0.5, 0.95, 0.99])
        assert int(q["n"][g]) == o["n"] == int(cnt[g])
        assert_bits_equal(q["value"][g].cpu().numpy(), np.array(o["value"]), f"config3 quantiles metric {g}")
        assert float(gmn[g]) == np.nanmin(x) and float(gmx[g]) == np.nanmax(x)


def test_config5_one_million_ragged_series(eng):
    """Config 5 at its stated size: 1 000 000 ragged series (~400 M samples), 15-minute windows at
    every sample, irregular timestamps, scrape gaps, NaN and stale markers."""
    t = eng.torch
    S, W = 1_000_000, 900_000
    ts, val, off = synth.ragged_device(S, seed=55, device=eng.device)
    N = ts.numel()
    assert 300_000_000 < N < 500_000_000
    mn, mx, mean, cnt = eng.rolling_ragged(ts, val, off, W)
    inc, rt, _ = eng.rate_ragged(ts, val, off, W, want_resets=False)
    # -- every 2000th series against the oracle
    h_off_all = off.cpu().numpy()
    rows = np.arange(0, S, 2000)
    idx = np.concatenate([np.arange(h_off_all[s], h_off_all[s + 1]) for s in rows])
    didx = t.as_tensor(idx, device=eng.device)
    h_ts, h_val = ts[didx].cpu().numpy(), val[didx].cpu().numpy()
    lens = h_off_all[rows + 1] - h_off_all[rows]
    h_off = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    omn, omx, omean, ocnt = O.rolling_ragged(h_ts, h_val, h_off, W)
    assert_exact_or_nan(mn[didx].cpu().numpy(), omn, "config5 min"); assert_exact_or_nan(mx[didx].cpu().numpy(), omx, "config5 max")
    assert np.array_equal(cnt[didx].cpu().numpy(), ocnt)
    assert_close(mean[didx].cpu().numpy(), omean, rtol=1e-12, what="config5 mean")
    oinc, ort, _ = O.rate_ragged(h_ts, h_val, h_off, W)
    assert_close(inc[didx].cpu().numpy(), oinc, rtol=1e-12, what="config5 increase")
    assert_close(rt[didx].cpu().numpy(), ort, rtol=1e-12, what="config5 rate")
    # -- the whole output: order relations and counts
    ok = ~t.isnan(mean)
    assert bool(((mn <= mean + 1e-9 * mean.abs()) & (mean <= mx + 1e-9 * mean.abs()))[ok].all())
    assert int(cnt.min()) >= 0 and int(cnt.max()) <= 70          # 900 s / 13 s
    assert bool((t.isnan(mean) == (cnt == 0)).all())             # no +-inf in this data: a window with a sample has a mean
    assert bool((~t.isnan(inc) | (cnt < 2)).all())
