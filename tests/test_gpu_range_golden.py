"""GPU: the CUDA path (through the C ABI) against ``tests/golden/range_golden.npz`` -- the vectors
pandas / numpy produced themselves for the range stages X3-X7 (tests/golden/make_range_golden.py).
Same arrays, same bars as tests/test_range_golden.py holds the oracle to."""
import json
import os

import numpy as np
import pytest

from k8s_rocm_metrics_dashboard_b200 import group_index
from oracle import prom_oracle as O
from tests.util import assert_bits_equal, assert_close, assert_exact_or_nan

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def G(golden_dir):
    return np.load(os.path.join(golden_dir, "range_golden.npz"))


def _np(t):
    return t.cpu().numpy()


@pytest.mark.parametrize("name", ["c1", "c2", "c4"])
def test_dense_configs_align_rolling_rate(eng, G, name):
    start, step, T, w, lookback = (int(x) for x in G[f"{name}/params"][:5])
    ts, val, off = G[f"{name}/ts"], G[f"{name}/val"], G[f"{name}/off"]
    gal = G[f"{name}/aligned"]
    al, idx = eng.align(ts, val, off, start, step, T, lookback)
    assert np.array_equal(_np(idx), G[f"{name}/src_idx"]), "alignment indices"
    assert_bits_equal(_np(al), gal, "aligned values")
    # the fused kernel and the separate kernels, both against pandas' own numbers
    out = eng.fused(ts, val, off, start, step, T, w, lookback)
    assert_bits_equal(_np(out["aligned"]), gal, "fused aligned")
    mn, mx, mean, cnt = eng.rolling(al, w)
    sc = O.rolling_scale(gal, w)
    for got_mn, got_mx, got_mean in ((mn, mx, mean), (out["roll_min"], out["roll_max"], out["roll_mean"])):
        assert_exact_or_nan(_np(got_mn), G[f"{name}/roll_min"], "rolling('900s').min")
        assert_exact_or_nan(_np(got_mx), G[f"{name}/roll_max"], "rolling('900s').max")
        assert_close(_np(got_mean), G[f"{name}/roll_mean"], scale=sc, rtol=1e-12, what="rolling('900s').mean")
    assert np.array_equal(_np(cnt), G[f"{name}/roll_count"]), "rolling('900s').count"
    mn, mx, mean, _ = eng.rolling(al, w, min_periods=w)
    assert_exact_or_nan(_np(mn), G[f"{name}/fix_min"], "rolling(60).min")
    assert_exact_or_nan(_np(mx), G[f"{name}/fix_max"], "rolling(60).max")
    assert_close(_np(mean), G[f"{name}/fix_mean"], scale=sc, rtol=1e-12, what="rolling(60).mean")
    inc, rt, nr = eng.rate(al, w, step)
    assert np.array_equal(_np(nr), G[f"{name}/n_resets"]), "reset counts"
    rs = O.rate_scale(gal, w)
    assert_close(_np(inc), G[f"{name}/increase"], scale=rs, rtol=1e-12, what="increase")
    assert_close(_np(rt), G[f"{name}/rate"], scale=rs / (step / 1000.0), rtol=1e-12, what="rate")
    assert_close(_np(out["rate"]), G[f"{name}/rate"], scale=rs / (step / 1000.0), rtol=1e-12, what="fused rate")


@pytest.mark.parametrize("name", ["c1", "c2", "c4"])
def test_dense_configs_groupby_and_quantiles(eng, G, name):
    gal = G[f"{name}/aligned"]
    labels = [json.loads(s) for s in G[f"{name}/labels"]]
    for tag, keys in (("card", ("instance", "gpu_id", "__name__")), ("metric", ("__name__",))):
        names = [str(x) for x in G[f"{name}/group_{tag}/names"]]
        gid = np.array([names.index("\x1f".join(l[k] for k in keys)) for l in labels], dtype=np.int32)
        sm, cnt, mn, mx, mean = eng.group_reduce(gal, gid, len(names))
        assert np.array_equal(_np(cnt), G[f"{name}/group_{tag}/count"]), tag
        assert_exact_or_nan(_np(mn), G[f"{name}/group_{tag}/min"], tag)
        assert_exact_or_nan(_np(mx), G[f"{name}/group_{tag}/max"], tag)
        assert_close(_np(sm), G[f"{name}/group_{tag}/sum"], rtol=1e-12, what=f"{tag} sum")
        assert_close(_np(mean), G[f"{name}/group_{tag}/mean"], rtol=1e-12, what=f"{tag} mean")
    gid = G[f"{name}/metric_id"]
    n_seg = int(gid.max()) + 1
    pg = group_index(gid, n_seg)
    sm, cnt, mn, mx, _ = eng.group_reduce(gal, None, n_seg, perm_goff=pg)
    for q in (eng.quantiles(gal, gid, n_seg, [0.5, 0.95, 0.99]),
              eng.quantiles(gal, gid, n_seg, [0.5, 0.95, 0.99], row_order=pg[0], summary=(cnt, mn, mx))):
        assert np.array_equal(_np(q["n"]), G[f"{name}/quantile/n"])
        assert np.array_equal(_np(q["lo"]), G[f"{name}/quantile/lo"]) and np.array_equal(_np(q["hi"]), G[f"{name}/quantile/hi"])
        assert_bits_equal(_np(q["x_lo"]), G[f"{name}/quantile/x_lo"], "x[lo]")
        assert_bits_equal(_np(q["x_hi"]), G[f"{name}/quantile/x_hi"], "x[hi]")
        assert_bits_equal(_np(q["value"]), G[f"{name}/quantile/value"], "quantile value")


def test_ragged_config5(eng, G):
    ts, val, off = G["c5/ts"], G["c5/val"], G["c5/off"]
    t0, step, T, w, lookback, W = (int(x) for x in G["c5/params"])
    mn, mx, mean, cnt = eng.rolling_ragged(ts, val, off, W)
    assert_exact_or_nan(_np(mn), G["c5/roll_min"]); assert_exact_or_nan(_np(mx), G["c5/roll_max"])
    assert np.array_equal(_np(cnt), G["c5/roll_count"])
    assert_close(_np(mean), G["c5/roll_mean"], rtol=1e-12, what="ragged mean")
    sc = np.full(len(ts), float(np.nanmax(np.abs(np.where(np.isfinite(val), val, 0.0)))))
    for want_resets in (True, False):                           # direct kernel / tiled kernel
        inc, rt, nr = eng.rate_ragged(ts, val, off, W, want_resets=want_resets)
        if want_resets:
            assert np.array_equal(_np(nr), G["c5/n_resets"])
        assert_close(_np(inc), G["c5/increase"], scale=sc, rtol=1e-12, what="ragged increase")
        assert_close(_np(rt), G["c5/rate"], scale=sc / 13.0, rtol=1e-12, what="ragged rate")
    gal = G["c5/aligned"]
    al, idx = eng.align(ts, val, off, t0, step, T, lookback)
    assert np.array_equal(_np(idx), G["c5/src_idx"]); assert_bits_equal(_np(al), gal)
    out = eng.fused(ts, val, off, t0, step, T, w, lookback)
    assert_bits_equal(_np(out["aligned"]), gal, "fused aligned")
    assert_exact_or_nan(_np(out["roll_min"]), G["c5/grid_roll_min"]); assert_exact_or_nan(_np(out["roll_max"]), G["c5/grid_roll_max"])
    assert_close(_np(out["roll_mean"]), G["c5/grid_roll_mean"], scale=O.rolling_scale(gal, w), rtol=1e-12)
    assert_close(_np(out["rate"]), G["c5/grid_rate"], scale=O.rate_scale(gal, w) / (step / 1000.0), rtol=1e-12)
    inc, rt, nr = eng.rate(al, w, step)
    assert np.array_equal(_np(nr), G["c5/grid_n_resets"])
    assert_close(_np(inc), G["c5/grid_increase"], scale=O.rate_scale(gal, w), rtol=1e-12)


@pytest.mark.parametrize("name", ["c2", "c5"])
def test_bucket_downsampling(eng, G, name):
    start, step, T = (int(x) for x in G[f"{name}/bucket_params"])
    last, mean, cnt = eng.bucket(G[f"{name}/ts"], G[f"{name}/val"], G[f"{name}/off"], start, step, T)
    assert np.array_equal(_np(cnt), G[f"{name}/bucket_count"]), "resample().count"
    assert_exact_or_nan(_np(last), G[f"{name}/bucket_last"], "resample().last")
    assert_close(_np(mean), G[f"{name}/bucket_mean"], rtol=1e-12, what="resample().mean")


def test_bucket_random_against_the_oracle(eng):
    from k8s_rocm_metrics_dashboard_b200 import synth
    d = synth.ragged(200, seed=9, mean_len=300, nan_frac=0.05, stale_frac=0.02, gap_frac=0.03)
    t0 = int(d["ts"].min())
    for start, step, T in ((t0 - 100_000, 60_000, 700), (t0 + 7, 15_000, 2000), (t0, 3_600_000, 5), (t0 + 10**9, 1000, 10)):
        last, mean, cnt = eng.bucket(d["ts"], d["val"], d["off"], start, step, T)
        ol, om, oc = O.bucket(d["ts"], d["val"], d["off"], start, step, T)
        assert np.array_equal(_np(cnt), oc)
        assert_exact_or_nan(_np(last), ol, "last")
        assert_bits_equal(np.nan_to_num(_np(mean)), np.nan_to_num(om), "mean: the same sequential sum")
