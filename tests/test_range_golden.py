"""CPU: the oracle (oracle/prom_oracle.py) against ``tests/golden/range_golden.npz`` -- vectors for the
range stages X3-X7 that pandas / numpy produced themselves (tests/golden/make_range_golden.py).
The reference holds no vector for these stages; this file is what pins the restatement, and
tests/test_gpu_range_golden.py checks the CUDA path against the same arrays."""
import json
import os

import numpy as np
import pandas as pd
import pytest

from oracle import prom_oracle as O
from tests.util import assert_bits_equal, assert_close, assert_exact_or_nan


@pytest.fixture(scope="module")
def G(golden_dir):
    return np.load(os.path.join(golden_dir, "range_golden.npz"))


def test_golden_was_made_by_this_pandas(G):
    v = [str(x) for x in G["versions"]]
    assert v[0] == pd.__version__ and v[1] == np.__version__, \
        f"range_golden.npz was made with pandas {v[0]} / numpy {v[1]}: regenerate it (tests/golden/make_range_golden.py)"


def _params(G, name):
    start, step, T, w, lookback = (int(x) for x in G[f"{name}/params"][:5])
    return start, step, T, w, lookback


@pytest.mark.parametrize("name", ["c1", "c2", "c4"])
def test_dense_configs_align_rolling_rate(G, name):
    start, step, T, w, lookback = _params(G, name)
    ts, val, off = G[f"{name}/ts"], G[f"{name}/val"], G[f"{name}/off"]
    al, idx = O.align(ts, val, off, start, step, T, lookback)
    assert np.array_equal(idx, G[f"{name}/src_idx"]), "alignment indices"
    assert_bits_equal(al, G[f"{name}/aligned"], "aligned values")
    mn, mx, mean, cnt = O.rolling(al, w)
    assert_exact_or_nan(mn, G[f"{name}/roll_min"], "rolling('900s').min")
    assert_exact_or_nan(mx, G[f"{name}/roll_max"], "rolling('900s').max")
    assert np.array_equal(cnt, G[f"{name}/roll_count"]), "rolling('900s').count"
    assert_close(mean, G[f"{name}/roll_mean"], scale=O.rolling_scale(al, w), rtol=1e-12, what="rolling('900s').mean")
    mn, mx, mean, _ = O.rolling(al, w, min_periods=w)
    assert_exact_or_nan(mn, G[f"{name}/fix_min"], "rolling(60).min")
    assert_exact_or_nan(mx, G[f"{name}/fix_max"], "rolling(60).max")
    assert_close(mean, G[f"{name}/fix_mean"], scale=O.rolling_scale(al, w), rtol=1e-12, what="rolling(60).mean")
    inc, rt, nr = O.rate(al, w, step)
    assert np.array_equal(nr, G[f"{name}/n_resets"]), "reset counts"
    sc = O.rate_scale(al, w)
    assert_close(inc, G[f"{name}/increase"], scale=sc, rtol=1e-12, what="increase")
    assert_close(rt, G[f"{name}/rate"], scale=sc / (step / 1000.0), rtol=1e-12, what="rate")


@pytest.mark.parametrize("name", ["c1", "c2", "c4"])
def test_dense_configs_groupby_and_quantiles(G, name):
    al = G[f"{name}/aligned"]
    labels = [json.loads(s) for s in G[f"{name}/labels"]]
    for tag, keys in (("card", ("instance", "gpu_id", "__name__")), ("metric", ("__name__",))):
        names = [str(x) for x in G[f"{name}/group_{tag}/names"]]
        gid = np.array([names.index("\x1f".join(l[k] for k in keys)) for l in labels], dtype=np.int32)
        sm, cnt, mn, mx, mean = O.group_reduce(al, gid, len(names))
        assert np.array_equal(cnt, G[f"{name}/group_{tag}/count"]), tag
        assert_exact_or_nan(mn, G[f"{name}/group_{tag}/min"], tag); assert_exact_or_nan(mx, G[f"{name}/group_{tag}/max"], tag)
        assert_close(sm, G[f"{name}/group_{tag}/sum"], rtol=1e-12, what=f"{tag} sum")
        assert_close(mean, G[f"{name}/group_{tag}/mean"], rtol=1e-12, what=f"{tag} mean")
    gid = G[f"{name}/metric_id"]
    for g in range(int(gid.max()) + 1):
        q = O.quantiles(al[gid == g], [0.5, 0.95, 0.99])
        assert q["n"] == int(G[f"{name}/quantile/n"][g])
        assert list(q["lo"]) == list(G[f"{name}/quantile/lo"][g]) and list(q["hi"]) == list(G[f"{name}/quantile/hi"][g])
        assert_bits_equal(q["x_lo"], G[f"{name}/quantile/x_lo"][g], "x[lo]")
        assert_bits_equal(q["x_hi"], G[f"{name}/quantile/x_hi"][g], "x[hi]")
        assert_bits_equal(q["value"], G[f"{name}/quantile/value"][g], "quantile value")


def test_ragged_config5(G):
    ts, val, off = G["c5/ts"], G["c5/val"], G["c5/off"]
    t0, step, T, w, lookback, W = (int(x) for x in G["c5/params"])
    mn, mx, mean, cnt = O.rolling_ragged(ts, val, off, W)
    assert_exact_or_nan(mn, G["c5/roll_min"]); assert_exact_or_nan(mx, G["c5/roll_max"])
    assert np.array_equal(cnt, G["c5/roll_count"])
    assert_close(mean, G["c5/roll_mean"], rtol=1e-12, what="ragged mean")
    inc, rt, nr = O.rate_ragged(ts, val, off, W)
    assert np.array_equal(nr, G["c5/n_resets"])
    sc = np.full(len(inc), float(np.nanmax(np.abs(np.where(np.isfinite(val), val, 0.0)))))
    assert_close(inc, G["c5/increase"], scale=sc, rtol=1e-12, what="ragged increase")
    assert_close(rt, G["c5/rate"], scale=sc / 13.0, rtol=1e-12, what="ragged rate")
    al, idx = O.align(ts, val, off, t0, step, T, lookback)
    assert np.array_equal(idx, G["c5/src_idx"]); assert_bits_equal(al, G["c5/aligned"])
    mn, mx, mean, cnt = O.rolling(al, w)
    assert_exact_or_nan(mn, G["c5/grid_roll_min"]); assert_exact_or_nan(mx, G["c5/grid_roll_max"])
    assert np.array_equal(cnt, G["c5/grid_roll_count"])
    assert_close(mean, G["c5/grid_roll_mean"], scale=O.rolling_scale(al, w), rtol=1e-12)
    inc, rt, nr = O.rate(al, w, step)
    assert np.array_equal(nr, G["c5/grid_n_resets"])
    assert_close(inc, G["c5/grid_increase"], scale=O.rate_scale(al, w), rtol=1e-12)
    assert_close(rt, G["c5/grid_rate"], scale=O.rate_scale(al, w) / (step / 1000.0), rtol=1e-12)


@pytest.mark.parametrize("name", ["c2", "c5"])
def test_bucket_downsampling(G, name):
    start, step, T = (int(x) for x in G[f"{name}/bucket_params"])
    last, mean, cnt = O.bucket(G[f"{name}/ts"], G[f"{name}/val"], G[f"{name}/off"], start, step, T)
    assert np.array_equal(cnt, G[f"{name}/bucket_count"]), "resample().count"
    assert_exact_or_nan(last, G[f"{name}/bucket_last"], "resample().last")
    assert_close(mean, G[f"{name}/bucket_mean"], rtol=1e-12, what="resample().mean")
