"""GPU: K5 group reductions and K6 radix-select quantiles against the oracle."""
import numpy as np
import pytest

from k8s_rocm_metrics_dashboard_b200 import group_index, synth
from oracle import prom_oracle as O
from tests.util import assert_bits_equal, assert_close, assert_exact_or_nan

pytestmark = pytest.mark.gpu


def _np(t):
    return t.cpu().numpy()


@pytest.mark.parametrize("S,T,G", [(12, 40, 3), (96, 1000, 7), (5, 1, 5), (300, 257, 1)])
def test_group_reduce_scalar_and_per_step(eng, S, T, G):
    rng = np.random.default_rng(S * T)
    x = rng.uniform(-5, 10, (S, T)); x[rng.random(x.shape) < 0.1] = np.nan
    gid = rng.integers(-1, G, S).astype(np.int32)
    gid[0] = G - 1
    for per_step in (False, True):
        sm, cnt, mn, mx, mean = eng.group_reduce(x, gid, G, per_step=per_step)
        osm, ocnt, omn, omx, omean = O.group_reduce(x, gid, G, per_step=per_step)
        assert np.array_equal(_np(cnt), ocnt)
        assert_exact_or_nan(_np(mn), omn, "min"); assert_exact_or_nan(_np(mx), omx, "max")
        # relative to the result, floored at the group's mean |x| (values of both signs may cancel)
        ax = np.where(np.isnan(x), 0.0, np.abs(x))
        floor = np.stack([ax[gid == g].sum(axis=0) if per_step else ax[gid == g].sum() for g in range(G)]) / np.maximum(ocnt, 1)
        assert_close(_np(sm), osm, scale=floor, rtol=1e-12, what="group sum")
        assert_close(_np(mean), omean, scale=floor, rtol=1e-12, what="group mean")


def test_group_reduce_empty_group_and_all_nan(eng):
    x = np.full((4, 10), np.nan); x[0, 3] = 2.0
    sm, cnt, mn, mx, mean = eng.group_reduce(x, np.array([0, 0, 1, 1], dtype=np.int32), 3)
    assert _np(cnt).tolist() == [1, 0, 0] and _np(sm).tolist() == [2.0, 0.0, 0.0]
    assert _np(mn)[0] == 2.0 and np.isnan(_np(mn)[1:]).all() and np.isnan(_np(mean)[2])


def test_group_reduce_per_step_is_sequential_like_pandas(eng):
    """Per-step sums visit the series in order: bit-identical to a left-to-right Python sum."""
    rng = np.random.default_rng(1)
    x = rng.uniform(0, 1e6, (50, 33))
    sm, cnt, *_ = eng.group_reduce(x, np.zeros(50, dtype=np.int32), 1, per_step=True)
    exp = np.zeros(33)
    for s in range(50):
        exp = exp + x[s]
    assert_bits_equal(_np(sm)[0], exp)


def _check_quantiles(eng, x, seg, n_seg, qs):
    from k8s_rocm_metrics_dashboard_b200 import group_index
    _check_quantiles_1(eng, x, seg, n_seg, qs, None)                                   # rows in storage order
    _check_quantiles_1(eng, x, seg, n_seg, qs, group_index(np.asarray(seg), n_seg)[0])  # rows grouped by segment


    _check_quantiles_1(eng, x, seg, n_seg, qs, None, _summary(x, seg, n_seg))          # pass 0 replaced by count/min/max
    _check_quantiles_1(eng, x, seg, n_seg, qs, group_index(np.asarray(seg), n_seg)[0], _summary(x, seg, n_seg))


def _summary(x, seg, n_seg):
    """count / min / max per segment the way a group reduction reports them (NaN when empty)."""
    cnt = np.zeros(n_seg, dtype=np.int64); mn = np.full(n_seg, np.nan); mx = np.full(n_seg, np.nan)
    for g in range(n_seg):
        v = x[np.asarray(seg) == g]; v = v[~np.isnan(v)]
        cnt[g] = v.size
        if v.size:
            mn[g] = v.min(); mx[g] = v.max()
    return cnt, mn, mx


def _check_quantiles_1(eng, x, seg, n_seg, qs, order, summary=None):
    out = eng.quantiles(x, seg, n_seg, qs, row_order=order, summary=summary)
    for g in range(n_seg):
        o = O.quantiles(x[np.asarray(seg) == g], qs)
        assert int(_np(out["n"])[g]) == o["n"]
        assert _np(out["lo"])[g].tolist() == o["lo"] and _np(out["hi"])[g].tolist() == o["hi"]
        assert_bits_equal(_np(out["x_lo"])[g], np.array(o["x_lo"]), "x_lo")
        assert_bits_equal(_np(out["x_hi"])[g], np.array(o["x_hi"]), "x_hi")
        assert_bits_equal(_np(out["value"])[g], np.array(o["value"]), "value")


def test_quantiles_exact_order_statistics(eng):
    rng = np.random.default_rng(4)
    x = rng.normal(size=(24, 700)) * np.exp(rng.normal(size=(24, 1)) * 5)
    x[rng.random(x.shape) < 0.05] = np.nan
    x[0, 0] = np.inf; x[1, 1] = -np.inf; x[2, 2] = -0.0; x[3, 3] = 0.0; x[4, 4] = 5e-324; x[5, 5] = -5e-324
    seg = (np.arange(24) % 3).astype(np.int32)
    _check_quantiles(eng, x, seg, 3, [0.5, 0.95, 0.99])
    _check_quantiles(eng, x, seg, 3, [0.0, 1.0, 0.25, 0.123456789])


def test_quantiles_heavy_ties_and_tiny_segments(eng):
    rng = np.random.default_rng(5)
    x = rng.integers(0, 101, (16, 500)).astype(np.float64)       # gfx_activity: 101 distinct values
    seg = np.array([0] * 8 + [1] * 7 + [-1], dtype=np.int32)
    _check_quantiles(eng, x, seg, 2, [0.5, 0.95, 0.99])
    y = np.full((3, 4), np.nan); y[0, 1] = 7.0; y[1, 0] = 1.0; y[1, 3] = 2.0
    _check_quantiles(eng, y, np.array([0, 1, 2], dtype=np.int32), 3, [0.5, 0.99])   # n = 1, 2, 0


def test_quantiles_config4_shape_scaled(eng):
    d = synth.config("config4", seed=2, scale=1 / 16)
    al, _ = O.align(d["ts"], d["val"], d["off"], d["start_ms"], d["step_ms"], d["n_steps"])
    names = sorted({l["__name__"] for l in d["labels"]})
    seg = np.array([names.index(l["__name__"]) for l in d["labels"]], dtype=np.int32)
    _check_quantiles(eng, al, seg, len(names), [0.5, 0.95, 0.99])


def test_quantiles_candidate_list_overflow_falls_back_to_full_passes(eng):
    """More than 2^20 equal values in one segment overflow its candidate list (kept from pass 2
    on); that segment must keep reading the full data, the other one uses its list."""
    rng = np.random.default_rng(9)
    x = np.full((320, 4096), 5.0)                       # 1.3 M ties
    x[:, ::97] = rng.normal(size=x[:, ::97].shape)
    y = rng.normal(size=(8, 4096))
    z = np.concatenate([x, y])
    seg = np.array([0] * 320 + [1] * 8, dtype=np.int32)
    _check_quantiles(eng, z, seg, 2, [0.5, 0.95, 0.99])
    _check_quantiles(eng, z, seg, 2, [0.001, 0.999])


def test_quantiles_many_segments_use_row_major_schedule(eng):
    """> 256 segments: CTAs sweep consecutive rows and switch histograms (one series per segment)."""
    rng = np.random.default_rng(12)
    x = rng.normal(size=(300, 257)); x[rng.random(x.shape) < 0.1] = np.nan
    seg = np.arange(300, dtype=np.int32)
    _check_quantiles(eng, x, seg, 300, [0.5, 0.9])


def test_quantiles_from_summary_signed_zero_constant_and_wide_segments(eng):
    """The count/min/max start: a zero bound may hide the other zero's sign; a constant segment is
    resolved by the bounds alone; a segment spanning both signs shares no key bits (all 6 passes);
    a gauge-like range within a factor of four needs passes 1..5 only."""
    rng = np.random.default_rng(21)
    x = np.empty((8, 300))
    x[0:2] = rng.choice([-0.0, 0.0, 1.5], size=(2, 300), p=[0.02, 0.02, 0.96])   # zeros of both signs below the ranks asked for
    x[2:4] = 42.125                                                  # min == max
    x[4:6] = rng.normal(size=(2, 300)) * 1e200                       # both signs, huge range
    x[6:8] = 30 + 65 * rng.random((2, 300))                          # gauge-like
    x[rng.random(x.shape) < 0.02] = np.nan
    seg = np.repeat(np.arange(4), 2).astype(np.int32)
    qs = [0.3, 0.5, 1.0]
    for zero in (0.0, -0.0):
        cnt, mn, mx = _summary(x, seg, 4)
        if mn[0] == 0.0:
            mn[0] = zero                                             # either sign is a legal report of the minimum
        out = eng.quantiles(x, seg, 4, qs, summary=(cnt, mn, mx))
        for g in range(4):
            o = O.quantiles(x[seg == g], qs)
            assert int(_np(out["n"])[g]) == o["n"]
            assert_bits_equal(_np(out["x_lo"])[g], np.array(o["x_lo"]), "x_lo")
            assert_bits_equal(_np(out["x_hi"])[g], np.array(o["x_hi"]), "x_hi")
            assert_bits_equal(_np(out["value"])[g], np.array(o["value"]), "value")


@pytest.mark.parametrize("start", ["summary", "pass0"])
def test_quantiles_heavy_ties_integer_gauges(eng, start):
    """Integer-valued gauges (the reference's amd_gpu_gfx_activity is 0..100): thousands of equal keys
    per target survive every digit.  Targets whose remaining keys are all one value are settled
    at once; a segment that mixes such ties with spread values runs its last passes on the
    candidate list."""
    rng = np.random.default_rng(5)
    S, T, G = 768, 2880, 12
    x = rng.integers(0, 101, (S, T)).astype(np.float64)
    x[1::12] = np.round(rng.uniform(0, 100, x[1::12].shape), 1)            # 1001 distinct values
    x[2::12] = np.where(rng.random(x[2::12].shape) < 0.9, 50.0, rng.normal(50, 1e-7, x[2::12].shape))   # ties + near-ties
    x[rng.random(x.shape) < 0.01] = np.nan
    gid = (np.arange(S) % G).astype(np.int32)
    pg = group_index(gid, G)
    qs = [0.5, 0.95, 0.99]
    if start == "summary":
        sm, cnt, mn, mx, _ = eng.group_reduce(x, None, G, perm_goff=pg)
        q = eng.quantiles(x, gid, G, qs, row_order=pg[0], summary=(cnt, mn, mx))
    else:
        q = eng.quantiles(x, gid, G, qs, row_order=pg[0])
    for g in range(G):
        o = O.quantiles(x[gid == g], qs)
        assert int(q["n"][g]) == o["n"]
        for k in ("x_lo", "x_hi", "value"):
            assert_bits_equal(q[k][g].cpu().numpy(), np.array(o[k]), f"{k} group {g}")
