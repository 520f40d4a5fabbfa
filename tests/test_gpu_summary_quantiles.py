"""pmx_summary_quantiles (SURVEY X6 + X7 in one sweep): the summary and the exact quantiles must be those
of pmx_group_reduce and of the per-pass selection bit for bit, on every kind of data -- and those of the
oracle (numpy.percentile's definition) where the oracle is fast enough."""
import numpy as np
import pytest

from k8s_rocm_metrics_dashboard_b200 import group_index, sharding
from oracle import prom_oracle as O
from tests.util import assert_bits_equal

pytestmark = pytest.mark.gpu
QS = (0.5, 0.95, 0.99)


def _both(eng, x, gid, n_seg, qs=QS):
    t = eng.torch
    dx = t.as_tensor(x, device=eng.device)
    perm, goff = group_index(gid, n_seg)
    pg = (t.as_tensor(perm, device=eng.device), t.as_tensor(goff, device=eng.device))
    seg = t.as_tensor(np.asarray(gid, dtype=np.int32), device=eng.device)
    summ, q, failed = eng.summary_quantiles(dx, seg, n_seg, qs, pg)
    ref_s = eng.group_reduce(dx, None, n_seg, perm_goff=pg)
    ref_q = eng.quantiles(dx, seg, n_seg, list(qs), row_order=pg[0], summary=(ref_s[1], ref_s[2], ref_s[3]))
    return summ, q, int(failed.item()), ref_s, ref_q


def _same(summ, q, ref_s, ref_q, what):
    for a, b, name in zip(summ, ref_s, ("sum", "count", "min", "max", "mean")):
        a, b = a.cpu().numpy(), b.cpu().numpy()
        if a.dtype == np.float64:
            assert_bits_equal(a, b, f"{what}: {name}")
        else:
            assert np.array_equal(a, b), f"{what}: {name}"
    for k in ("n", "lo", "hi"):
        assert np.array_equal(q[k].cpu().numpy(), ref_q[k].cpu().numpy()), f"{what}: {k}"
    for k in ("x_lo", "x_hi", "value"):
        assert_bits_equal(q[k].cpu().numpy(), ref_q[k].cpu().numpy(), f"{what}: {k}")


def _data(kind, S, T, rng):
    if kind == "uniform":
        return rng.uniform(20.0, 95.0, (S, T)).round(3)
    if kind == "integers":                                         # gauges in whole per cent: ~100 distinct values
        return rng.integers(0, 101, (S, T)).astype(np.float64)
    if kind == "constant":
        return np.full((S, T), 42.5)
    if kind == "counters":
        return np.cumsum(rng.uniform(0, 50, (S, T)), axis=1) + rng.uniform(0, 1e9, (S, 1))
    if kind == "zeros":                                            # idle devices: mostly zero, of both signs
        x = np.where(rng.random((S, T)) < 0.7, 0.0, rng.normal(0, 1, (S, T)))
        x[rng.random((S, T)) < 0.2] = -0.0
        return x
    if kind == "signed":
        return rng.normal(0.0, 1e3, (S, T))
    if kind == "two_values":
        return np.where(rng.random((S, T)) < 0.5, 1.0, 2.0)
    if kind == "nan_heavy":
        x = rng.uniform(0, 1, (S, T))
        x[rng.random((S, T)) < 0.6] = np.nan
        return x
    if kind == "infs":
        x = rng.normal(0, 1, (S, T))
        x[rng.random((S, T)) < 0.02] = np.inf
        x[rng.random((S, T)) < 0.02] = -np.inf
        return x
    if kind == "wide":                                             # many binary exponents
        return np.exp(rng.uniform(-40, 40, (S, T))) * rng.choice([-1.0, 1.0], (S, T))
    raise KeyError(kind)


KINDS = ["uniform", "integers", "constant", "counters", "zeros", "signed", "two_values", "nan_heavy", "infs", "wide"]


@pytest.mark.parametrize("kind", KINDS)
def test_one_sweep_equals_the_per_pass_selection(eng, kind):
    rng = np.random.default_rng(KINDS.index(kind) + 100)
    S, T, M = 96, 3001, 6                                          # odd T: rows start off a 32-byte boundary
    x = _data(kind, S, T, rng)
    x[rng.random((S, T)) < 0.01] = np.nan
    gid = (np.arange(S) % M).astype(np.int32)
    summ, q, failed, ref_s, ref_q = _both(eng, x, gid, M)
    assert failed == 0, f"{kind}: {failed} brackets failed"
    _same(summ, q, ref_s, ref_q, kind)
    # ... and the oracle
    for g in range(M):
        o = O.quantiles(x[gid == g], QS)
        assert int(q["n"][g]) == o["n"]
        for k in ("x_lo", "x_hi", "value"):
            if kind == "zeros":                                    # np.sort does not order -0.0 and +0.0; the keys do
                assert np.array_equal(q[k][g].cpu().numpy(), np.array(o[k])), f"{kind}: oracle {k}"
            else:
                assert_bits_equal(q[k][g].cpu().numpy(), np.array(o[k]), f"{kind}: oracle {k}")


def test_small_empty_and_unlisted_segments(eng):
    rng = np.random.default_rng(7)
    S, T, M = 40, 17, 9
    x = rng.normal(0, 1, (S, T))
    gid = rng.integers(0, 6, S).astype(np.int32)                   # groups 6..8 stay empty
    gid[:3] = -1                                                   # series outside every group
    x[gid == 4] = np.nan                                           # a group with no value at all
    one = np.flatnonzero(gid == 5)
    x[one] = np.nan
    if len(one):
        x[one[0], 3] = 7.25                                        # a group with a single value
    summ, q, failed, ref_s, ref_q = _both(eng, x, gid, M)
    assert failed == 0
    _same(summ, q, ref_s, ref_q, "small")
    for qs in ((0.0,), (1.0,), (0.0, 1.0), (0.25, 0.5, 0.75, 0.999)):
        summ, q, failed, ref_s, ref_q = _both(eng, x, gid, M, qs)
        assert failed == 0
        _same(summ, q, ref_s, ref_q, f"levels {qs}")


def test_config2_sized_segment_on_device_data(eng):
    """One metric of config 2 (512 series x 5760 steps = 2.9 M values per segment) in three kinds."""
    from k8s_rocm_metrics_dashboard_b200 import synth
    t = eng.torch
    S, T, M = 2048, 5760, 4
    ts, val, off = synth.dense_device(S, T, M, seed=5, device=eng.device)
    x = val.reshape(S, T).clone()
    x[1::4] = t.round(x[1::4])                                     # one segment of whole numbers
    x[2::4] = 0.0                                                  # one constant segment
    gid = (np.arange(S) % M).astype(np.int32)
    summ, q, failed, ref_s, ref_q = _both(eng, x, gid, M)
    assert failed == 0
    _same(summ, q, ref_s, ref_q, "config-2 sized")


def test_overflowing_list_is_reported_and_summarize_falls_back(eng):
    """With 2048 segments a list holds 16384 keys; a segment of 300 000 distinct values puts ~25 000 inside
    its brackets: the entry reports it, ShardedEngine.summarize then runs the per-pass selection."""
    t = eng.torch
    rng = np.random.default_rng(11)
    S, T, M = 64, 18750, 2048
    x = rng.normal(0, 1, (S, T))
    gid = (np.arange(S) % 4).astype(np.int32)                      # four segments of 300 000 values, 2044 empty
    summ, q, failed, ref_s, ref_q = _both(eng, x, gid, M)
    assert failed > 0
    sh = sharding.ShardedEngine(eng, local_only=True)
    assert not sh.one_sweep_pays(S, T, M)                          # thousands of short segments: summarize would not even try
    sh.ONE_SWEEP_MIN_VALUES_PER_SEGMENT = 0                        # ... so make it
    perm, goff = group_index(gid, M)
    pg = (t.as_tensor(perm, device=eng.device), t.as_tensor(goff, device=eng.device))
    s2, q2 = sh.summarize(t.as_tensor(x, device=eng.device), t.as_tensor(gid, device=eng.device), M, pg, QS)
    _same(s2, q2, ref_s, ref_q, "fallback")
    # deferred: the same call says so afterwards instead
    sh.summarize(t.as_tensor(x, device=eng.device), t.as_tensor(gid, device=eng.device), M, pg, QS, defer_check=True)
    assert eng.deferred_ok(wait=True) is False


def test_more_than_four_levels_use_the_passes(eng):
    rng = np.random.default_rng(12)
    t = eng.torch
    x = rng.normal(0, 1, (24, 500))
    gid = (np.arange(24) % 3).astype(np.int32)
    perm, goff = group_index(gid, 3)
    pg = (t.as_tensor(perm, device=eng.device), t.as_tensor(goff, device=eng.device))
    qs = (0.1, 0.25, 0.5, 0.75, 0.9)
    sh = sharding.ShardedEngine(eng, local_only=True)
    _, q = sh.summarize(t.as_tensor(x, device=eng.device), t.as_tensor(gid, device=eng.device), 3, pg, qs)
    for g in range(3):
        o = O.quantiles(x[gid == g], qs)
        assert_bits_equal(q["value"][g].cpu().numpy(), np.array(o["value"]), "five levels")
    with pytest.raises(Exception):
        eng.summary_quantiles(t.as_tensor(x, device=eng.device), t.as_tensor(gid, device=eng.device), 3, qs, pg)


def test_thinned_sample_of_rows_that_move_together(eng):
    """More than 98 304 sample slots per segment: the bounds use a thinned sample.  Counters that rise
    together make every stretch of time one narrow cluster of values, so a thinning with a fixed
    phase would miss whole ranges (config 3's counters did); the phase is hashed."""
    t = eng.torch
    S, T, M = 512, 32768, 2                                        # 256 rows x 512 slots = 131 072 slots per segment
    g = t.Generator(device=eng.device); g.manual_seed(21)
    slope = 4000.0 + 2000.0 * t.rand((S, 1), generator=g, device=eng.device, dtype=t.float64)
    x = slope * t.arange(T, device=eng.device, dtype=t.float64)[None, :] + t.rand((S, T), generator=g, device=eng.device, dtype=t.float64)
    gid = (np.arange(S) % M).astype(np.int32)
    perm, goff = group_index(gid, M)
    pg = (t.as_tensor(perm, device=eng.device), t.as_tensor(goff, device=eng.device))
    seg = t.as_tensor(gid, device=eng.device)
    for qs in (QS, (0.01, 0.25, 0.75, 0.999)):
        summ, q, failed = eng.summary_quantiles(x, seg, M, qs, pg)
        assert int(failed.item()) == 0
        ref_s = eng.group_reduce(x, None, M, perm_goff=pg)
        ref_q = eng.quantiles(x, seg, M, list(qs), row_order=pg[0], summary=(ref_s[1], ref_s[2], ref_s[3]))
        _same(summ, q, ref_s, ref_q, f"thinned {qs}")


def test_brackets_from_the_raw_samples(eng):
    """pmx_summary_brackets_csr + pmx_summary_sweep: brackets taken from the series' raw samples (ragged CSR,
    gaps, NaN) instead of the aligned matrix give the same exact results -- they are only hints."""
    from k8s_rocm_metrics_dashboard_b200 import synth
    t = eng.torch
    d = synth.dense(2, 8, 6, 3000, seed=31, nan_frac=0.01)
    S, T, M = len(d["off"]) - 1, 3000, 6
    # thin the series unevenly: the raw samples are no longer one per grid step
    keep = np.ones(len(d["ts"]), dtype=bool)
    rng = np.random.default_rng(5)
    keep[rng.random(len(keep)) < 0.15] = False
    lens = np.array([keep[int(d["off"][s]):int(d["off"][s + 1])].sum() for s in range(S)])
    ts, val, off = d["ts"][keep], d["val"][keep], np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    al, _ = eng.align(ts, val, off, d["start_ms"], d["step_ms"], T, want_idx=False)
    gid = (np.arange(S) % M).astype(np.int32)
    perm, goff = group_index(gid, M)
    pg = (t.as_tensor(perm, device=eng.device), t.as_tensor(goff, device=eng.device))
    seg = t.as_tensor(gid, device=eng.device)
    eng.summary_brackets_csr(t.as_tensor(val, device=eng.device), t.as_tensor(off, device=eng.device), T, M, QS, pg)
    summ, q, failed = eng.summary_quantiles(al, seg, M, QS, pg, brackets_ready=True)
    assert int(failed.item()) == 0
    s1, q1, f1 = eng.summary_quantiles(al, seg, M, QS, pg)
    assert int(f1.item()) == 0
    _same(summ, q, s1, q1, "brackets from raw samples")
    x = al.cpu().numpy()
    for g in range(M):
        o = O.quantiles(x[gid == g], QS)
        assert_bits_equal(q["value"][g].cpu().numpy(), np.array(o["value"]), "oracle")
