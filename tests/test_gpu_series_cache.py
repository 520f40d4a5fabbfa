"""SeriesCache: ingesting a range piece by piece and producing only the grid
points that became due must give what one pass over the whole range gives."""
import numpy as np
import pytest

from k8s_rocm_metrics_dashboard_b200 import synth
from k8s_rocm_metrics_dashboard_b200.cache import SeriesCache
from oracle import prom_oracle as O
from tests.util import assert_bits_equal, assert_close

pytestmark = pytest.mark.gpu

KEYS = ("aligned", "roll_min", "roll_max", "roll_mean", "rate")


def _piece(d, lo_ms, hi_ms, skip=()):
    """CSR of the samples with lo_ms < ts <= hi_ms, as a matrix body (series without samples left out)."""
    lab, ts, val, off = [], [], [], [0]
    for s in range(len(d["off"]) - 1):
        if s in skip:
            continue
        a, b = int(d["off"][s]), int(d["off"][s + 1])
        t = d["ts"][a:b]
        keep = (t > lo_ms) & (t <= hi_ms)
        if keep.any():
            lab.append(d["labels"][s]); ts.append(t[keep]); val.append(d["val"][a:b][keep]); off.append(off[-1] + int(keep.sum()))
    if not lab:
        return b'{"status":"success","data":{"resultType":"matrix","result":[]}}'
    return synth.to_matrix_json(lab, np.concatenate(ts), np.concatenate(val), np.array(off, dtype=np.int64))


def _whole(eng, d, w, n_steps):
    out = eng.fused(d["ts"], d["val"], d["off"], d["start_ms"], d["step_ms"], n_steps, w)
    return {k: v.cpu().numpy() for k, v in out.items()}


def _compare(got, ref, order, w, step_ms):
    al = ref["aligned"][order]
    assert_bits_equal(got["aligned"], al, "aligned")
    for k in ("roll_min", "roll_max"):
        assert_bits_equal(got[k], ref[k][order], k)
    # sums are formed from different tile positions in the two runs: 1e-12 of the window's magnitude
    assert_close(got["roll_mean"], ref["roll_mean"][order], scale=O.rolling_scale(al, w), what="roll_mean")
    assert_close(got["rate"], ref["rate"][order], rtol=1e-12, what="cache rate")


@pytest.mark.parametrize("w", [20, 60])
def test_incremental_refresh_equals_one_pass(eng, w):
    d = synth.dense(1, 4, 6, 400, seed=17, nan_frac=0.02)          # scrape phases, NaN samples, counters
    n_steps, step, start = 400, d["step_ms"], d["start_ms"]
    ref = _whole(eng, d, w, n_steps)
    cache = SeriesCache(eng, start, step, w, capacity=4 * w + 300)
    cols = {k: [] for k in KEYS}
    t_prev = start - 10 * step
    for upto_step in (0, 1, 2, 37, 38, 150, 151, 399):            # single refreshes and long gaps
        upto = start + upto_step * step
        cache.ingest(_piece(d, t_prev - 2 * step, upto))           # overlapping queries: duplicates are skipped
        t_prev = upto
        r = cache.refresh(upto)
        if r is None:                                              # nothing scraped yet: the points stay due
            assert cache.n_series == 0
            continue
        assert r["first_step"] + len(r["ts"]) == upto_step + 1
        for k in KEYS:
            cols[k].append(r[k].cpu().numpy())
    assert cache.refresh(start + 399 * step) is None
    order = [next(i for i, l in enumerate(d["labels"]) if l == lab) for lab in cache.labels]
    got = {k: np.concatenate(v, axis=1) for k, v in cols.items()}
    assert got["aligned"].shape == (24, 400)
    _compare(got, ref, order, w, step)
    # ... and against the oracle itself, not only against the one-pass kernel
    oref = O.aggregate_range(d["ts"], d["val"], d["off"], start, step, n_steps, w * step)
    _compare(got, {k: oref[k] for k in KEYS}, order, w, step)


def test_series_that_start_later_and_device_tokenizer(eng):
    d = synth.dense(1, 2, 5, 200, seed=18, phase=False)
    step, start, w = d["step_ms"], d["start_ms"], 16
    late = {3, 7}
    d2 = {k: v for k, v in d.items()}
    # the reference for a late series: its samples before step 90 removed
    keep = np.ones(len(d["ts"]), dtype=bool)
    for s in late:
        a, b = int(d["off"][s]), int(d["off"][s + 1])
        keep[a:b] = d["ts"][a:b] > start + 90 * step
    lens = np.array([keep[int(d["off"][s]):int(d["off"][s + 1])].sum() for s in range(10)])
    d2["ts"], d2["val"], d2["off"] = d["ts"][keep], d["val"][keep], np.concatenate([[0], np.cumsum(lens)])
    ref = _whole(eng, d2, w, 200)
    cache = SeriesCache(eng, start, step, w, max_series=2)         # the slab has to grow
    cols = {k: [] for k in KEYS}
    rows_seen = []
    t_prev = start - step
    for upto_step in range(9, 200, 10):
        upto = start + upto_step * step
        cache.ingest(_piece(d2, t_prev, upto), tokenizer="device" if upto_step > 50 else "host")
        t_prev = upto
        r = cache.refresh(upto)
        rows_seen.append(r["aligned"].shape[0])
        for k in KEYS:
            cols[k].append(r[k].cpu().numpy())
    assert rows_seen[0] == 8 and rows_seen[-1] == 10
    order = [next(i for i, l in enumerate(d["labels"]) if l == lab) for lab in cache.labels]
    # rows that did not exist yet in early refreshes are all-NaN there
    got = {}
    for k in KEYS:
        parts = []
        for c in cols[k]:
            full = np.full((10, c.shape[1]), np.nan)
            full[:c.shape[0]] = c
            parts.append(full)
        got[k] = np.concatenate(parts, axis=1)
    _compare(got, ref, order, w, step)


def test_capacity_guard(eng):
    d = synth.dense(1, 1, 1, 300, seed=19, phase=False)
    cache = SeriesCache(eng, d["start_ms"], d["step_ms"], 16, capacity=64)
    with pytest.raises(OverflowError):
        cache.ingest(_piece(d, d["start_ms"] - 1, d["start_ms"] + 299 * d["step_ms"]))
    with pytest.raises(ValueError):
        SeriesCache(eng, d["start_ms"], d["step_ms"], 60, capacity=10)


def test_sharded_caches_partition_the_series_and_summarise_together(eng):
    """Two caches that split the series by label hash (what two ranks hold): together they store every
    series exactly once, and each produces the fused kernel's rows for its own shard."""
    from k8s_rocm_metrics_dashboard_b200 import sharding
    d = synth.dense(2, 4, 6, 200, seed=23, nan_frac=0.02)
    step, start, w = d["step_ms"], d["start_ms"], 20
    ref = O.aggregate_range(d["ts"], d["val"], d["off"], start, step, 200, w * step)
    caches = [SeriesCache(eng, start, step, w, capacity=300, rank=r, world=2) for r in range(2)]
    body = _piece(d, start - step, start + 199 * step)
    stored = [c.ingest(body) for c in caches]
    owner = sharding.partition(d["labels"], 2)
    assert [c.n_series for c in caches] == [int((owner == r).sum()) for r in range(2)]
    assert sum(stored) == int(((d["ts"] > start - step) & (d["ts"] <= start + 199 * step)).sum())
    for r, c in enumerate(caches):
        out = c.refresh(start + 199 * step)
        order = [next(i for i, l in enumerate(d["labels"]) if l == lab) for lab in c.labels]
        assert all(owner[i] == r for i in order)
        got = {k: out[k].cpu().numpy() for k in KEYS}
        _compare(got, {k: ref[k] for k in KEYS}, order, w, step)
    # one rank alone: the summary of the new points through ShardedEngine
    whole = SeriesCache(eng, start, step, w, capacity=300)
    whole.ingest(body)
    names = sorted({l["__name__"] for l in d["labels"]})
    res, summary, q = whole.refresh_summary(start + 199 * step, sharding.ShardedEngine(eng), names)
    al = res["aligned"].cpu().numpy()
    gid = np.array([names.index(l["__name__"]) for l in whole.labels])
    for g in range(len(names)):
        o = O.quantiles(al[gid == g], [0.5, 0.95, 0.99])
        assert int(summary[1][g]) == o["n"]
        assert_bits_equal(q["value"][g].cpu().numpy(), np.array(o["value"]), "cache summary quantiles")
