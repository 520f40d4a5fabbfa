"""The several-rank form of the one-sweep selection (pmx_bsel_*) with the ranks EMULATED on one GPU: every
"rank" is an Engine of its own (its own ctx and scratch) holding a shard of the series, and the four
collectives are done by hand (concatenate = all-gather, add = all-reduce).  Every rank must return the
summary and the exact quantiles of the UNION, bit for bit what one sweep over the union gives -- also
with an empty rank and with shards of very different distributions."""
import ctypes as C

import numpy as np
import pytest

from k8s_rocm_metrics_dashboard_b200 import Engine, _abi, group_index
from oracle import prom_oracle as O
from tests.test_gpu_summary_quantiles import QS, _data, _same
from tests.util import assert_bits_equal

pytestmark = pytest.mark.gpu


def _run_ranks(engines, shards, gids, M, qs, gathered_hist=False):
    """shards[r]: [S_r, T] host matrix of rank r (S_r may be 0); returns per rank (summary, q, failed)."""
    t = engines[0].torch
    W, n_q = len(engines), len(qs)
    w = [C.c_int(0) for _ in range(3)]
    _abi.check(engines[0].lib.pmx_bsel_widths(C.byref(w[0]), C.byref(w[1]), C.byref(w[2])))
    bw, pw, rw = (int(v.value) for v in w)
    st = []
    for e, x, gid in zip(engines, shards, gids):
        dev = e.device
        dx = t.as_tensor(np.ascontiguousarray(x), device=dev).reshape(len(gid), x.shape[1])
        perm, goff = group_index(gid, M)
        st.append({"x": dx, "seg": t.as_tensor(gid, device=dev), "perm": t.as_tensor(perm, device=dev),
                   "goff": t.as_tensor(goff, device=dev), "q": e._quantile_levels(qs),
                   "bounds": t.empty(M * bw, dtype=t.float64, device=dev),
                   "packed": t.empty(M * (pw + n_q * 1024), dtype=t.int64, device=dev),
                   "hist2": t.empty(M * n_q * 2 * 2048, dtype=t.int32, device=dev),
                   "rows": t.empty(M * n_q * rw, dtype=t.int64, device=dev)})
    T = shards[0].shape[1]
    p = lambda e, a: e._ptr(a)
    for e, s in zip(engines, st):
        _abi.check(e.lib.pmx_bsel_begin(e.ctx, p(e, s["x"]), s["x"].shape[0], T, p(e, s["seg"]), M, p(e, s["perm"]), p(e, s["goff"]),
                                        p(e, s["q"]), n_q, p(e, s["bounds"]), e._stream()))
    bounds_all = t.cat([s["bounds"] for s in st])
    for e, s in zip(engines, st):
        _abi.check(e.lib.pmx_bsel_sweep(e.ctx, p(e, s["x"]), s["x"].shape[0], T, p(e, s["seg"]), M, p(e, s["perm"]), p(e, s["goff"]),
                                        n_q, p(e, bounds_all), W, p(e, s["packed"]), e._stream()))
    packed_all = t.cat([s["packed"] for s in st])
    res = []
    for e, s in zip(engines, st):
        sm, mn, mx, mean = (t.empty(M, dtype=t.float64, device=e.device) for _ in range(4))
        cnt = t.empty(M, dtype=t.int64, device=e.device)
        _abi.check(e.lib.pmx_bsel_hist(e.ctx, s["x"].shape[0], T, M, p(e, s["q"]), n_q, p(e, packed_all), W, p(e, sm), p(e, cnt),
                                       p(e, mn), p(e, mx), p(e, mean), p(e, s["hist2"]), e._stream()))
        res.append([(sm, cnt, mn, mx, mean)])
    if gathered_hist:                                              # all-gather: the rows call adds the copies
        hist2, n_hist = t.cat([s["hist2"] for s in st]), W
    else:                                                          # all-reduce
        hist2, n_hist = st[0]["hist2"].clone(), 1
        for s in st[1:]:
            hist2 += s["hist2"]
    for e, s in zip(engines, st):
        _abi.check(e.lib.pmx_bsel_rows(e.ctx, s["x"].shape[0], T, M, n_q, p(e, hist2), n_hist, p(e, s["rows"]), e._stream()))
    rows_all = t.cat([s["rows"] for s in st])
    for e, s, r in zip(engines, st, res):
        outs = {"n": t.empty(M, dtype=t.int64, device=e.device), "lo": t.empty((M, n_q), dtype=t.int64, device=e.device),
                "hi": t.empty((M, n_q), dtype=t.int64, device=e.device), "x_lo": t.empty((M, n_q), dtype=t.float64, device=e.device),
                "x_hi": t.empty((M, n_q), dtype=t.float64, device=e.device), "value": t.empty((M, n_q), dtype=t.float64, device=e.device)}
        failed = t.empty(1, dtype=t.int32, device=e.device)
        _abi.check(e.lib.pmx_bsel_finish(e.ctx, s["x"].shape[0], T, M, n_q, p(e, rows_all), W, p(e, r[0][1]), p(e, outs["n"]),
                                         p(e, outs["lo"]), p(e, outs["hi"]), p(e, outs["x_lo"]), p(e, outs["x_hi"]), p(e, outs["value"]),
                                         p(e, failed), e._stream()))
        r.extend([outs, int(failed.item())])
    return res


@pytest.fixture(scope="module")
def ranks(eng):
    return [eng, Engine(0), Engine(0)]


@pytest.mark.parametrize("kind", ["uniform", "integers", "counters", "zeros", "nan_heavy", "wide"])
def test_three_emulated_ranks_one_of_them_empty(eng, ranks, kind):
    rng = np.random.default_rng(500 + len(kind))
    S, T, M = 90, 3500, 5
    x = _data(kind, S, T, rng)
    gid = (np.arange(S) % M).astype(np.int32)
    owner = np.where(np.arange(S) % 3 == 0, 0, 2)                  # rank 1 holds nothing; ranks 0 and 2 share 1 : 2
    shards = [x[owner == r] for r in range(3)]
    gids = [gid[owner == r] for r in range(3)]
    res = _run_ranks(ranks, shards, gids, M, QS)
    t = eng.torch
    perm, goff = group_index(gid, M)
    pg = (t.as_tensor(perm, device=eng.device), t.as_tensor(goff, device=eng.device))
    dx, seg = t.as_tensor(x, device=eng.device), t.as_tensor(gid, device=eng.device)
    ref_s = eng.group_reduce(dx, None, M, perm_goff=pg)
    ref_q = eng.quantiles(dx, seg, M, list(QS), row_order=pg[0], summary=(ref_s[1], ref_s[2], ref_s[3]))
    for r, (summ, q, failed) in enumerate(res):
        assert failed == 0, f"rank {r}: {failed} brackets failed"
        # count / min / max and every quantile exact; sums are folded per rank, then in rank order: 1e-12
        assert np.array_equal(summ[1].cpu().numpy(), ref_s[1].cpu().numpy())
        assert_bits_equal(summ[2].cpu().numpy(), ref_s[2].cpu().numpy(), "min")
        assert_bits_equal(summ[3].cpu().numpy(), ref_s[3].cpu().numpy(), "max")
        a, b = summ[0].cpu().numpy(), ref_s[0].cpu().numpy()
        assert np.all(np.abs(a - b) <= 1e-12 * np.maximum(np.abs(b), 1e-300) + 0.0) or kind in ("infs",)
        for k in ("n", "lo", "hi"):
            assert np.array_equal(q[k].cpu().numpy(), ref_q[k].cpu().numpy()), f"rank {r}: {k}"
        for k in ("x_lo", "x_hi", "value"):
            assert_bits_equal(q[k].cpu().numpy(), ref_q[k].cpu().numpy(), f"rank {r} {kind}: {k}")


def test_ranks_with_different_distributions(eng, ranks):
    """Rank 0 holds low values, rank 2 high ones: the hull of their brackets is wide, the result still exact."""
    rng = np.random.default_rng(77)
    T, M = 6000, 2
    lo = rng.normal(10.0, 1.0, (16, T)); hi = rng.normal(50.0, 5.0, (24, T))
    shards = [lo, np.zeros((0, T)), hi]
    gids = [(np.arange(16) % M).astype(np.int32), np.zeros(0, dtype=np.int32), (np.arange(24) % M).astype(np.int32)]
    res = _run_ranks(ranks, shards, gids, M, QS, gathered_hist=True)
    x = np.concatenate([lo, hi]); gid = np.concatenate([gids[0], gids[2]])
    for summ, q, failed in res:
        assert failed == 0
        for g in range(M):
            o = O.quantiles(x[gid == g], QS)
            assert int(q["n"][g]) == o["n"]
            assert_bits_equal(q["value"][g].cpu().numpy(), np.array(o["value"]), "value vs oracle")
