"""GPU: pmx_group_reduce_noalloc, the entry point for callers that cannot allocate device memory
at call time (all memory already owned by another allocator, or no tolerance for the device-wide
synchronisation an allocation implies).  Called through the C ABI, twice in a row on different
inputs, and compared with pmx_group_reduce, which keeps its partials in allocated ctx scratch.
Collected last on purpose: a -x run has then already run every other test."""
import ctypes as C

import numpy as np
import pytest

from k8s_rocm_metrics_dashboard_b200 import _abi, group_index

pytestmark = pytest.mark.gpu


def _noalloc(eng, x, perm, goff, G):
    t = eng.torch
    dx = t.as_tensor(x, dtype=t.float64, device=eng.device)
    dp = t.as_tensor(perm, dtype=t.int32, device=eng.device)
    dg = t.as_tensor(goff, dtype=t.int64, device=eng.device)
    sm, mn, mx, mean = (t.empty(G, dtype=t.float64, device=eng.device) for _ in range(4))
    cnt = t.empty(G, dtype=t.int64, device=eng.device)
    p = lambda v: C.c_void_p(v.data_ptr())
    with t.cuda.device(eng.device):
        _abi.check(eng.lib.pmx_group_reduce_noalloc(p(dx), x.shape[0], x.shape[1], p(dp), p(dg), G, p(sm), p(cnt),
                                                    p(mn), p(mx), p(mean),
                                                    C.c_void_p(t.cuda.current_stream(eng.device).cuda_stream)))
    t.cuda.synchronize()
    return [v.cpu().numpy() for v in (sm, cnt, mn, mx, mean)]


def _same(a, b, tol):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    assert np.array_equal(np.isnan(a), np.isnan(b))
    ok = ~np.isnan(b)
    assert np.all(np.abs(a[ok] - b[ok]) <= tol * np.maximum(np.abs(b[ok]), 1.0))


def test_noalloc_matches_group_reduce_on_two_calls_in_a_row(eng):
    rng = np.random.default_rng(21)
    for S, T, G in [(700, 513, 9), (96, 2000, 5)]:             # second call reuses (and overwrites) the array
        x = rng.uniform(-5, 10, (S, T)); x[rng.random(x.shape) < 0.1] = np.nan
        x[3] = np.nan
        gid = rng.integers(-1, G, S).astype(np.int32); gid[0] = G - 1
        perm, goff = group_index(gid, G)
        before = eng.scratch_bytes()
        got = _noalloc(eng, x, perm, goff, G)
        assert eng.scratch_bytes() == before                   # the ctx allocated nothing for it
        ref = [v.cpu().numpy() for v in eng.group_reduce(x, gid, G)]
        assert np.array_equal(got[1], ref[1])                  # counts
        _same(got[2], ref[2], 0.0); _same(got[3], ref[3], 0.0)  # min / max exact
        _same(got[0], ref[0], 1e-12); _same(got[4], ref[4], 1e-12)
        # ... and against the oracle, not only against the sibling entry point
        from oracle import prom_oracle as O
        osm, ocnt, omn, omx, omean = O.group_reduce(x, gid, G)
        assert np.array_equal(got[1], ocnt)
        _same(got[2], omn, 0.0); _same(got[3], omx, 0.0)
        _same(got[0], osm, 1e-12); _same(got[4], omean, 1e-12)


def test_noalloc_rejects_more_series_than_its_fixed_array(eng):
    rc = eng.lib.pmx_group_reduce_noalloc(None, 65537, 1, None, None, 1, None, None, None, None, None, None)
    assert rc == _abi.E_LIMIT
    assert b"65536" in eng.lib.pmx_last_error()
