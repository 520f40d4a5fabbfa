"""The several-GPU form of the quantile selection on one GPU: a one-rank NCCL group makes
Engine.quantiles take the per-pass protocol with its collectives (histogram all-reduce, row
all-gather, pmx_quantile_tail_filter / pmx_quantile_tail_merge) instead of the single-call path.
The N > 1 runs are tests/multi_gpu_check.py under torchrun."""
import numpy as np
import pytest

from k8s_rocm_metrics_dashboard_b200 import group_index
from oracle import prom_oracle as O
from tests.util import assert_bits_equal

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def world1():
    import torch.distributed as dist
    if dist.is_initialized():
        yield dist.group.WORLD
        return
    dist.init_process_group(backend="nccl", init_method="tcp://127.0.0.1:29577", world_size=1, rank=0)
    yield dist.group.WORLD
    dist.destroy_process_group()


def _check(eng, group, x, seg, n_seg, qs, summary):
    out = eng.quantiles(x, seg, n_seg, qs, group=group, row_order=group_index(np.asarray(seg), n_seg)[0], summary=summary)
    for g in range(n_seg):
        o = O.quantiles(x[np.asarray(seg) == g], qs)
        assert int(out["n"].cpu().numpy()[g]) == o["n"]
        assert_bits_equal(out["x_lo"].cpu().numpy()[g], np.array(o["x_lo"]), "x_lo")
        assert_bits_equal(out["x_hi"].cpu().numpy()[g], np.array(o["x_hi"]), "x_hi")
        assert_bits_equal(out["value"].cpu().numpy()[g], np.array(o["value"]), "value")


def _summary(x, seg, n_seg):
    cnt = np.zeros(n_seg, dtype=np.int64); mn = np.full(n_seg, np.nan); mx = np.full(n_seg, np.nan)
    for g in range(n_seg):
        v = x[np.asarray(seg) == g]; v = v[~np.isnan(v)]
        cnt[g] = v.size
        if v.size:
            mn[g] = v.min(); mx[g] = v.max()
    return cnt, mn, mx


def test_group_protocol_settles_segments_from_gathered_rows(eng, world1):
    rng = np.random.default_rng(31)
    x = np.concatenate([30 + 65 * rng.random((40, 900)),                      # gauge-like
                        np.cumsum(rng.random((40, 900)) * 1e4, axis=1),       # counters
                        rng.normal(size=(8, 900))])                           # both signs
    x[rng.random(x.shape) < 0.01] = np.nan
    seg = np.array([0] * 40 + [1] * 40 + [2] * 7 + [-1], dtype=np.int32)
    for summary in (None, _summary(x, seg, 3)):
        _check(eng, world1, x, seg, 3, [0.5, 0.95, 0.99], summary)
        _check(eng, world1, x, seg, 3, [0.0, 1.0], summary)


def test_group_protocol_falls_back_to_passes_when_rows_do_not_fit(eng, world1):
    """More repeated values than a row (1023 keys) or the merge buffer holds: the merge reports the
    segment as unresolved and passes 3.. run with their all-reduces."""
    rng = np.random.default_rng(32)
    x = np.full((64, 2048), 7.25)                          # 131 072 copies of one value ...
    x[:, ::5] = rng.normal(size=x[:, ::5].shape) + 7       # ... among values around it
    y = rng.normal(size=(4, 2048))
    z = np.concatenate([x, y])
    seg = np.array([0] * 64 + [1] * 4, dtype=np.int32)
    for summary in (None, _summary(z, seg, 2)):
        _check(eng, world1, z, seg, 2, [0.5, 0.9], summary)


@pytest.mark.parametrize("kind", ["uniform", "integers", "zeros", "counters", "nan_heavy", "constant"])
def test_bracket_selection_over_ranks_equals_one_sweep(eng, world1, kind):
    """Engine.summary_quantiles_ranks (five calls, four collectives) on a one-rank group gives what
    pmx_summary_quantiles and the per-pass selection give, bit for bit."""
    from tests.test_gpu_summary_quantiles import KINDS, QS, _data, _same
    t = eng.torch
    rng = np.random.default_rng(KINDS.index(kind) + 300)
    S, T, M = 64, 4100, 5
    x = _data(kind, S, T, rng)
    gid = (np.arange(S) % M).astype(np.int32)
    gid[5] = -1
    perm, goff = group_index(gid, M)
    pg = (t.as_tensor(perm, device=eng.device), t.as_tensor(goff, device=eng.device))
    dx, seg = t.as_tensor(x, device=eng.device), t.as_tensor(gid, device=eng.device)
    summ, q, failed = eng.summary_quantiles_ranks(dx, seg, M, QS, pg, world1)
    assert int(failed.item()) == 0
    ref_s = eng.group_reduce(dx, None, M, perm_goff=pg)
    ref_q = eng.quantiles(dx, seg, M, list(QS), row_order=pg[0], summary=(ref_s[1], ref_s[2], ref_s[3]))
    _same(summ, q, ref_s, ref_q, f"ranks {kind}")
    s1, q1, f1 = eng.summary_quantiles(dx, seg, M, QS, pg)
    assert int(f1.item()) == 0
    _same(summ, q, s1, q1, f"ranks vs one sweep {kind}")


def test_peer_memory_exchange_on_a_one_rank_group(eng, world1):
    """sharding.PeerExchange (stores into symmetric memory + device-side barrier instead of NCCL all-gathers)
    under Engine.summary_quantiles_ranks: same results as with the NCCL calls.  One rank here; the N > 1 runs
    are tests/multi_gpu_check.py and bench.py's parity_check."""
    from k8s_rocm_metrics_dashboard_b200 import sharding
    from tests.test_gpu_summary_quantiles import QS, _data, _same
    t = eng.torch
    rng = np.random.default_rng(900)
    S, T, M = 48, 3000, 4
    x = _data("uniform", S, T, rng)
    gid = (np.arange(S) % M).astype(np.int32)
    perm, goff = group_index(gid, M)
    pg = (t.as_tensor(perm, device=eng.device), t.as_tensor(goff, device=eng.device))
    dx, seg = t.as_tensor(x, device=eng.device), t.as_tensor(gid, device=eng.device)
    try:
        peer = sharding.PeerExchange(eng, world1, eng.bsel_region_bytes(M, len(QS)))
    except Exception as e:  # noqa: BLE001
        pytest.skip(f"symmetric memory not available here: {e}")
    for _ in range(3):                                             # the regions are reused step after step
        s1, q1, f1 = eng.summary_quantiles_ranks(dx, seg, M, QS, pg, world1, peer=peer)
    s0, q0, f0 = eng.summary_quantiles_ranks(dx, seg, M, QS, pg, world1)
    assert int(f1.item()) == 0 and int(f0.item()) == 0
    _same(s1, q1, s0, q0, "peer exchange vs NCCL")


def test_stage_graphs_of_the_several_rank_selection(eng, world1):
    """stage_graphs=True: the launches between two exchanges replayed from CUDA graphs (captured on the third
    call) give the eager results, also after the matrix was refilled in place."""
    from tests.test_gpu_summary_quantiles import QS, _data, _same
    t = eng.torch
    S, T, M = 48, 3000, 4
    gid = (np.arange(S) % M).astype(np.int32)
    perm, goff = group_index(gid, M)
    pg = (t.as_tensor(perm, device=eng.device), t.as_tensor(goff, device=eng.device))
    seg = t.as_tensor(gid, device=eng.device)
    dx = t.empty((S, T), dtype=t.float64, device=eng.device)
    for it, kind in enumerate(["uniform", "integers", "counters", "zeros", "uniform"]):
        x = _data(kind, S, T, np.random.default_rng(950 + it))
        dx.copy_(t.as_tensor(x, device=eng.device))
        s1, q1, f1 = eng.summary_quantiles_ranks(dx, seg, M, QS, pg, world1, stage_graphs=True)
        got_s = [v.clone() for v in s1]; got_q = {k: v.clone() for k, v in q1.items()}
        s0, q0, f0 = eng.summary_quantiles_ranks(dx, seg, M, QS, pg, world1)
        assert int(f1.item()) == 0 and int(f0.item()) == 0
        _same(got_s, got_q, s0, q0, f"stage graphs, call {it} ({kind})")
    graphs = [v for k, v in eng._bsel[(M, len(QS), 1)]["graphs"].items() if k[0] != "res"]
    assert graphs and all(g[1] is not None for g in graphs), "a stage was not captured"
