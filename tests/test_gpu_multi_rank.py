"""GPU: the sharded aggregate step against the oracle on the union of all shards
(tests/multi_gpu_check.py): in this process at world size 1, and under torchrun with one rank per
GPU when the box has more than one."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_parity_world1(eng):
    from k8s_rocm_metrics_dashboard_b200 import sharding
    from tests.multi_gpu_check import parity_check
    res = parity_check(eng, sharding.ShardedEngine(eng), 0, 1, n_nodes=2, n_steps=700)
    assert res["ok"]


def test_parity_all_gpus_of_the_box():
    import torch
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("one GPU visible: the multi-rank run needs at least two")
    n = 2 if n < 4 else (4 if n < 8 else 8)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr", "127.0.0.1",
           "--master-port", "29541", os.path.join(ROOT, "tests", "multi_gpu_check.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0 and "multi_gpu_check ok" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]


def test_graph_replay_of_the_summary_equals_the_eager_step(eng):
    """aggregate_step(graph=True): the launches behind the fused kernel replayed from a CUDA graph give
    the bits of the eager step, also after the input buffers were refilled with other data."""
    import numpy as np
    import torch as t
    from k8s_rocm_metrics_dashboard_b200 import group_index, sharding, synth
    S, T, M = 96, 1500, 12
    sh = sharding.ShardedEngine(eng, local_only=True)
    sh.ONE_SWEEP_MIN_VALUES_PER_SEGMENT = 0                        # these segments are short: take the one-sweep path anyway
    gid = (np.arange(S) % M).astype(np.int32)
    seg = t.as_tensor(gid, device=eng.device)
    perm, goff = group_index(gid, M)
    pg = (t.as_tensor(perm, device=eng.device), t.as_tensor(goff, device=eng.device))
    out = {k: t.empty((S, T), dtype=t.float64, device=eng.device) for k in ("aligned", "roll_min", "roll_max", "roll_mean", "rate")}
    ts, val, off = synth.dense_device(S, T, M, seed=1, device=eng.device)
    for seed in (1, 2, 3):
        ts2, val2, _ = synth.dense_device(S, T, M, seed=seed, device=eng.device)
        ts.copy_(ts2); val.copy_(val2)
        _, s_g, q_g = sh.aggregate_step(ts, val, off, synth.START_MS, synth.STEP_MS, T, 60, seg, M, pg, out=out, graph=True)
        got = [x.clone() for x in s_g] + [q_g["value"].clone(), q_g["x_lo"].clone()]
        _, s_e, q_e = sh.aggregate_step(ts, val, off, synth.START_MS, synth.STEP_MS, T, 60, seg, M, pg, out=out)
        exp = list(s_e) + [q_e["value"], q_e["x_lo"]]
        for a, b in zip(got, exp):
            assert t.equal(a.view(t.int64) if a.dtype == t.float64 else a, b.view(t.int64) if b.dtype == t.float64 else b)
    assert any(g[0] is not None for g in sh._graphs.values()), "the capture fell back to the eager path"
