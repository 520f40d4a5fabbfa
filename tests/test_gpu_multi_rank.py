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
