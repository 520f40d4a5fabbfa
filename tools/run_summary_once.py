"""Run the summary part of the aggregate step (series partials, fold, quantile passes) a few times on
config-2 shaped data (target for ncu)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from k8s_rocm_metrics_dashboard_b200 import Engine, synth, group_index, sharding
eng = Engine(0); dev = eng.device
S, T, M = 6144, 5760, 12
ts, val, off = synth.dense_device(S, T, M, seed=1, device=dev)
out = eng.fused(ts, val, off, synth.START_MS, synth.STEP_MS, T, 60)
gid = (np.arange(S) % M).astype(np.int32)
perm, goff = group_index(gid, M)
pg = (torch.as_tensor(perm, device=dev), torch.as_tensor(goff, device=dev))
seg = torch.as_tensor(gid, device=dev)
sh = sharding.ShardedEngine(eng, local_only=True)
for _ in range(3):
    sh.summarize(out["aligned"], seg, M, pg, (0.5, 0.95, 0.99), defer_check=True)
torch.cuda.synchronize()
print("ok", eng.deferred_ok(wait=True))
