"""The several-rank form of the summary on a one-rank NCCL group, config-2 shaped (target for ncu)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.distributed as dist
from k8s_rocm_metrics_dashboard_b200 import Engine, synth, group_index
dist.init_process_group(backend="nccl", init_method="tcp://127.0.0.1:29579", world_size=1, rank=0)
eng = Engine(0); dev = eng.device
S, T, M = 6144, 5760, 12
ts, val, off = synth.dense_device(S, T, M, seed=1, device=dev)
out = eng.fused(ts, val, off, synth.START_MS, synth.STEP_MS, T, 60)
gid = (np.arange(S) % M).astype(np.int32)
perm, goff = group_index(gid, M)
pg = (torch.as_tensor(perm, device=dev), torch.as_tensor(goff, device=dev))
seg = torch.as_tensor(gid, device=dev)
for _ in range(3):
    s, q, f = eng.summary_quantiles_ranks(out["aligned"], seg, M, (0.5, 0.95, 0.99), pg, dist.group.WORLD)
torch.cuda.synchronize()
print("ok", int(f.item()))
dist.destroy_process_group()
