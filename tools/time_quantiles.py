"""Time the exact-quantile selection (p50/p95/p99 per metric) and the group reduction on config-2 shaped data."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from k8s_rocm_metrics_dashboard_b200 import Engine, synth, group_index
eng = Engine(0); dev = eng.device
S, T = 6144, 5760
ts, val, off = synth.dense_device(S, T, 12, seed=1, device=dev)
out = eng.fused(ts, val, off, synth.START_MS, synth.STEP_MS, T, 60, 300000)
al = out["aligned"]
seg = (np.arange(S) % 12).astype(np.int32)
pg = group_index(seg, 12)
def timeit(f, n=20):
    for _ in range(3): f()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): f()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n
print("quantiles  %.3f ms" % timeit(lambda: eng.quantiles(al, seg, 12, [0.5, 0.95, 0.99], row_order=pg[0])))
print("group_reduce %.3f ms" % timeit(lambda: eng.group_reduce(al, None, 12, perm_goff=pg)))
