import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from k8s_rocm_metrics_dashboard_b200 import Engine, synth, group_index
eng = Engine(0); dev = eng.device; t = torch
S, T, M = 16384, 40320, 16
ts, val, off = synth.dense_device(S, T, M, seed=3000, device=dev)
al, _ = eng.align(ts, val, off, synth.START_MS, synth.STEP_MS, T, want_idx=False)
del ts, val
gid = (np.arange(S) % M).astype(np.int32)
perm, goff = group_index(gid, M)
pg = (t.as_tensor(perm, device=dev), t.as_tensor(goff, device=dev))
seg = t.as_tensor(gid, device=dev)
summ, q, failed = eng.summary_quantiles(al, seg, M, (0.5, 0.95, 0.99), pg)
print("failed", failed.item())
print(q["value"])
print(q["lo"][:4], summ[1][:4])
ref_s = eng.group_reduce(al, None, M, perm_goff=pg)
ref_q = eng.quantiles(al, seg, M, [0.5, 0.95, 0.99], row_order=pg[0], summary=(ref_s[1], ref_s[2], ref_s[3]))
print(ref_q["value"])
print("equal", t.equal(ref_q["value"].view(t.int64), q["value"].view(t.int64)))
