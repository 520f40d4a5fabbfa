import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from k8s_rocm_metrics_dashboard_b200 import Engine, synth
eng = Engine(0)
d = synth.ragged(60_000, seed=5)
ts, val, off = (torch.as_tensor(d[k], device=eng.device) for k in ("ts", "val", "off"))
for _ in range(3):
    eng.rolling_ragged(ts, val, off, 900_000)
    eng.rate_ragged(ts, val, off, 900_000, want_resets=False)
torch.cuda.synchronize()
print("ok", int(d["off"][-1]))
