"""Run the fused kernel a few times on config-2 shaped data (target for ncu)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from k8s_rocm_metrics_dashboard_b200 import Engine, synth
eng = Engine(0); dev = eng.device
S, T = 6144, 5760
W = int(sys.argv[1]) if len(sys.argv) > 1 else 60
ts, val, off = synth.dense_device(S, T, 12, seed=1, device=dev)
out = {k: torch.empty((S, T), dtype=torch.float64, device=dev) for k in ("aligned", "roll_min", "roll_max", "roll_mean", "rate")}
for _ in range(4):
    eng.fused(ts, val, off, synth.START_MS, synth.STEP_MS, T, W, out=out)
torch.cuda.synchronize()
print("ok")
