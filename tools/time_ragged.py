"""Time the ragged rolling / rate kernels on config-5 shaped data generated on the device (N series)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from k8s_rocm_metrics_dashboard_b200 import Engine, synth
eng = Engine(0)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 400_000
ts, val, off = synth.ragged_device(n, seed=5, device=eng.device)
N = ts.numel()
def ev(f, k=5):
    f(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(k): f()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / k
r = ev(lambda: eng.rolling_ragged(ts, val, off, 900_000))
q = ev(lambda: eng.rate_ragged(ts, val, off, 900_000, want_resets=False))
print(f"{n} series, {N / 1e6:.1f} M samples: rolling {r:.3f} ms = {N / r / 1e6:.2f} G samples/s, rate {q:.3f} ms = {N / q / 1e6:.2f} G samples/s")
