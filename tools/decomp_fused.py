import sys, ctypes as C, torch, numpy as np
import os; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from k8s_rocm_metrics_dashboard_b200 import Engine, synth, _abi
eng = Engine(0); dev = eng.device
S, T, W = 6144, 5760, 60
ts, val, off = synth.dense_device(S, T, 12, seed=1, device=dev)
outs = [torch.empty((S, T), dtype=torch.float64, device=dev) for _ in range(5)]
p = lambda x: C.c_void_p(x.data_ptr()) if x is not None else None
def run(mask, w=W):
    o = [outs[i] if mask[i] else None for i in range(5)]
    def f():
        _abi.check(eng.lib.pmx_fused_align_rolling_rate(eng.ctx, p(ts), p(val), p(off), S, synth.START_MS, synth.STEP_MS, T, 300000, w, 1,
                   p(o[0]), p(o[1]), p(o[2]), p(o[3]), p(o[4]), C.c_void_p(torch.cuda.current_stream().cuda_stream)))
    for _ in range(3): f()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(10): f()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / 10
for name, mask in [("aligned only", (1,0,0,0,0)), ("aligned+min", (1,1,0,0,0)), ("aligned+min+max+mean", (1,1,1,1,0)), ("aligned+rate", (1,0,0,0,1)), ("rate only", (0,0,0,0,1)), ("all", (1,1,1,1,1))]:
    ms = run(mask)
    nb = 16 + 8 * sum(mask)
    print(f"{name:24s} {ms:7.3f} ms   {S*T*nb/ms/1e6:7.0f} GB/s")
for w in (16, 60, 120, 240, 480):
    print("all, w =", w, f"{run((1,1,1,1,1), w):7.3f} ms")
