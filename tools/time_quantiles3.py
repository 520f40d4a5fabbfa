"""Time the selection pieces on the config-3 per-GPU shard."""
import os, sys, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from k8s_rocm_metrics_dashboard_b200 import Engine, synth, group_index, _abi
eng = Engine(0); dev = eng.device
S, T, M = 16384, 40320, 16
if len(sys.argv) > 1: S, T, M = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
ts, val, off = synth.dense_device(S, T, M, seed=3000, device=dev)
al, _ = eng.align(ts, val, off, synth.START_MS, synth.STEP_MS, T, want_idx=False)
del ts, val
seg = (np.arange(S) % M).astype(np.int32)
pg = group_index(seg, M)
pg = (torch.as_tensor(pg[0], device=dev), torch.as_tensor(pg[1], device=dev))
dseg = torch.as_tensor(seg, device=dev)
def ev(f, n=3):
    f(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): f()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n
sm, cnt, mn, mx, mean = eng.group_reduce(al, None, M, perm_goff=pg)
print("group_reduce %.3f ms" % ev(lambda: eng.group_reduce(al, None, M, perm_goff=pg)))
print("quantiles (summary) %.3f ms" % ev(lambda: eng.quantiles(al, dseg, M, [0.5, 0.95, 0.99], row_order=pg[0], summary=(cnt, mn, mx))))
print("quantiles (select)  %.3f ms" % ev(lambda: eng.quantiles(al, dseg, M, [0.5, 0.95, 0.99], row_order=pg[0])))
# pieces
t = torch
state, hist = eng.quantile_buffers(M, 3)
dq = t.tensor([0.5, 0.95, 0.99], dtype=t.float64, device=dev)
p = lambda x: C.c_void_p(x.data_ptr()) if x is not None else None
st = lambda: C.c_void_p(t.cuda.current_stream().cuda_stream)
def piece(name, f):
    torch.cuda.synchronize(); a, b = t.cuda.Event(enable_timing=True), t.cuda.Event(enable_timing=True)
    a.record(); f(); b.record(); torch.cuda.synchronize(); print(f"  {name:12s} {a.elapsed_time(b):9.3f} ms")
piece("init", lambda: _abi.check(eng.lib.pmx_quantile_init(eng.ctx, p(cnt), p(mn), p(mx), M, p(dq), 3, p(state), st())))
for ps in (1, 2):
    piece(f"hist{ps}", lambda: _abi.check(eng.lib.pmx_quantile_hist(eng.ctx, p(al), S, T, p(dseg), M, p(pg[0]), pg[0].numel(), 3, ps, p(state), p(hist), st())))
    piece(f"pick{ps}", lambda: _abi.check(eng.lib.pmx_quantile_pick(eng.ctx, p(hist), M, p(dq), 3, ps, p(state), st())))
piece("tail", lambda: _abi.check(eng.lib.pmx_quantile_tail(eng.ctx, p(al), S, T, p(dseg), M, 3, 3, p(state), st())))
print("state[:,1] resolved bits:", [(int(x) >> 40) & 1 for x in state[:, 1].cpu()])
