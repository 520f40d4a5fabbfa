"""Development check: run a set of fused / rolling / rate calls, save or compare the raw outputs.
   PMX_WINDOW_IMPL=f python tools/dev_compare.py save ; python tools/dev_compare.py cmp"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from k8s_rocm_metrics_dashboard_b200 import Engine, synth
mode = sys.argv[1]
eng = Engine(0); dev = eng.device
res = {}
def put(name, out):
    for k, v in out.items():
        if v is not None: res[f"{name}/{k}"] = v.cpu()
for w in (9, 12, 16, 60, 61, 63, 64, 120, 250, 480):
    ts, val, off = synth.dense_device(96, 5760, 12, seed=w, device=dev)
    put(f"dense_w{w}", eng.fused(ts, val, off, synth.START_MS, synth.STEP_MS, 5760, w))
for T in (1021, 1022, 1023, 4097):
    d = synth.dense(1, 3, 3, T, seed=T, nan_frac=0.02, kinds="ttk")
    put(f"rowalign_T{T}", eng.fused(d["ts"], d["val"], d["off"], d["start_ms"], d["step_ms"], T, 60))
for w in (9, 60, 64, 250):
    d = synth.ragged(120, seed=w, mean_len=500, nan_frac=0.03, stale_frac=0.02, gap_frac=0.04)
    t0 = int(d["ts"].min())
    put(f"ragged_w{w}", eng.fused(d["ts"], d["val"], d["off"], t0 - 45_000, 15_000, 2600 + w, w))
rng = np.random.default_rng(1)
x = np.round(rng.uniform(-50, 100, (7, 5000)), 3); x[rng.random(x.shape) < 0.03] = np.nan
x[0, 100] = np.inf; x[1, 200] = -np.inf; x[2, :2100] = np.nan
for w in (9, 60, 240):
    mn, mx, mean, cnt = eng.rolling(x, w)
    inc, rt, _ = eng.rate(x, w, 15000, want_resets=False)
    put(f"dense_roll_w{w}", dict(mn=mn, mx=mx, mean=mean, cnt=cnt.double(), inc=inc, rt=rt))
    mn, mx, mean, cnt = eng.rolling(x, w, min_periods=w // 2)
    put(f"dense_roll_mp_w{w}", dict(mn=mn, mx=mx, mean=mean, cnt=cnt.double()))
torch.cuda.synchronize()
path = "/tmp/dev_compare.pt"
if mode == "save":
    torch.save(res, path); print("saved", len(res))
else:
    ref = torch.load(path)
    bad = 0
    for k, v in res.items():
        r = ref[k]
        same = (v.view(torch.int64) == r.view(torch.int64)) | (torch.isnan(v) & torch.isnan(r))
        n = int((~same).sum())
        if n:
            err = ((v - r).abs() / r.abs().clamp(min=1e-300))[~same]
            nanmis = int((torch.isnan(v) != torch.isnan(r)).sum())
            print(f"{k:32s} differ {n:8d} of {v.numel():9d}  max rel {float(err[~torch.isnan(err)].max()) if (~torch.isnan(err)).any() else float('nan'):.3e}  nan-mismatch {nanmis}")
            if ("rate" in k or k.endswith("/rt") or k.endswith("/inc")) and nanmis == 0 and float(err[~torch.isnan(err)].max()) < 1e-9: continue
            bad += 1
    print("compare done; suspicious outputs:", bad)
