"""Where the end-to-end-from-text step spends its time (pieces timed with a device sync after each)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from k8s_rocm_metrics_dashboard_b200 import Engine, synth, group_index
eng = Engine(0)
d = synth.dense(8, 8, 12, 5760, seed=7, nan_frac=1e-3, kinds="ttk")
body = synth.to_matrix_json(d["labels"], d["ts"], d["val"], d["off"])
pinned = torch.frombuffer(bytearray(body), dtype=torch.uint8).pin_memory()
def T(name, f, acc):
    torch.cuda.synchronize(); t0 = time.perf_counter(); r = f(); torch.cuda.synchronize()
    acc[name] = acc.get(name, 0.0) + (time.perf_counter() - t0) * 1e3
    return r
acc = {}
out = None
for it in range(6):
    if it == 1: acc = {}
    tok = T("tokenize_device (H2D + 3 kernels + status/spans D2H)", lambda: eng.tokenize_device(pinned), acc)
    labels = T("labels (pmx_labels_intern)", lambda: tok.labels(), acc)
    def groups():
        names = sorted({l["__name__"] for l in labels})
        gid = np.array([names.index(l["__name__"]) for l in labels], dtype=np.int32)
        return names, gid, group_index(gid, len(names))
    names, gid, pg = T("group ids (host)", groups, acc)
    ts, val = T("parse", lambda: eng.parse(tok), acc)
    out = T("fused", lambda: eng.fused(ts, val, tok.series_off, synth.START_MS, synth.STEP_MS, 5760, 60, 300000, out=out), acc)
    def summ():
        sm, cnt, mn, mx, mean = eng.group_reduce(out["aligned"], None, len(names), perm_goff=pg)
        q = eng.quantiles(out["aligned"], gid, len(names), [0.5, 0.95, 0.99], row_order=pg[0], summary=(cnt, mn, mx))
        return torch.stack([sm, mean, q["value"][:, 0], q["value"][:, 1], q["value"][:, 2]]).cpu()
    T("summary + quantiles + D2H", summ, acc)
for k, v in acc.items(): print(f"{v / 5:8.3f} ms  {k}")
print(f"{sum(acc.values()) / 5:8.3f} ms  total; body {len(body) / 1e6:.1f} MB, {len(d['ts'])} samples")
