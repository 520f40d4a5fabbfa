timeout 900 python -m pytest tests/test_gpu_parse.py tests/test_gpu_tokenize_device.py tests/test_gpu_instant.py -x -q 2>&1 | tail -6
python - <<'PY'
import sys, time, ctypes as C
sys.path.insert(0, '.')
import numpy as np, torch
from k8s_rocm_metrics_dashboard_b200 import Engine, synth, tokenize, _abi
eng = Engine(0); dev = eng.device
d1 = synth.dense(4, 8, 12, 5760, seed=3, nan_frac=1e-3, kinds="ttk")
body = synth.to_matrix_json(d1["labels"], d1["ts"], d1["val"], d1["off"])
tok = tokenize(body)
n1 = tok.n_samples
text = torch.from_numpy(tok.flat.copy()).to(dev)
voff = torch.from_numpy(tok.val_off.view(np.int32).copy()).to(dev); toff = torch.from_numpy(tok.ts_off.view(np.int32).copy()).to(dev)
pv, pt = torch.empty(n1, dtype=torch.float64, device=dev), torch.empty(n1, dtype=torch.int64, device=dev)
bad = torch.empty(2, dtype=torch.int64, device=dev)
p = lambda x: C.c_void_p(x.data_ptr()); st = lambda: C.c_void_p(torch.cuda.current_stream().cuda_stream)
def parse(): _abi.check(eng.lib.pmx_parse_samples(eng.ctx, p(text), text.numel(), p(toff), p(voff), n1, p(pt), p(pv), p(bad), st()))
for _ in range(3): parse()
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(10): parse()
b.record(); torch.cuda.synchronize()
ms = a.elapsed_time(b) / 10
print(f"K1 parse: {n1/1e6:.2f} M samples, body {len(body)/1e6:.0f} MB, {ms:.3f} ms, {n1/ms/1e6:.1f} G samples/s, {len(body)/ms/1e6:.0f} GB/s of text")
assert np.array_equal(pt.cpu().numpy(), d1["ts"]) and np.array_equal(pv.cpu().numpy().view(np.uint64), d1["val"].view(np.uint64))
print("bits ok", bad.cpu().tolist())
PY
