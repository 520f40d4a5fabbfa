"""Run the device tokenizer (X1) a few times on a config-2 shaped body (target for ncu)."""
import os, sys, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from k8s_rocm_metrics_dashboard_b200 import Engine, synth, _abi
eng = Engine(0); dev = eng.device
d1 = synth.dense(8, 8, 12, 5760, seed=3, nan_frac=1e-3, kinds="ttk")
body = synth.to_matrix_json(d1["labels"], d1["ts"], d1["val"], d1["off"])
dtok = eng.tokenize_device(body)
n1 = dtok.n_samples
cap_s, cap_r = len(body) // 8 + 1, len(body) // 24 + 1
b_ts, b_val = torch.empty(cap_s, dtype=torch.int32, device=dev), torch.empty(cap_s, dtype=torch.int32, device=dev)
b_so = torch.empty(cap_r + 1, dtype=torch.int64, device=dev)
b_mo, b_ml = torch.empty(cap_r, dtype=torch.int32, device=dev), torch.empty(cap_r, dtype=torch.int32, device=dev)
b_st = torch.empty(4, dtype=torch.int64, device=dev)
pp = lambda x: C.c_void_p(x.data_ptr())
def run():
    _abi.check(eng.lib.pmx_tokenize_device(eng.ctx, pp(dtok.d_text), len(body), pp(b_ts), pp(b_val), cap_s, pp(b_so), pp(b_mo),
                                           pp(b_ml), cap_r, pp(b_st), C.c_void_p(torch.cuda.current_stream().cuda_stream)))
for _ in range(3): run()
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(10): run()
b.record(); torch.cuda.synchronize()
ms = a.elapsed_time(b) / 10
st = b_st.cpu().tolist()
print(f"X1 device tokenizer: {n1/1e6:.2f} M samples, body {len(body)/1e6:.0f} MB, {ms:.3f} ms, {len(body)/ms/1e6:.0f} GB/s of text; status {st}")
