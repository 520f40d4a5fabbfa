"""Under torchrun: where a several-rank aggregate step spends its time (CUDA events between the stages of
Engine.summary_quantiles_ranks, averaged over 50 steps, printed by rank 0).  PMX_PEER=0 forces NCCL."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.distributed as dist
from k8s_rocm_metrics_dashboard_b200 import Engine, synth, group_index, sharding, _abi
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
eng = Engine(local); dev = eng.device
S, T, M = 6144, 5760, 12
ts, val, off = synth.dense_device(S, T, M, seed=1 + rank, device=dev)
out = eng.fused(ts, val, off, synth.START_MS, synth.STEP_MS, T, 60)
x = out["aligned"]
gid = (np.arange(S) % M).astype(np.int32)
perm, goff = group_index(gid, M)
pg = (torch.as_tensor(perm, device=dev), torch.as_tensor(goff, device=dev))
seg = torch.as_tensor(gid, device=dev)
sh = sharding.ShardedEngine(eng)
use_peer = os.environ.get("PMX_PEER", "1") == "1"
peer = sh._peer(M, 3) if use_peer else None
marks = []
orig = {}
def wrap(name):
    f = getattr(eng.lib, name)
    def g(*a):
        r = f(*a)
        e = torch.cuda.Event(enable_timing=True); e.record(); marks.append((name, e))
        return r
    return f, g
class Lib:
    def __init__(self, lib): self._lib = lib
    def __getattr__(self, n):
        f = getattr(self._lib, n)
        if not n.startswith("pmx_bsel") and n != "pmx_peer_put": return f
        def g(*a):
            r = f(*a)
            e = torch.cuda.Event(enable_timing=True); e.record(); marks.append((n, e))
            return r
        return g
eng.lib = Lib(eng.lib)
acc = {}
for it in range(55):
    marks.clear()
    e0 = torch.cuda.Event(enable_timing=True); e0.record()
    eng.summary_quantiles_ranks(x, seg, M, (0.5, 0.95, 0.99), pg, dist.group.WORLD, peer=peer)
    e1 = torch.cuda.Event(enable_timing=True); e1.record()
    torch.cuda.synchronize()
    if it < 5: continue
    prev = e0
    for k, (n, e) in enumerate(marks):
        key = f"{k:02d} up to the end of {n}"
        acc[key] = acc.get(key, 0.0) + prev.elapsed_time(e); prev = e
    acc["99 rest (last exchange + finish)"] = acc.get("99 rest (last exchange + finish)", 0.0) + prev.elapsed_time(e1)
    acc["total"] = acc.get("total", 0.0) + e0.elapsed_time(e1)
if rank == 0:
    print("peer exchange" if peer is not None else "NCCL", "world", world)
    for k in sorted(acc): print(f"{acc[k] / 50 * 1000:9.1f} us  {k}")
dist.barrier(); dist.destroy_process_group()
