"""Run under torchrun on >= 2 GPUs (not collected by pytest):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
        --master-port 29533 tests/multi_gpu_check.py

Every rank takes its label-hash shard of a small config, runs the fused kernel locally and the
NCCL-combined per-metric summary and exact quantiles; rank 0 compares with the oracle on the union."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from k8s_rocm_metrics_dashboard_b200 import Engine, group_index, sharding, synth  # noqa: E402
from oracle import prom_oracle as O  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    eng = Engine(local)
    sh = sharding.ShardedEngine(eng)
    d = synth.dense(6, 8, 6, 1200, seed=11, nan_frac=0.01)
    labels = d["labels"]
    names = sorted({l["__name__"] for l in labels})
    gid = np.array([names.index(l["__name__"]) for l in labels], dtype=np.int32)
    owner = sharding.partition(labels, world)
    ts, val, off, mine = sharding.shard_csr(d["ts"], d["val"], d["off"], owner, rank)
    G = len(names)
    perm, goff = group_index(gid[mine], G)
    pg = (torch.as_tensor(perm, device=eng.device), torch.as_tensor(goff, device=eng.device))
    qs = (0.5, 0.95, 0.99)
    out, summ, q = sh.aggregate_step(ts, val, off, d["start_ms"], d["step_ms"], 1200, 60,
                                     torch.as_tensor(gid[mine], device=eng.device), G, pg, qs)
    torch.cuda.synchronize()
    ref = O.aggregate_range(d["ts"], d["val"], d["off"], d["start_ms"], d["step_ms"], 1200, 900_000)
    # local shard rows against the oracle's rows
    assert np.array_equal(out["aligned"].cpu().numpy().view(np.uint64), ref["aligned"][mine].view(np.uint64))
    osm, ocnt, omn, omx, omean = O.group_reduce(ref["aligned"], gid, G)
    sm, cnt, mn, mx, mean = [v.cpu().numpy() for v in summ]
    assert np.array_equal(cnt, ocnt) and np.array_equal(mn, omn) and np.array_equal(mx, omx)
    assert np.allclose(sm, osm, rtol=1e-12) and np.allclose(mean, omean, rtol=1e-12)
    for g in range(G):
        o = O.quantiles(ref["aligned"][gid == g], qs)
        assert int(q["n"][g]) == o["n"]
        assert np.array_equal(q["value"][g].cpu().numpy().view(np.uint64), np.array(o["value"]).view(np.uint64))
        assert np.array_equal(q["x_lo"][g].cpu().numpy().view(np.uint64), np.array(o["x_lo"]).view(np.uint64))
    dist.barrier()
    if rank == 0:
        print(f"multi_gpu_check ok: world={world}, shard sizes via label hash, summary + exact quantiles match the oracle")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
