"""Multi-rank parity check.  ``parity_check`` is called by bench.py at N > 1 (its result goes into the
JSON line) and by tests/test_gpu_multi_rank.py; the file also runs under torchrun on >= 2 GPUs:

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


def parity_check(eng, sh, rank, world, n_nodes=6, n_steps=1200):
    """Every rank takes its label-hash shard of a small seeded config, runs the aggregate step
    (fused kernel locally, per-metric summary and exact quantiles over all ranks) and compares
    with the oracle on the UNION of all shards.  Returns a dict for bench.py's JSON line; raises
    AssertionError on a mismatch.  Works at world size 1 too (no process group needed)."""
    d = synth.dense(n_nodes, 8, 6, n_steps, seed=11, nan_frac=0.01)
    labels = d["labels"]
    names = sorted({l["__name__"] for l in labels})
    gid = np.array([names.index(l["__name__"]) for l in labels], dtype=np.int32)
    owner = sharding.partition(labels, world)
    ts, val, off, mine = sharding.shard_csr(d["ts"], d["val"], d["off"], owner, rank)
    G = len(names)
    perm, goff = group_index(gid[mine], G)
    pg = (torch.as_tensor(perm, device=eng.device), torch.as_tensor(goff, device=eng.device))
    qs = (0.5, 0.95, 0.99)
    seg = torch.as_tensor(gid[mine], device=eng.device)
    ref = {"aligned": O.align(d["ts"], d["val"], d["off"], d["start_ms"], d["step_ms"], n_steps)[0]}
    osm, ocnt, omn, omx, omean = O.group_reduce(ref["aligned"], gid, G)
    worst = 0.0
    for defer in (False, True):                                  # the synchronous and the deferred form of the step
        out, summ, q = sh.aggregate_step(ts, val, off, d["start_ms"], d["step_ms"], n_steps, 60, seg, G, pg, qs,
                                         defer_check=defer)
        torch.cuda.synchronize()
        if defer:
            assert eng.deferred_ok(wait=True), "a segment did not fit the key exchange"
        # local shard rows against the oracle's rows
        assert np.array_equal(out["aligned"].cpu().numpy().view(np.uint64), ref["aligned"][mine].view(np.uint64))
        sm, cnt, mn, mx, mean = [v.cpu().numpy() for v in summ]
        assert np.array_equal(cnt, ocnt) and np.array_equal(mn, omn) and np.array_equal(mx, omx)
        err = float(np.max(np.abs(sm - osm) / np.maximum(np.abs(osm), 1e-300)))
        worst = max(worst, err, float(np.max(np.abs(mean - omean) / np.maximum(np.abs(omean), 1e-300))))
        assert worst <= 1e-12, f"sum/mean relative error {worst}"
        for g in range(G):
            o = O.quantiles(ref["aligned"][gid == g], qs)
            assert int(q["n"][g]) == o["n"]
            for k in ("value", "x_lo", "x_hi"):
                assert np.array_equal(q[k][g].cpu().numpy().view(np.uint64), np.array(o[k]).view(np.uint64)), k
    return {"world": world, "ok": True, "series_union": len(labels), "series_local": int(len(mine)), "steps": n_steps,
            "checked": "aligned rows bit-exact; per-metric count/min/max exact, sum/mean rel<=1e-12; p50/p95/p99 "
                       "order statistics and values bit-exact vs the oracle on the union of all shards",
            "worst_sum_mean_rel_err": worst}


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    eng = Engine(local)
    sh = sharding.ShardedEngine(eng)
    res = parity_check(eng, sh, rank, world)
    dist.barrier()
    if rank == 0:
        print(f"multi_gpu_check ok: {res}")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
