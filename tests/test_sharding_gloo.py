"""CPU, world_size 2 over gloo: the multi-GPU layer's host logic (partition, partial combine,
distributed radix-select protocol).  The per-rank compute is a numpy stand-in built from the
oracle (test double for the CUDA kernels); the combine code under test is the product's."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from k8s_rocm_metrics_dashboard_b200 import sharding, synth
from oracle import prom_oracle as O

BINS = 2048
SHIFT = [53, 42, 31, 20, 9, 0]
BITS = [11, 11, 11, 11, 11, 9]


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


class NumpySelect:
    """numpy restatement of pmx_quantile_hist / pick / finish on one rank's values."""

    def __init__(self, x, seg, n_seg, qs):
        self.seg, self.n_seg, self.qs = np.asarray(seg), n_seg, list(qs)
        self.keys = [O.f64_to_key(x[self.seg == g][~np.isnan(x[self.seg == g])]) for g in range(n_seg)]
        self.n_t = 2 * len(qs)
        self.prefix = np.zeros((n_seg, self.n_t), dtype=np.uint64)
        self.k = np.zeros((n_seg, self.n_t), dtype=np.int64)
        self.n = np.zeros(n_seg, dtype=np.int64)

    def hist(self, p):
        n_h = 1 if p == 0 else self.n_t
        h = np.zeros((self.n_seg * n_h, BINS), dtype=np.int64)
        for g in range(self.n_seg):
            keys = self.keys[g]
            d = ((keys >> np.uint64(SHIFT[p])) & np.uint64((1 << BITS[p]) - 1)).astype(np.int64)
            for t in range(n_h):
                m = np.ones(len(keys), dtype=bool) if p == 0 else (keys >> np.uint64(64 - 11 * p)) == self.prefix[g, t]
                h[g * n_h + t] = np.bincount(d[m], minlength=BINS)
        return torch.from_numpy(h)

    def pick(self, p, h):
        h = h.numpy()
        n_h = 1 if p == 0 else self.n_t
        for g in range(self.n_seg):
            for t in range(self.n_t):
                row = h[g * n_h + (0 if p == 0 else t)]
                if p == 0:
                    n = int(row.sum()); self.n[g] = n
                    if n == 0:
                        continue
                    pos = (n - 1) * self.qs[t // 2]
                    lo = int(np.floor(pos)); hi = min(lo + 1, n - 1)
                    k = hi if t & 1 else lo
                elif self.n[g] == 0:
                    continue
                else:
                    k = int(self.k[g, t])
                cum = np.cumsum(row)
                b = int(np.searchsorted(cum, k, side="right"))
                self.prefix[g, t] = (self.prefix[g, t] << np.uint64(BITS[p])) | np.uint64(b)
                self.k[g, t] = k - (cum[b - 1] if b else 0)

    def finish(self):
        out = []
        for g in range(self.n_seg):
            vals = []
            for i, q in enumerate(self.qs):
                if self.n[g] == 0:
                    vals.append((np.nan, np.nan, np.nan)); continue
                n = int(self.n[g]); pos = (n - 1) * q; lo = int(np.floor(pos)); t = pos - lo
                a = float(O.key_to_f64(self.prefix[g, 2 * i])); b = float(O.key_to_f64(self.prefix[g, 2 * i + 1]))
                vals.append((a, b, O.lerp(a, b, t)))
            out.append(vals)
        return out


def _worker(rank, world, port, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        d = synth.dense(3, 4, 6, 200, seed=4, nan_frac=0.05)
        labels = d["labels"]; S = len(labels)
        x = d["val"].reshape(S, 200)
        owner = sharding.partition(labels, world)
        assert owner.min() >= 0 and owner.max() < world
        ts, val, off, mine = sharding.shard_csr(d["ts"], d["val"], d["off"], owner, rank)
        assert np.array_equal(val.view(np.uint64), x[mine].ravel().view(np.uint64))
        names = sorted({l["__name__"] for l in labels})
        gid = np.array([names.index(l["__name__"]) for l in labels], dtype=np.int32)
        G = len(names)
        # local partials (stand-in for pmx_group_reduce on this rank's shard), then the product's combine
        sm, cnt, mn, mx, _ = O.group_reduce(x[mine], gid[mine], G)
        sm, cnt, mn, mx, mean = sharding.combine_group_partials(
            torch.from_numpy(sm.copy()), torch.from_numpy(cnt.copy()), torch.from_numpy(mn.copy()),
            torch.from_numpy(mx.copy()), None)
        osm, ocnt, omn, omx, omean = O.group_reduce(x, gid, G)
        assert np.array_equal(cnt.numpy(), ocnt)
        assert np.array_equal(mn.numpy(), omn) and np.array_equal(mx.numpy(), omx)
        assert np.allclose(sm.numpy(), osm, rtol=1e-12) and np.allclose(mean.numpy(), omean, rtol=1e-12)
        # distributed exact quantiles: local histograms, all-reduce, identical pick on every rank
        qs = [0.5, 0.95, 0.99]
        sel = NumpySelect(x[mine], gid[mine], G, qs)
        res = sharding.distributed_quantiles(sel.hist, sel.pick, sel.finish, G, len(qs))
        for g in range(G):
            o = O.quantiles(x[gid == g], qs)
            for i in range(len(qs)):
                a, b, v = res[g][i]
                assert np.float64(a).view(np.uint64) == np.float64(o["x_lo"][i]).view(np.uint64)
                assert np.float64(b).view(np.uint64) == np.float64(o["x_hi"][i]).view(np.uint64)
                assert np.float64(v).view(np.uint64) == np.float64(o["value"][i]).view(np.uint64)
        ret[rank] = int(len(mine))
    finally:
        dist.destroy_process_group()


def test_world_size_2_partition_combine_and_quantile_protocol():
    world = 2
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), ret), nprocs=world, join=True)
    assert sum(ret.values()) == 3 * 4 * 6 and all(v > 0 for v in ret.values())


def test_partition_is_stable_and_label_order_independent():
    labels = synth.series_labels(5, 8, 6)
    a = sharding.partition(labels, 8)
    b = sharding.partition([dict(reversed(list(l.items()))) for l in labels], 8)
    assert np.array_equal(a, b)
    counts = np.bincount(a, minlength=8)
    assert counts.min() > 0.5 * counts.mean()
    assert sharding.label_hash({"a": "1", "b": "2"}) == sharding.label_hash({"b": "2", "a": "1"})
