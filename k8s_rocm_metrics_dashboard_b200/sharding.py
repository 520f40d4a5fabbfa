"""Multi-GPU layer: one process per GPU, series sharded by label set.

Series are independent for decode / align / rolling / rate, so ranks never exchange samples.
The only exchange steps are
  * per-group partials (sum, count, min, max) -- one all-gather of 4 x ``n_groups`` numbers, folded
    locally from identical data on every rank, for groups that span ranks (per-metric groups);
  * exact quantiles -- for the first passes the 2048-bin histograms of every (segment, target) are
    summed with one all-reduce per pass (16 KiB per histogram) and every rank picks the same bin
    on its own device (pmx_quantile_pick); after pass 2 the few keys that still match a target are
    all-gathered and every rank settles the segments from the union (Engine.quantiles).  No
    sample moves between ranks apart from those few keys.
``torch.distributed`` (NCCL over NVLink/NVSwitch on GPUs, gloo in the CPU tests) is the plumbing.
The combine logic is written against tensors and a process group so that the world_size-2
gloo tests exercise exactly this code with CPU tensors.
"""
from __future__ import annotations

import json
import zlib

import numpy as np


def label_hash(labels: dict) -> int:
    """Stable hash of a sorted label set (independent of process, rank and PYTHONHASHSEED)."""
    canon = json.dumps(sorted(labels.items()), separators=(",", ":"), ensure_ascii=True)
    return zlib.crc32(canon.encode("ascii"))


def partition(label_sets, world_size: int):
    """rank of every series: hash(sorted label set) mod world_size."""
    return np.fromiter((label_hash(l) % world_size for l in label_sets), dtype=np.int32, count=len(label_sets))


def shard_csr(ts, val, off, owner, rank):
    """The CSR arrays of the series owned by ``rank`` (host arrays), plus their global indices."""
    mine = np.flatnonzero(owner == rank)
    lens = (off[1:] - off[:-1])[mine]
    new_off = np.zeros(len(mine) + 1, dtype=np.int64)
    np.cumsum(lens, out=new_off[1:])
    if len(mine):
        idx = np.concatenate([np.arange(off[s], off[s + 1]) for s in mine]) if new_off[-1] else np.zeros(0, dtype=np.int64)
    else:
        idx = np.zeros(0, dtype=np.int64)
    return ts[idx], val[idx], new_off, mine


def combine_group_partials(sm, cnt, mn, mx, group=None, engine=None):
    """Combine per-group partials of all ranks and return (sum, count, min, max, mean).

    Inputs are torch tensors of shape [G] (or [G,T]) as produced by ``Engine.group_reduce`` on the
    local shard: empty groups carry sum 0, count 0, min/max NaN.  With ``engine`` and device
    tensors the exchange is two kernels of the library around one all-gather
    (``Engine.group_combine``); the tensor form below is what the CPU (gloo) tests run."""
    import torch
    import torch.distributed as dist
    if engine is not None and sm.is_cuda and dist.is_initialized() and dist.get_world_size(group) > 1:
        return engine.group_combine(sm.contiguous(), cnt.contiguous(), mn.contiguous(), mx.contiguous(),
                                    group or _world_group())
    inf = float("inf")
    mn = torch.where(cnt > 0, mn, torch.full_like(mn, inf))
    mx = torch.where(cnt > 0, mx, torch.full_like(mx, -inf))
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        # one collective: the exchange is latency-bound (a few KB), so everything travels in one
        # all-gather and is folded locally, from identical data on every rank (sums are then
        # bit-identical across ranks whatever the collective's internal order)
        world = dist.get_world_size(group)
        packed = torch.stack([sm, cnt.to(sm.dtype), mn, mx]).contiguous()   # counts exact in float64 below 2^53
        parts = [torch.empty_like(packed) for _ in range(world)]
        dist.all_gather(parts, packed, group=group)         # (all_gather_into_tensor is not in gloo)
        gathered = torch.stack(parts)
        both = gathered[:, 0:2].sum(dim=0)                  # same data and layout on every rank: same result
        sm, cnt = both[0], both[1].round().to(cnt.dtype)
        mn = gathered[:, 2].min(dim=0).values
        mx = gathered[:, 3].max(dim=0).values
    nan = torch.full_like(mn, float("nan"))
    any_ = cnt > 0
    mean = torch.where(any_, sm / cnt.to(sm.dtype).clamp(min=1), nan)
    return sm, cnt, torch.where(any_, mn, nan), torch.where(any_, mx, nan), mean


def distributed_quantiles(hist_fn, pick_fn, finish_fn, n_seg, n_q, group=None, passes=6):
    """The radix-select protocol, independent of where hist/pick run.

    hist_fn(pass) -> int64 tensor [n_hist, 2048] of LOCAL counts (n_hist = n_seg for pass 0,
    n_seg*2*n_q afterwards); pick_fn(pass, reduced_hist) advances the per-target state;
    finish_fn() returns the result.  This is the per-pass form, which Engine.quantiles runs for
    the first passes (and for all of them when a segment's remaining keys do not fit the gathered
    rows); the gloo tests plug in a numpy restatement."""
    import torch.distributed as dist
    for p in range(passes):
        h = hist_fn(p)
        if dist.is_initialized() and dist.get_world_size(group) > 1:
            dist.all_reduce(h, op=dist.ReduceOp.SUM, group=group)
        pick_fn(p, h)
    return finish_fn()


class PeerExchange:
    """All-gathers of a few hundred KiB as stores into peer memory (``pmx_peer_put``) followed by a
    device-side barrier, instead of NCCL calls of 20-40 us fixed latency each.  Every rank's receive
    buffer is symmetric memory (``torch.distributed._symmetric_memory``): CUDA VMM handles are
    exchanged once at construction, after that a rank "sends" by storing its block into slot ``rank``
    of every rank's buffer over NVLink.  Every named exchange has a region of its own; one step uses
    each region once, and a rank that starts the next step has passed this step's last barrier, so a
    region is never written while another rank still reads it.  Construction raises when symmetric
    memory cannot be set up (callers then stay with NCCL)."""

    def __init__(self, engine, group, regions):
        """``regions``: {name: bytes per rank}."""
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm
        t = engine.torch
        self.eng, self.group = engine, group
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        self.off, total = {}, 0
        for name, nbytes in regions.items():
            nbytes = (int(nbytes) + 15) & ~15
            self.off[name] = (total, nbytes)
            total += self.world * nbytes
        self.buf = symm.empty(total, dtype=t.uint8, device=engine.device)
        self.hdl = symm.rendezvous(self.buf, group)
        self.peers = t.tensor([int(p) for p in self.hdl.buffer_ptrs], dtype=t.int64, device=engine.device)
        self.channel = {name: i for i, name in enumerate(regions)}

    def all_gather(self, name, src):
        """``src``: contiguous device tensor of this region's size (or smaller, a multiple of 16 bytes).  Returns
        the gathered blocks, rank-major, as a tensor of ``src``'s dtype viewing this rank's buffer."""
        t = self.eng.torch
        off, cap = self.off[name]
        nbytes = src.numel() * src.element_size()
        if nbytes > cap or nbytes % 16:
            raise ValueError(f"PeerExchange.all_gather({name}): {nbytes} bytes do not fit the region of {cap} (multiples of 16)")
        _abi_check(self.eng.lib.pmx_peer_put(self.eng._ptr(src), self.eng._ptr(self.peers), self.world, self.rank, off, nbytes,
                                             self.eng._stream()))
        self.hdl.barrier(channel=self.channel[name])
        self.eng.launches += 1
        return self.buf[off:off + self.world * nbytes].view(src.dtype)


def _abi_check(rc):
    from . import _abi
    _abi.check(rc)


class ShardedEngine:
    """Engine + process group: the aggregate step on this rank's shard with cluster-wide summaries."""

    def __init__(self, engine, group=None, local_only=False, peer_exchange=True):
        """``local_only``: behave as one rank whatever the process group says (no exchange).
        ``peer_exchange``: several ranks exchange through peer memory (``PeerExchange``) where symmetric
        memory is available; False keeps the NCCL calls."""
        import torch.distributed as dist
        self.eng = engine
        self.group = group
        alone = local_only or not dist.is_initialized()
        self.world = 1 if alone else dist.get_world_size(group)
        self.rank = 0 if alone else dist.get_rank(group)
        self._qbuf = {}
        self._graphs = {}
        self._peers = {}
        self.peer_exchange = peer_exchange

    # The one-sweep selection spends a cluster of CTAs per segment on the brackets and another per
    # (segment, level) on the last keys, ~40 us each whatever the segment holds: it pays for long
    # segments (per-metric groups: millions of values), not for thousands of short ones (per-card groups).
    ONE_SWEEP_MIN_VALUES_PER_SEGMENT = 1 << 18

    def one_sweep_pays(self, n_series, n_steps, n_seg):
        """Several ranks must all take the same path, and their shards differ in size: there the answer
        depends on the number of segments alone."""
        if self.world > 1:
            return 0 < n_seg <= 64
        return n_seg > 0 and n_series * n_steps >= n_seg * self.ONE_SWEEP_MIN_VALUES_PER_SEGMENT

    def _peer(self, n_seg, n_q):
        """The peer-memory exchange for this shape, set up at first use (a collective: every rank calls it at the
        same point); None -- on every rank alike -- where symmetric memory is not available: NCCL then."""
        key = (n_seg, n_q)
        if not self.peer_exchange:
            return None
        if key not in self._peers:
            import torch.distributed as dist
            t = self.eng.torch
            group = self.group or _world_group()
            peer = None
            try:
                peer = PeerExchange(self.eng, group, self.eng.bsel_region_bytes(n_seg, n_q))
            except Exception:  # noqa: BLE001  -- no symmetric memory here
                peer = None
            ok = t.tensor([1 if peer is not None else 0], dtype=t.int32, device=self.eng.device)
            dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=group)
            self._peers[key] = peer if int(ok.item()) == 1 else None
        return self._peers[key]

    @staticmethod
    def _qsummary(summary):
        """(count, min, max) of the combined group summary: where the quantile selection starts."""
        return summary[1], summary[2], summary[3]

    def summarize(self, x, seg, n_seg, perm_goff, qs=(0.5, 0.95, 0.99), defer_check=False, _unchecked=False,
                  brackets_ready=False, stage_graphs=False):
        """Per-segment sum / count / min / max / mean and exact quantiles of the [S, T] device matrix ``x`` over
        ALL ranks (every rank calls this with its own rows; S may be 0).  Returns (summary, quantiles).
        One rank and at most four levels: one sweep of ``x`` (``Engine.summary_quantiles``); its brackets
        are verified on the device, and the per-pass selection runs if the check fails (with
        ``defer_check`` the check is left to ``Engine.deferred_ok``, see ``Engine.quantiles``).
        ``brackets_ready``: ``Engine.summary_brackets_csr`` (one rank) or ``summary_begin_csr_ranks`` (several; then
        on EVERY rank) already ran for these series."""
        eng = self.eng
        if 1 <= len(qs) <= 4 and self.one_sweep_pays(x.shape[0], x.shape[1], n_seg):
            if self.world == 1:
                summary, q, failed = eng.summary_quantiles(x, seg, n_seg, qs, perm_goff, brackets_ready=brackets_ready)
            else:                                       # ``failed`` is the same on every rank: they fall back together
                summary, q, failed = eng.summary_quantiles_ranks(x, seg, n_seg, qs, perm_goff, self.group or _world_group(),
                                                                 begun=brackets_ready, peer=self._peer(n_seg, len(qs)),
                                                                 stage_graphs=stage_graphs)
            if _unchecked:
                return summary, q, failed
            if defer_check:
                eng.defer_flag(failed)
                return summary, q
            if int(failed.item()) == 0:
                return summary, q
        sm, cnt, mn, mx, mean = eng.group_reduce(x, None, n_seg, perm_goff=perm_goff)
        key = (n_seg, len(qs))
        if key not in self._qbuf:
            self._qbuf[key] = eng.quantile_buffers(n_seg, len(qs))
        if self.world > 1:
            summary = combine_group_partials(sm, cnt, mn, mx, self.group, engine=eng)
            q = eng.quantiles(x, seg, n_seg, list(qs), group=self.group or _world_group(), buffers=self._qbuf[key],
                              row_order=perm_goff[0], summary=self._qsummary(summary), defer_check=defer_check)
        else:
            summary = (sm, cnt, mn, mx, mean)          # one rank: the kernel's own result is the summary
            q = eng.quantiles(x, seg, n_seg, list(qs), buffers=self._qbuf[key], row_order=perm_goff[0],
                              summary=self._qsummary(summary))
        return (summary, q, None) if _unchecked else (summary, q)

    def summarize_graph(self, x, seg, n_seg, perm_goff, qs=(0.5, 0.95, 0.99), defer_check=False, brackets_ready=False):
        """``summarize`` replayed from a CUDA graph (one rank): the small launches behind the fused kernel
        are launch-bound, so they are captured once per (buffer, shape) and replayed.  ``x``, ``seg`` and
        ``perm_goff`` must be the same device tensors on every call (the graph reads them in
        place); the returned tensors are the graph's own and are overwritten by the next replay.
        Falls back to the eager path when capture is not possible, and to the eager per-pass selection
        when the replayed sweep reports a failed bracket."""
        t = self.eng.torch
        key = (x.data_ptr(), tuple(x.shape), seg.data_ptr(), n_seg, perm_goff[0].data_ptr(), perm_goff[1].data_ptr(), tuple(qs),
               brackets_ready)
        hit = self._graphs.get(key)
        if hit is None:
            for _ in range(2):                                  # scratch grows and kernels load outside the capture
                before = self.eng.launches
                self.summarize(x, seg, n_seg, perm_goff, qs, _unchecked=True, brackets_ready=brackets_ready)
            n_launches = self.eng.launches - before             # kernels of ours inside one replay
            t.cuda.synchronize()
            try:
                g = t.cuda.CUDAGraph()
                with t.cuda.graph(g, capture_error_mode="thread_local"):
                    res = self.summarize(x, seg, n_seg, perm_goff, qs, _unchecked=True, brackets_ready=brackets_ready)
                self.eng.launches -= n_launches                 # the capture launched nothing
                hit = (g, res, (x, seg, perm_goff), n_launches)  # the inputs stay alive with the graph
            except Exception:  # noqa: BLE001  -- capture refused: stay eager
                t.cuda.synchronize()
                hit = (None, None, None, 0)
            self._graphs[key] = hit
        if hit[0] is None:
            return self.summarize(x, seg, n_seg, perm_goff, qs, defer_check=defer_check, brackets_ready=brackets_ready)
        hit[0].replay()
        self.eng.launches += hit[3]
        summary, q, failed = hit[1]
        if failed is not None:
            if defer_check:
                self.eng.defer_flag(failed)
            elif int(failed.item()) != 0:
                return self.summarize(x, seg, n_seg, perm_goff, qs)   # (finds the failure again and runs the passes)
        return summary, q

    def aggregate_step(self, ts, val, off, start_ms, step_ms, n_steps, window_steps, seg, n_seg, perm_goff,
                       qs=(0.5, 0.95, 0.99), out=None, lookback_ms=300_000, host=None, defer_check=False,
                       graph=False):
        """fused align+rolling+rate on the shard, then per-metric sum/count/min/max/mean and exact
        quantiles of the aligned values over ALL ranks.  ``seg``: metric id per local series.
        With ``host`` = {"dev": device buffers, "out": pinned host outputs} the inputs are pinned host
        tensors and the copies are pipelined with the kernel (Engine.fused_from_host).
        ``defer_check``: see ``Engine.quantiles`` -- no host synchronisation inside the step; the
        caller asks ``Engine.deferred_ok()`` later.
        ``graph`` (one rank, with ``out`` given): the launches behind the fused kernel are replayed
        from a CUDA graph (``summarize_graph``)."""
        eng = self.eng
        if host is not None:        # ts/val/off are pinned host tensors; results also go to host["out"]
            out = eng.fused_from_host(ts, val, off, start_ms, step_ms, n_steps, window_steps, host["dev"],
                                      host["out"], lookback_ms)
        else:
            # the brackets of the quantile selection come from the raw samples on a side stream while the
            # fused kernel runs (they are hints; the sweep verifies them)
            pre = (1 <= len(qs) <= 4 and getattr(val, "is_cuda", False) and getattr(off, "is_cuda", False) and n_steps > 0
                   and (self.world > 1 or off.numel() > 1) and self.one_sweep_pays(off.numel() - 1, n_steps, n_seg))
            if pre:
                t = eng.torch
                cur = t.cuda.current_stream()
                if self.__dict__.get("_side") is None:
                    self._side = t.cuda.Stream()
                self._side.wait_stream(cur)
                with t.cuda.stream(self._side):
                    if self.world == 1:
                        eng.summary_brackets_csr(val, off, n_steps, n_seg, qs, perm_goff)
                    else:
                        eng.summary_begin_csr_ranks(val, off, n_steps, n_seg, qs, perm_goff, self.group or _world_group())
                ready = self._side.record_event()
            out = eng.fused(ts, val, off, start_ms, step_ms, n_steps, window_steps, lookback_ms, out=out)
            if pre:
                cur.wait_event(ready)
        # (one rank only: with NCCL collectives inside the capture the step gained 2 % on two GPUs and the
        # processes no longer shut down cleanly)
        pre = host is None and pre
        if graph and self.world == 1 and host is None:
            summary, q = self.summarize_graph(out["aligned"], seg, n_seg, perm_goff, qs, defer_check=defer_check, brackets_ready=pre)
        else:
            # several ranks with ``graph``: the launches between two exchanges are replayed from graphs
            summary, q = self.summarize(out["aligned"], seg, n_seg, perm_goff, qs, defer_check, brackets_ready=pre,
                                        stage_graphs=graph and self.world > 1 and host is None)
        return out, summary, q



def _world_group():
    import torch.distributed as dist
    return dist.group.WORLD
