"""Per-GPU series cache: the dashboard's 5 s refresh (app.py:24, :486) ingests
only the samples that are new and recomputes only the grid points that are new.

The reference re-queries and re-processes everything on every refresh.  Here every series keeps
its recent samples in a device-resident slab (fixed capacity per series, oldest samples dropped),
keyed by its label set.  ``ingest`` appends the samples of a ``query_range`` body that covers only
the time since the last refresh; ``refresh`` runs the fused kernel over just the grid points that
have become due plus the ``window_steps - 1`` points before them (their windows reach back that
far), on a compact CSR gathered from the tails of the slabs.  Work per refresh is proportional to
series x (window + new points) instead of series x range.

The slab append, the tail gather, decode, alignment, windows and rates are libpmx.so kernels.
"""
from __future__ import annotations

import json

import numpy as np

from .engine import DEFAULT_LOOKBACK_MS, Engine, tokenize


class SeriesCache:
    def __init__(self, eng: Engine, start_ms, step_ms, window_steps, lookback_ms=DEFAULT_LOOKBACK_MS,
                 capacity=None, max_series=1024, rank=0, world=1):
        """``capacity``: samples kept per series (default: enough for one window plus the lookback
        at one sample per step, doubled).  ``max_series``: initial number of slab rows (grows).
        ``rank`` / ``world``: this cache holds the series with ``label_hash % world == rank`` -- the
        partition of ``sharding.partition`` -- and ignores the others in what it is given."""
        self.eng = eng
        self.rank, self.world = int(rank), int(world)
        self.t = eng.torch
        self.start_ms, self.step_ms, self.w, self.lookback_ms = int(start_ms), int(step_ms), int(window_steps), int(lookback_ms)
        need = self.w + self.lookback_ms // self.step_ms + 2
        self.cap = int(capacity) if capacity else 2 * need
        if self.cap < need:
            raise ValueError(f"capacity {self.cap} cannot hold a window plus the lookback ({need} samples)")
        self.index = {}                 # canonical label string -> slab row
        self.labels = []
        self._rows = 0
        self._alloc(max_series)
        self.next_step = 0              # first grid point not yet produced

    # ------------------------------------------------------------------ storage
    def _alloc(self, rows):
        t, dev = self.t, self.eng.device
        ts = t.zeros((rows, self.cap), dtype=t.int64, device=dev)
        val = t.full((rows, self.cap), float("nan"), dtype=t.float64, device=dev)
        cnt = t.zeros((rows,), dtype=t.int64, device=dev)
        if self._rows:
            ts[:self._rows] = self.ts[:self._rows]; val[:self._rows] = self.val[:self._rows]; cnt[:self._rows] = self.cnt[:self._rows]
        self.ts, self.val, self.cnt = ts, val, cnt

    def _row_of(self, lab):
        k = json.dumps(lab, sort_keys=True, separators=(",", ":"))
        r = self.index.get(k)
        if r is None:
            r = self._rows
            if r == self.ts.shape[0]:
                self._alloc(2 * r)
            self.index[k] = r
            self.labels.append(lab)
            self._rows += 1
        return r

    @property
    def n_series(self):
        return self._rows

    # ------------------------------------------------------------------ ingest
    def ingest(self, body, tokenizer="host"):
        """Append the samples of a ``query_range`` matrix body (host walk or ``"device"`` tokenizer).
        Samples not newer than what a series already holds are skipped, so refresh queries may
        overlap.  Returns the number of samples stored."""
        eng, t = self.eng, self.t
        tok = eng.tokenize_device(body) if tokenizer == "device" else tokenize(body)
        if tok.n_samples == 0:
            return 0
        ts_new, val_new = eng.parse(tok)
        return self.append(tok.labels(), ts_new, val_new, tok.series_off)

    def append(self, labels, ts_new, val_new, series_off):
        """The same from decoded samples: CSR over ``labels`` (device tensors, host offsets).  One call
        of the library (pmx_slab_append: a plan kernel and an in-place move) and one read of its two
        status words.  With ``world > 1`` only the series whose label hash belongs to this rank are
        stored (the others are another rank's shard)."""
        eng, t, dev = self.eng, self.t, self.eng.device
        from . import _abi
        off = np.asarray(series_off, dtype=np.int64)
        n_in = len(labels)
        if n_in == 0 or int(off[-1]) == 0:
            return 0
        if self.world > 1:
            from .sharding import label_hash
            rows = np.fromiter((self._row_of(l) if label_hash(l) % self.world == self.rank else -1 for l in labels),
                               dtype=np.int64, count=n_in)
        else:
            rows = np.fromiter((self._row_of(l) for l in labels), dtype=np.int64, count=n_in)
        mine = rows[rows >= 0]
        if len(np.unique(mine)) != len(mine):
            raise ValueError("the same label set appears twice in one body")
        # one upload: [offsets | rows]; scratch and status live with the cache
        host = np.concatenate([off, rows])
        d_in = t.as_tensor(host, device=dev)
        if getattr(self, "_work_n", 0) < n_in:
            self._work, self._work_n = eng._empty((2 * n_in,), t.int64), n_in
            self._status = eng._empty((2,), t.int64)
        # a sample still needed by a grid point that has not been produced must not fall out
        needed_from = self.start_ms + (self.next_step - (self.w - 1)) * self.step_ms - self.lookback_ms
        p = eng._ptr
        _abi.check(eng.lib.pmx_slab_append(eng.ctx, p(self.ts), p(self.val), p(self.cnt), self.cap, p(ts_new), p(val_new),
                                           p(d_in), p(d_in[n_in + 1:]), n_in, needed_from, p(self._work), p(self._status),
                                           eng._stream()))
        eng.launches += 2
        lost, stored = self._status.cpu().tolist()
        if lost:
            raise OverflowError("SeriesCache: capacity too small for the samples between two refreshes; "
                                "call refresh() more often or raise capacity")
        return int(stored)

    # ------------------------------------------------------------------ refresh
    def refresh(self, upto_ms):
        """Produce the grid points that have become due: every g = start + j*step with
        next_step <= j and g <= upto_ms.  Returns None when there is none, else a dict with
        ``first_step``, ``ts`` (int64[k] grid times) and [n_series, k] device tensors ``aligned``,
        ``roll_min``, ``roll_max``, ``roll_mean``, ``rate`` (rows in ``self.labels`` order)."""
        eng, t, dev = self.eng, self.t, self.eng.device
        j_end = (int(upto_ms) - self.start_ms) // self.step_ms
        k = j_end - self.next_step + 1
        if k <= 0 or self._rows == 0:
            return None
        a = max(0, self.next_step - (self.w - 1))                   # the windows of the new points reach back to here
        n = j_end - a + 1
        g0 = self.start_ms + a * self.step_ms
        S = self._rows
        # a sample can serve a grid point >= g0 only if it is newer than g0 - lookback: the tails of
        # the slab rows from there on, gathered into one CSR (two kernels; the
        # compact buffers are sized for full rows, so no count has to come back to the host)
        from . import _abi
        p = eng._ptr
        first, off = eng._empty((S,), t.int64), eng._empty((S + 1,), t.int64)
        if getattr(self, "_tail_work", None) is None:
            self._tail_work = t.zeros((1,), dtype=t.int32, device=dev)
        _abi.check(eng.lib.pmx_slab_tail_offsets(eng.ctx, p(self.ts), p(self.cnt), S, self.cap, g0 - self.lookback_ms,
                                                 p(first), p(off), p(self._tail_work), eng._stream()))
        if getattr(self, "_compact_rows", 0) < S:
            self._ts_c, self._val_c = eng._empty((S * self.cap,), t.int64), eng._empty((S * self.cap,), t.float64)
            self._compact_rows = S
        ts_c, val_c = self._ts_c, self._val_c
        _abi.check(eng.lib.pmx_slab_tail_gather(eng.ctx, p(self.ts), p(self.val), p(first), p(off), S, self.cap,
                                                p(ts_c), p(val_c), eng._stream()))
        eng.launches += 2
        out = eng.fused(ts_c, val_c, off, g0, self.step_ms, n, self.w, self.lookback_ms)
        res = {key: v[:, n - k:] for key, v in out.items()}
        res["first_step"] = self.next_step
        res["ts"] = self.start_ms + self.step_ms * np.arange(self.next_step, j_end + 1, dtype=np.int64)
        self.next_step = j_end + 1
        return res

    def refresh_summary(self, upto_ms, sharded, names, qs=(0.5, 0.95, 0.99)):
        """``refresh`` plus the per-metric summary of the NEW grid points over all ranks: sum / count /
        min / max / mean and exact quantiles through ``sharded`` (a ``ShardedEngine``; its process
        group spans the ranks whose caches partition the series).  ``names``: the metric names in
        the order of the summary rows, the same list on every rank.  Every rank must call this
        together (ranks without a due point or without series take part with empty data)."""
        t, eng = self.t, self.eng
        from .engine import group_index
        res = self.refresh(upto_ms)
        where = {n: i for i, n in enumerate(names)}
        if res is None:
            k = max(0, (int(upto_ms) - self.start_ms) // self.step_ms - self.next_step + 1)
            x = eng._empty((0, max(k, 1)), t.float64)
            gid = np.zeros(0, dtype=np.int32)
        else:
            x = res["aligned"].contiguous()
            gid = np.fromiter((where[l.get("__name__", "")] for l in self.labels), dtype=np.int32, count=len(self.labels))
        pg = group_index(gid, len(names))
        summary, q = sharded.summarize(x, gid, len(names), pg, qs)
        return res, summary, q

