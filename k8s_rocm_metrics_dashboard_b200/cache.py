"""Per-GPU series cache: the dashboard's 5 s refresh (app.py:24, :486) ingests
only the samples that are new and recomputes only the grid points that are new.

The reference re-queries and re-processes everything on every refresh.  Here every series keeps
its recent samples in a device-resident slab (fixed capacity per series, oldest samples dropped),
keyed by its label set.  ``ingest`` appends the samples of a ``query_range`` body that covers only
the time since the last refresh; ``refresh`` runs the fused kernel over just the grid points that
have become due plus the ``window_steps - 1`` points before them (their windows reach back that
far), on a compact CSR gathered from the tails of the slabs.  Work per refresh is proportional to
series x (window + new points) instead of series x range.

The slab append is a handful of torch index operations; the tail gather, decode, alignment, windows
and rates are libpmx.so kernels.
"""
from __future__ import annotations

import json

import numpy as np

from .engine import DEFAULT_LOOKBACK_MS, Engine, tokenize


class SeriesCache:
    def __init__(self, eng: Engine, start_ms, step_ms, window_steps, lookback_ms=DEFAULT_LOOKBACK_MS,
                 capacity=None, max_series=1024):
        """``capacity``: samples kept per series (default: enough for one window plus the lookback
        at one sample per step, doubled).  ``max_series``: initial number of slab rows (grows)."""
        self.eng = eng
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
        """The same from decoded samples: CSR over ``labels`` (device tensors, host offsets)."""
        eng, t, dev = self.eng, self.t, self.eng.device
        off = np.asarray(series_off, dtype=np.int64)
        lens = np.diff(off)
        rows = np.fromiter((self._row_of(l) for l in labels), dtype=np.int64, count=len(labels))
        if len(np.unique(rows)) != len(rows):
            raise ValueError("the same label set appears twice in one body")
        keep = lens > 0
        rows, lens, off0 = rows[keep], lens[keep], off[:-1][keep]
        if len(rows) == 0:
            return 0
        d_rows = t.as_tensor(rows, device=dev)
        d_len = t.as_tensor(lens, device=dev)
        d_off = t.as_tensor(off0, device=dev)
        cnt_old = self.cnt[d_rows]
        # skip what the series already holds: new samples are ascending, so it is a prefix
        last = t.where(cnt_old > 0, self.ts[d_rows, (cnt_old - 1).clamp(min=0)],
                       t.full_like(cnt_old, -(1 << 62)))
        sample_series = t.repeat_interleave(t.arange(len(rows), device=dev), d_len)
        flat = t.repeat_interleave(d_off, d_len) + (t.arange(int(lens.sum()), device=dev)
                                                    - t.repeat_interleave(t.cumsum(d_len, 0) - d_len, d_len))
        stale = (ts_new[flat] <= last[sample_series]).to(t.int64)
        skip = t.zeros_like(d_len).index_add_(0, sample_series, stale)
        d_off, d_len = d_off + skip, d_len - skip
        total = cnt_old + d_len
        new_total = total.clamp(max=self.cap)
        drop = total - new_total                                   # oldest samples that fall out of the slab
        if bool((drop > 0).any()):
            # a sample still needed by a grid point that has not been produced must not fall out
            needed_from = self.start_ms + (self.next_step - (self.w - 1)) * self.step_ms - self.lookback_ms
            newest_lost = self.ts[d_rows, (drop - 1).clamp(min=0, max=self.cap - 1)]    # meaningful while drop <= cnt_old
            lost = (drop > 0) & ((drop > cnt_old) | (newest_lost > needed_from))
            if bool(lost.any()):
                raise OverflowError("SeriesCache: capacity too small for the samples between two refreshes; "
                                    "call refresh() more often or raise capacity")
        pos = t.arange(self.cap, device=dev)[None, :]
        q = pos + drop[:, None]                                    # index into [old | new] of what lands at pos
        from_old = q < cnt_old[:, None]
        valid = pos < new_total[:, None]
        old_ts = self.ts[d_rows[:, None], q.clamp(max=self.cap - 1)]
        old_val = self.val[d_rows[:, None], q.clamp(max=self.cap - 1)]
        nidx = (d_off[:, None] + (q - cnt_old[:, None])).clamp(min=0, max=ts_new.numel() - 1)
        new_ts = t.where(from_old, old_ts, ts_new[nidx])
        new_val = t.where(from_old, old_val, val_new[nidx])
        self.ts[d_rows] = t.where(valid, new_ts, t.zeros_like(new_ts))
        self.val[d_rows] = t.where(valid, new_val, t.full_like(new_val, float("nan")))
        self.cnt[d_rows] = new_total
        return int(d_len.sum().item())

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
        # the slab rows from there on, gathered into one CSR (two kernels and a cumulative sum; the
        # compact buffers are sized for full rows, so no count has to come back to the host)
        from . import _abi
        p = eng._ptr
        first, lens = eng._empty((S,), t.int64), eng._empty((S,), t.int64)
        _abi.check(eng.lib.pmx_slab_tail_lengths(eng.ctx, p(self.ts), p(self.cnt), S, self.cap, g0 - self.lookback_ms,
                                                 p(first), p(lens), eng._stream()))
        off = t.zeros((S + 1,), dtype=t.int64, device=dev)
        t.cumsum(lens, 0, out=off[1:])
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
