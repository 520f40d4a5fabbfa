"""Engine: one pmx_ctx on one GPU; torch is used for device memory and streams only.

Every method hands raw device pointers to libpmx.so through the C ABI (include/pmx.h).  There
is no CPU path: constructing an Engine without a CUDA device raises.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

import numpy as np

from . import _abi

DEFAULT_LOOKBACK_MS = 300_000


class LabelTable:
    """The label sets of all series, interned (pmx_labels_intern): ``strings`` holds every distinct
    label name / value once, series s owns the pairs ``pair_off[s]:pair_off[s+1]`` of
    (``key_id``, ``val_id``).  Behaves like the list of dicts ``json.loads`` would give (len, index,
    iteration build the dict of one series on demand); ``column(name)`` is the vectorised form."""

    def __init__(self, strings, pair_off, key_id, val_id):
        self.strings, self.pair_off, self.key_id, self.val_id = strings, pair_off, key_id, val_id
        self._id = {s: i for i, s in enumerate(strings)}

    def __len__(self):
        return len(self.pair_off) - 1

    def __getitem__(self, s):
        if isinstance(s, slice):
            return [self[i] for i in range(*s.indices(len(self)))]
        if s < 0:
            s += len(self)
        a, b = int(self.pair_off[s]), int(self.pair_off[s + 1])
        st = self.strings
        return {st[k]: st[v] for k, v in zip(self.key_id[a:b].tolist(), self.val_id[a:b].tolist())}

    def __iter__(self):
        return (self[i] for i in range(len(self)))

    def __eq__(self, other):
        try:
            return len(self) == len(other) and all(a == b for a, b in zip(self, other))
        except TypeError:
            return NotImplemented

    __hash__ = None

    def column(self, name):
        """int32[n_series]: the string id of label ``name`` in every series, -1 where it is missing."""
        out = np.full(len(self), -1, dtype=np.int32)
        k = self._id.get(name)
        if k is None:
            return out
        hit = np.flatnonzero(self.key_id == k)
        series = np.searchsorted(self.pair_off, hit, side="right") - 1
        out[series] = self.val_id[hit]
        return out

    def values(self, name, missing=""):
        """The label ``name`` of every series as a list of str (``missing`` where absent)."""
        ids = self.column(name)
        st = self.strings
        return [st[i] if i >= 0 else missing for i in ids.tolist()]


def _label_table(body, spans):
    """pmx_labels_intern on the metric spans of a body (bytes or uint8 array).  ``spans``: a list of
    (offset, length) or the two uint32 arrays themselves (what the tokenizers produce: no per-series
    Python object on the way)."""
    lib = _abi.lib()
    arrays = isinstance(spans, tuple) and len(spans) == 2 and isinstance(spans[0], np.ndarray)
    n = len(spans[0]) if arrays else len(spans)
    if isinstance(body, (bytes, bytearray)):
        raw, raw_len = body, len(body)
    else:                                                    # uint8 array (e.g. the view of a pinned tensor): no copy
        arr = np.ascontiguousarray(body, dtype=np.uint8)
        raw, raw_len = C.c_void_p(arr.ctypes.data), arr.size
    if arrays:
        m_off, m_len = (np.ascontiguousarray(a, dtype=np.uint32) for a in spans)
    else:
        m_off = np.fromiter((o for o, _ in spans), dtype=np.uint32, count=n)
        m_len = np.fromiter((l for _, l in spans), dtype=np.uint32, count=n)
    total = int(m_len.sum())
    cap_pairs = total // 6 + 1                               # "k":"v", is at least 8 bytes; be generous
    pair_off = np.empty(n + 1, dtype=np.int64)
    key_id, val_id = np.empty(cap_pairs, dtype=np.int32), np.empty(cap_pairs, dtype=np.int32)
    strings = np.empty(max(total, 1), dtype=np.uint8)
    cap_distinct = cap_pairs * 2 + 1
    str_off = np.empty(cap_distinct + 1, dtype=np.int64)
    res = _abi.LabelsResult()
    rc = lib.pmx_labels_intern(raw, raw_len, m_off.ctypes.data, m_len.ctypes.data, n, pair_off.ctypes.data,
                               key_id.ctypes.data, val_id.ctypes.data, cap_pairs, strings.ctypes.data, strings.size,
                               str_off.ctypes.data, cap_distinct, C.byref(res))
    if rc != _abi.OK:
        raise ValueError(lib.pmx_last_error().decode("utf-8", "replace"))
    blob = strings[:res.strings_len].tobytes()
    so = str_off[:res.n_strings + 1].tolist()
    st = [blob[so[i]:so[i + 1]].decode("utf-8", "surrogatepass") for i in range(res.n_strings)]
    return LabelTable(st, pair_off, key_id[:res.n_pairs].copy(), val_id[:res.n_pairs].copy())


@dataclass
class Tokens:
    """Host-side result of pmx_tokenize: the flat byte buffer and per-sample offsets."""
    flat: np.ndarray            # uint8: body + decoded escaped tokens
    ts_off: np.ndarray          # uint32[N]
    val_off: np.ndarray         # uint32[N]
    series_off: np.ndarray      # int64[S+1]
    metric_spans: list          # [(off, len)] of each series' "metric" object in the body
    result_type: str            # "matrix" | "vector"
    span_arrays: tuple = None   # the same spans as (uint32[S] offsets, uint32[S] lengths)

    @property
    def n_samples(self):
        return len(self.ts_off)

    @property
    def n_series(self):
        return len(self.series_off) - 1

    def labels(self):
        """The label sets, interned (``LabelTable``: indexes and iterates like a list of dicts)."""
        return _label_table(self.flat, self.span_arrays if self.span_arrays is not None else self.metric_spans)


@dataclass
class DeviceTokens:
    """Result of pmx_tokenize_device: the body and the per-sample offsets stay in device memory;
    only the CSR offsets and the label spans (a few bytes per series) come back to the host."""
    body: bytes                 # the host copy the labels are read from
    d_text: object              # uint8 device tensor holding the body
    d_ts_off: object            # int32 device tensor (uint32 offsets), N entries
    d_val_off: object
    series_off: np.ndarray      # int64[S+1] (host)
    metric_spans: list
    result_type: str
    span_arrays: tuple = None   # the same spans as (uint32[S] offsets, uint32[S] lengths)

    @property
    def n_samples(self):
        return int(self.series_off[-1]) if len(self.series_off) else 0

    @property
    def n_series(self):
        return len(self.series_off) - 1

    def labels(self):
        """The label sets, interned (``LabelTable``: indexes and iterates like a list of dicts)."""
        return _label_table(self.body, self.span_arrays if self.span_arrays is not None else self.metric_spans)


_TOKD_ERRORS = {1: "not a successful Prometheus matrix/vector answer in compact encoding",
                2: "'[' count does not match samples + series (white space between tokens, or a shape other than [t,\"v\"])",
                3: "capacity exceeded", 4: "malformed sample near byte {where}",
                5: "unterminated label object near byte {where}"}


def tokenize(body: bytes, threads=None) -> Tokens:
    """X1 on the host (C++, no Python objects per sample).  Raises ValueError on a malformed body.
    Bodies of 1 MiB and more are cut at series boundaries and walked by ``threads`` host threads
    (default: the host's cores, at most 32); ``threads=1`` forces the single walk."""
    if isinstance(body, str):
        body = body.encode("utf-8")
    lib = _abi.lib()
    n = len(body)
    if threads is None:
        import os
        threads = min(32, os.cpu_count() or 1)
    cap_s, cap_r = n // 8 + 1, n // 24 + 1
    ts_off = np.empty(cap_s, dtype=np.uint32)
    val_off = np.empty(cap_s, dtype=np.uint32)
    series_off = np.empty(cap_r + 1, dtype=np.int64)
    m_off = np.empty(cap_r, dtype=np.uint32)
    m_len = np.empty(cap_r, dtype=np.uint32)
    extra = np.empty(n + 16, dtype=np.uint8)
    res = _abi.TokenizeResult()
    if threads > 1 and n >= (1 << 20):
        rc = lib.pmx_tokenize_mt(body, n, ts_off.ctypes.data, val_off.ctypes.data, cap_s,
                                 series_off.ctypes.data, m_off.ctypes.data, m_len.ctypes.data, cap_r,
                                 extra.ctypes.data, extra.size, threads, C.byref(res))
    else:
        rc = lib.pmx_tokenize(body, n, ts_off.ctypes.data, val_off.ctypes.data, cap_s,
                              series_off.ctypes.data, m_off.ctypes.data, m_len.ctypes.data, cap_r,
                              extra.ctypes.data, extra.size, C.byref(res))
    if rc == _abi.E_KEY:
        raise KeyError("result")
    if rc != _abi.OK:
        raise ValueError(lib.pmx_last_error().decode("utf-8", "replace"))
    ns, nr, ne = res.n_samples, res.n_series, res.extra_len
    flat = np.frombuffer(body, dtype=np.uint8)
    if ne:
        flat = np.concatenate([flat, extra[:ne]])
    mo, ml = m_off[:nr].astype(np.uint32), m_len[:nr].astype(np.uint32)
    return Tokens(flat, ts_off[:ns].copy(), val_off[:ns].copy(), series_off[:nr + 1].copy(),
                  list(zip(mo.tolist(), ml.tolist())), "matrix" if res.result_type == 0 else "vector", (mo, ml))


def group_index(group_id, n_groups):
    """Host helper (pmx_group_index): group id per series -> (perm int32[S'], goff int64[G+1])."""
    gid = np.ascontiguousarray(group_id, dtype=np.int32)
    perm = np.empty(max(1, len(gid)), dtype=np.int32)
    goff = np.empty(n_groups + 1, dtype=np.int64)
    _abi.check(_abi.lib().pmx_group_index(gid.ctypes.data, len(gid), n_groups, perm.ctypes.data, goff.ctypes.data))
    return perm[:int(goff[-1])], goff


class Engine:
    def __init__(self, device=None):
        import torch
        if not torch.cuda.is_available():
            raise RuntimeError("k8s_rocm_metrics_dashboard_b200.Engine needs a CUDA device (B200, sm_100a); "
                               "there is no CPU fallback")
        self.torch = torch
        self.device = torch.device("cuda", torch.cuda.current_device() if device is None else device)
        self.lib = _abi.lib()
        ctx = C.c_void_p()
        with torch.cuda.device(self.device):
            _abi.check(self.lib.pmx_ctx_create(self.device.index, C.byref(ctx)))
        self.ctx = ctx
        self.launches = 0           # kernels of ours launched (bench's gpu_launches claim)

    def close(self):
        if getattr(self, "ctx", None):
            self.lib.pmx_ctx_destroy(self.ctx)
            self.ctx = None

    def __del__(self):
        try:
            self.close()
        except Exception:  # noqa: BLE001
            pass

    # ------------------------------------------------------------------ helpers
    def _stream(self):
        return C.c_void_p(self.torch.cuda.current_stream(self.device).cuda_stream)

    def _empty(self, shape, dtype):
        return self.torch.empty(shape, dtype=dtype, device=self.device)

    def _dev(self, a, dtype):
        t = self.torch
        if isinstance(a, t.Tensor):
            return a.to(device=self.device, dtype=dtype).contiguous()
        return t.as_tensor(np.ascontiguousarray(a), dtype=dtype).to(self.device)

    @staticmethod
    def _ptr(t):
        return C.c_void_p(t.data_ptr()) if t is not None else None

    def scratch_bytes(self):
        return int(self.lib.pmx_ctx_scratch_bytes(self.ctx))

    # ------------------------------------------------------------------ K1
    def tokenize_device(self, body) -> DeviceTokens:
        """X1 on the GPU: one H2D copy of the body, then pmx_tokenize_device.  Raises ValueError when
        the body is not in Prometheus' compact encoding (use ``tokenize`` for general JSON).
        ``body``: bytes / str, or a uint8 CPU tensor -- a PINNED one (e.g. filled by
        ``response.raw.readinto``) is copied at the full PCIe rate instead of through a staging
        buffer (131 MB: 2.4 ms instead of ~11 ms)."""
        t = self.torch
        if isinstance(body, str):
            body = body.encode("utf-8")
        if isinstance(body, t.Tensor):
            if body.dtype != t.uint8 or body.is_cuda or body.dim() != 1:
                raise ValueError("tokenize_device: a tensor body must be a 1-D uint8 CPU tensor")
            n = body.numel()
            text = body.to(self.device, non_blocking=True)
            body = body.numpy()                            # host view for the labels (shares the memory)
        else:
            n = len(body)
            import warnings
            with warnings.catch_warnings():                # read-only bytes view; it is only read
                warnings.simplefilter("ignore", UserWarning)
                text = t.frombuffer(body, dtype=t.uint8).to(self.device, non_blocking=True)
        cap_s, cap_r = n // 8 + 1, n // 24 + 1
        ts_off, val_off = self._empty((cap_s,), t.int32), self._empty((cap_s,), t.int32)
        series_off = self._empty((cap_r + 1,), t.int64)
        m_off, m_len = self._empty((cap_r,), t.int32), self._empty((cap_r,), t.int32)
        status = self._empty((4,), t.int64)
        _abi.check(self.lib.pmx_tokenize_device(self.ctx, self._ptr(text), n, self._ptr(ts_off), self._ptr(val_off), cap_s,
                                                self._ptr(series_off), self._ptr(m_off), self._ptr(m_len), cap_r,
                                                self._ptr(status), self._stream()))
        self.launches += 3
        ns, nr, err, where = status.cpu().tolist()
        if err:
            raise ValueError("pmx_tokenize_device: " + _TOKD_ERRORS.get(err, f"error {err}").format(where=where))
        soff = series_off[:nr + 1].cpu().numpy()
        mo, ml = m_off[:nr].cpu().numpy().view(np.uint32), m_len[:nr].cpu().numpy().view(np.uint32)
        spans = list(zip(mo.tolist(), ml.tolist()))
        kind = "matrix" if bytes(body[42:48]) == b"matrix" else "vector"
        return DeviceTokens(body, text, ts_off[:ns], val_off[:ns], soff, spans, kind, (mo, ml))

    def parse(self, tok, want_ts=True):
        """Device decode of all value (and timestamp) tokens.  Returns (ts int64[N] | None, val float64[N]).
        Raises ValueError naming the first malformed token, like float() at app.py:188.
        ``tok``: Tokens (host offsets, uploaded here) or DeviceTokens (already resident)."""
        t = self.torch
        if isinstance(tok, DeviceTokens):
            return self._parse_device(tok, want_ts)
        import warnings
        with warnings.catch_warnings():                    # the body is a read-only bytes view; it is only read
            warnings.simplefilter("ignore", UserWarning)
            text = t.from_numpy(tok.flat).to(self.device)
        n = tok.n_samples
        voff = t.from_numpy(tok.val_off.view(np.int32)).to(self.device)
        toff = t.from_numpy(tok.ts_off.view(np.int32)).to(self.device) if want_ts else None
        ts, val, b = self._parse_launch(text, toff, voff, n, want_ts)
        if b[0] < n:
            o = int(tok.val_off[b[0]])
            end = o
            while end < len(tok.flat) and tok.flat[end] != 0x22:
                end += 1
            err = ValueError(f"could not convert string to float: {bytes(tok.flat[o:end]).decode('utf-8', 'replace')!r}")
            err.sample_index = int(b[0])
            raise err
        if want_ts and b[1] < n:
            raise ValueError(f"malformed timestamp in sample {b[1]}")
        return ts, val

    def _parse_launch(self, text, toff, voff, n, want_ts):
        """pmx_parse_samples: one launch for both tokens of every sample.  Returns (ts | None, val,
        [first bad value index, first bad timestamp index]) -- an index >= n means none."""
        t = self.torch
        val = self._empty((n,), t.float64)
        ts = self._empty((n,), t.int64) if want_ts else None
        bad = self._empty((2,), t.int64)
        _abi.check(self.lib.pmx_parse_samples(self.ctx, self._ptr(text), text.numel(), self._ptr(toff), self._ptr(voff), n,
                                              self._ptr(ts), self._ptr(val), self._ptr(bad), self._stream()))
        self.launches += 1
        return ts, val, bad.cpu().tolist()

    def parse_async(self, tok, want_ts=True):
        """``parse`` of DeviceTokens in two halves: the kernel is launched and (ts | None, val, finish)
        returned at once; ``finish()`` waits for the kernel's verdict and raises what ``parse`` would
        raise.  Host work that does not need the numbers (the label table) runs between the two;
        nothing may USE ts / val before ``finish()`` has returned."""
        t = self.torch
        if not isinstance(tok, DeviceTokens):
            raise TypeError("parse_async takes DeviceTokens (Engine.tokenize_device)")
        n = tok.n_samples
        val = self._empty((n,), t.float64)
        ts = self._empty((n,), t.int64) if want_ts else None
        bad = self._empty((2,), t.int64)
        _abi.check(self.lib.pmx_parse_samples(self.ctx, self._ptr(tok.d_text), tok.d_text.numel(),
                                              self._ptr(tok.d_ts_off if want_ts else None), self._ptr(tok.d_val_off), n,
                                              self._ptr(ts), self._ptr(val), self._ptr(bad), self._stream()))
        self.launches += 1
        pool = self.__dict__.setdefault("_bad_pool", [])
        host = pool.pop() if pool else t.empty((2,), dtype=t.int64).pin_memory()
        host.copy_(bad, non_blocking=True)
        ev = t.cuda.Event(); ev.record()

        def finish():
            ev.synchronize()
            b = host.tolist()
            pool.append(host)
            self._raise_bad_token(tok, b, n, want_ts)
        return ts, val, finish

    def _parse_device(self, tok, want_ts):
        n = tok.n_samples
        ts, val, b = self._parse_launch(tok.d_text, tok.d_ts_off if want_ts else None, tok.d_val_off, n, want_ts)
        self._raise_bad_token(tok, b, n, want_ts)
        return ts, val

    def _raise_bad_token(self, tok, b, n, want_ts):
        if b[0] < n:
            o = int(tok.d_val_off[b[0]].item()) & 0xffffffff
            tail = bytes(tok.body[o:o + 4096])               # (bytes, or the uint8 view of a pinned body)
            end = tail.find(b'"')
            err = ValueError(f"could not convert string to float: {tail[:end if end >= 0 else len(tail)].decode('utf-8', 'replace')!r}")
            err.sample_index = int(b[0])
            raise err
        if want_ts and b[1] < n:
            raise ValueError(f"malformed timestamp in sample {b[1]}")

    # ------------------------------------------------------------------ K2
    def align(self, ts, val, off, start_ms, step_ms, n_steps, lookback_ms=DEFAULT_LOOKBACK_MS, want_idx=True):
        t = self.torch
        ts, val, off = self._dev(ts, t.int64), self._dev(val, t.float64), self._dev(off, t.int64)
        S = off.numel() - 1
        out = self._empty((S, n_steps), t.float64)
        idx = self._empty((S, n_steps), t.int32) if want_idx else None
        _abi.check(self.lib.pmx_align(self.ctx, self._ptr(ts), self._ptr(val), self._ptr(off), S, start_ms, step_ms,
                                      n_steps, lookback_ms, self._ptr(out), self._ptr(idx), self._stream()))
        self.launches += 1
        return out, idx

    def bucket(self, ts, val, off, start_ms, step_ms, n_steps):
        """X3, second form: ``resample('<step>')`` bins [t, t+step) labelled left -> (last, mean, count)
        of the non-NaN samples of every bin, [S, n_steps] each."""
        t = self.torch
        ts, val, off = self._dev(ts, t.int64), self._dev(val, t.float64), self._dev(off, t.int64)
        S = off.numel() - 1
        last, mean = self._empty((S, n_steps), t.float64), self._empty((S, n_steps), t.float64)
        cnt = self._empty((S, n_steps), t.int32)
        for a in range(0, S, 65535):                                  # grid.y limit of the kernel
            b = min(S, a + 65535)
            _abi.check(self.lib.pmx_bucket(self.ctx, self._ptr(ts), self._ptr(val), C.c_void_p(off.data_ptr() + 8 * a), b - a,
                                           start_ms, step_ms, n_steps, C.c_void_p(last.data_ptr() + 8 * a * n_steps),
                                           C.c_void_p(mean.data_ptr() + 8 * a * n_steps),
                                           C.c_void_p(cnt.data_ptr() + 4 * a * n_steps), self._stream()))
            self.launches += 1
        return last, mean, cnt

    # ------------------------------------------------------------------ K3
    def rolling(self, x, window_steps, min_periods=1, want_count=True):
        t = self.torch
        x = self._dev(x, t.float64)
        S, T = x.shape
        mn, mx, mean = (self._empty((S, T), t.float64) for _ in range(3))
        cnt = self._empty((S, T), t.int32) if want_count else None
        _abi.check(self.lib.pmx_rolling(self.ctx, self._ptr(x), S, T, window_steps, min_periods, self._ptr(mn),
                                        self._ptr(mx), self._ptr(mean), self._ptr(cnt), self._stream()))
        self.launches += 1
        return mn, mx, mean, cnt

    def rolling_ragged(self, ts, val, off, window_ms, min_periods=1):
        t = self.torch
        ts, val, off = self._dev(ts, t.int64), self._dev(val, t.float64), self._dev(off, t.int64)
        N, S = ts.numel(), off.numel() - 1
        mn, mx, mean = (self._empty((N,), t.float64) for _ in range(3))
        cnt = self._empty((N,), t.int32)
        _abi.check(self.lib.pmx_rolling_ragged(self.ctx, self._ptr(ts), self._ptr(val), self._ptr(off), S, N,
                                               window_ms, min_periods, self._ptr(mn), self._ptr(mx),
                                               self._ptr(mean), self._ptr(cnt), self._stream()))
        self.launches += 1
        return mn, mx, mean, cnt

    # ------------------------------------------------------------------ K4
    def rate(self, x, window_steps, step_ms, mode=0, want_resets=True):
        """Increase / rate (and the number of counter resets) per trailing window of a dense
        matrix.  ``want_resets=False`` skips the n_resets output (4 bytes per point less to write)."""
        t = self.torch
        x = self._dev(x, t.float64)
        S, T = x.shape
        inc, rt = self._empty((S, T), t.float64), self._empty((S, T), t.float64)
        nres = self._empty((S, T), t.int32) if want_resets else None
        _abi.check(self.lib.pmx_rate(self.ctx, self._ptr(x), S, T, window_steps, step_ms, mode, self._ptr(inc),
                                     self._ptr(rt), self._ptr(nres), self._stream()))
        self.launches += 1
        return inc, rt, nres

    def rate_ragged(self, ts, val, off, window_ms, want_resets=True, mode=0):
        """Increase / rate at every sample.  ``want_resets=False`` skips n_resets (and lets the
        tiled kernel serve the call; the reset counts come from the direct kernel)."""
        t = self.torch
        ts, val, off = self._dev(ts, t.int64), self._dev(val, t.float64), self._dev(off, t.int64)
        N, S = ts.numel(), off.numel() - 1
        inc, rt = self._empty((N,), t.float64), self._empty((N,), t.float64)
        nres = self._empty((N,), t.int32) if want_resets else None
        _abi.check(self.lib.pmx_rate_ragged(self.ctx, self._ptr(ts), self._ptr(val), self._ptr(off), S, N, window_ms, mode,
                                            self._ptr(inc), self._ptr(rt), self._ptr(nres), self._stream()))
        self.launches += 1
        return inc, rt, nres

    # ------------------------------------------------------------------ fused
    def fused(self, ts, val, off, start_ms, step_ms, n_steps, window_steps, lookback_ms=DEFAULT_LOOKBACK_MS,
              min_periods=1, out=None):
        """X3+X4+X5 in one kernel.  ``out`` may hold preallocated [S, n_steps] float64 tensors
        (keys aligned, roll_min, roll_max, roll_mean, rate) to avoid allocation in a timed loop."""
        t = self.torch
        if window_steps < 1:
            raise ValueError(f"window_steps must be >= 1 (got {window_steps}: a window shorter than the step)")
        ts, val, off = self._dev(ts, t.int64), self._dev(val, t.float64), self._dev(off, t.int64)
        S = off.numel() - 1
        if window_steps - 1 > _abi.FUSED_MAX_WINDOW_STEPS - 1:
            # longer than the fused kernel's tile allows (e.g. 1 h at a 5 s step): the three kernels
            al, _ = self.align(ts, val, off, start_ms, step_ms, n_steps, lookback_ms, want_idx=False)
            mn, mx, mean, _ = self.rolling(al, window_steps, min_periods, want_count=False)
            _, rt, _ = self.rate(al, window_steps, step_ms, want_resets=False)
            res = {"aligned": al, "roll_min": mn, "roll_max": mx, "roll_mean": mean, "rate": rt}
            if out is not None:
                for k, v in res.items():
                    out[k].copy_(v)
                return out
            return res
        if out is None:
            out = {k: self._empty((S, n_steps), t.float64)
                   for k in ("aligned", "roll_min", "roll_max", "roll_mean", "rate")}
        _abi.check(self.lib.pmx_fused_align_rolling_rate(
            self.ctx, self._ptr(ts), self._ptr(val), self._ptr(off), S, start_ms, step_ms, n_steps, lookback_ms,
            window_steps, min_periods, self._ptr(out["aligned"]), self._ptr(out["roll_min"]),
            self._ptr(out["roll_max"]), self._ptr(out["roll_mean"]), self._ptr(out["rate"]), self._stream()))
        self.launches += 1
        return out

    def fused_from_host(self, h_ts, h_val, h_off, start_ms, step_ms, n_steps, window_steps, dev, h_out,
                        lookback_ms=DEFAULT_LOOKBACK_MS, min_periods=1, n_chunks=8):
        """The fused kernel fed from PINNED host arrays, results returned to pinned host arrays.

        Series are cut into ``n_chunks`` slices; the H2D copy of slice k+1, the kernel on slice k and
        the D2H copy of slice k-1 run on three streams, so the step costs about max(H2D, D2H) of
        the PCIe link instead of their sum.  ``dev``: dict of preallocated device tensors
        (ts, val, off, aligned, roll_min, roll_max, roll_mean, rate); ``h_out``: dict of pinned host
        tensors for the five outputs.  Returns the device outputs (valid after the current stream
        has been synchronised; the call makes the current stream wait for all three)."""
        t = self.torch
        if not hasattr(self, "_streams"):
            self._streams = [t.cuda.Stream(device=self.device) for _ in range(3)]
        s_in, s_run, s_out = self._streams
        cur = t.cuda.current_stream(self.device)
        for s in self._streams:
            s.wait_stream(cur)
        S = h_off.numel() - 1
        keys = ("aligned", "roll_min", "roll_max", "roll_mean", "rate")
        with t.cuda.stream(s_in):
            dev["off"].copy_(h_off, non_blocking=True)
        off_list = h_off.tolist() if S <= 65536 else None
        bounds = [S * k // n_chunks for k in range(n_chunks + 1)]
        for k in range(n_chunks):
            a, b = bounds[k], bounds[k + 1]
            if a == b:
                continue
            lo = off_list[a] if off_list is not None else int(h_off[a])
            hi = off_list[b] if off_list is not None else int(h_off[b])
            with t.cuda.stream(s_in):
                dev["ts"][lo:hi].copy_(h_ts[lo:hi], non_blocking=True)
                dev["val"][lo:hi].copy_(h_val[lo:hi], non_blocking=True)
                ev_in = t.cuda.Event(); ev_in.record(s_in)
            with t.cuda.stream(s_run):
                s_run.wait_event(ev_in)
                # offsets stay absolute: the slice is addressed through off + a over the full ts/val buffers
                _abi.check(self.lib.pmx_fused_align_rolling_rate(
                    self.ctx, self._ptr(dev["ts"]), self._ptr(dev["val"]), C.c_void_p(dev["off"].data_ptr() + 8 * a),
                    b - a, start_ms, step_ms, n_steps, lookback_ms, window_steps, min_periods,
                    *(C.c_void_p(dev[key].data_ptr() + 8 * a * n_steps) for key in keys),
                    C.c_void_p(s_run.cuda_stream)))
                self.launches += 1
                ev_run = t.cuda.Event(); ev_run.record(s_run)
            with t.cuda.stream(s_out):
                s_out.wait_event(ev_run)
                for key in keys:
                    h_out[key][a:b].copy_(dev[key][a:b], non_blocking=True)
        for s in self._streams:
            cur.wait_stream(s)
        return {key: dev[key] for key in keys}

    # ------------------------------------------------------------------ K5
    def group_reduce(self, x, group_id, n_groups, per_step=False, perm_goff=None):
        """NaN-skipping sum/count/min/max/mean per group.  group_id: int per series (host)."""
        t = self.torch
        x = self._dev(x, t.float64)
        S, T = x.shape
        perm, goff = perm_goff if perm_goff is not None else group_index(group_id, n_groups)
        d_perm, d_goff = self._dev(perm, t.int32), self._dev(goff, t.int64)
        shape = (n_groups, T) if per_step else (n_groups,)
        sm, mn, mx, mean = (self._empty(shape, t.float64) for _ in range(4))
        cnt = self._empty(shape, t.int64)
        _abi.check(self.lib.pmx_group_reduce(self.ctx, self._ptr(x), S, T, self._ptr(d_perm), self._ptr(d_goff),
                                             n_groups, 1 if per_step else 0, self._ptr(sm), self._ptr(cnt),
                                             self._ptr(mn), self._ptr(mx), self._ptr(mean), self._stream()))
        self.launches += 1 if per_step else 2
        return sm, cnt, mn, mx, mean

    def group_combine(self, sm, cnt, mn, mx, group):
        """Partials of ``group_reduce`` on every rank -> the same reduction over all ranks' series:
        pmx_group_pack, ONE all-gather, pmx_group_combine (rank order, identical bits everywhere)."""
        import torch.distributed as dist
        t = self.torch
        world = dist.get_world_size(group)
        n = sm.numel()
        key = (n, world)
        cache = self.__dict__.setdefault("_combuf", {})
        if key not in cache:
            cache[key] = (self._empty((4 * n,), t.float64), self._empty((world * 4 * n,), t.float64))
        packed, gathered = cache[key]
        _abi.check(self.lib.pmx_group_pack(self._ptr(sm), self._ptr(cnt), self._ptr(mn), self._ptr(mx), n,
                                           self._ptr(packed), self._stream()))
        dist.all_gather_into_tensor(gathered, packed, group=group)
        o_sm, o_mn, o_mx, o_mean = (t.empty_like(sm) for _ in range(4))
        o_cnt = t.empty_like(cnt)
        _abi.check(self.lib.pmx_group_combine(self._ptr(gathered), world, n, self._ptr(o_sm), self._ptr(o_cnt),
                                              self._ptr(o_mn), self._ptr(o_mx), self._ptr(o_mean), self._stream()))
        self.launches += 2
        return o_sm, o_cnt, o_mn, o_mx, o_mean

    # ------------------------------------------------------------------ instant-path steps
    def pivot_gather(self, val, cell_src):
        t = self.torch
        val = self._dev(val, t.float64)
        src = self._dev(cell_src, t.int32)
        out = self._empty((src.numel(),), t.float64)
        _abi.check(self.lib.pmx_pivot_gather(self.ctx, self._ptr(val), self._ptr(src), src.numel(), self._ptr(out),
                                             self._stream()))
        self.launches += 1
        return out

    def ratio_pct(self, mat, num_col, den_col, out_col):
        """mat[:, out_col] = mat[:, num_col] / mat[:, den_col] * 100 on a row-major [G, C] matrix.
        Returns True if any denominator was zero."""
        t = self.torch
        G, Cc = mat.shape
        flag = self._empty((1,), t.int32)
        base = mat.data_ptr()
        _abi.check(self.lib.pmx_ratio_pct(self.ctx, C.c_void_p(base + 8 * num_col), Cc, C.c_void_p(base + 8 * den_col),
                                          Cc, G, C.c_void_p(base + 8 * out_col), Cc, self._ptr(flag), self._stream()))
        self.launches += 1
        return bool(flag.item())

    def where_gt(self, x, threshold):
        t = self.torch
        x = self._dev(x, t.float64)
        out = self.torch.empty_like(x)
        _abi.check(self.lib.pmx_where_gt(self.ctx, self._ptr(x), x.numel(), float(threshold), self._ptr(out),
                                         self._stream()))
        self.launches += 1
        return out

    # ------------------------------------------------------------------ K6
    def summary_brackets_csr(self, val, off, n_steps, n_seg, qs, perm_goff):
        """First half of ``summary_quantiles`` from the series' raw samples (CSR ``val`` / ``off``): sample and
        brackets, on the current stream -- e.g. a side stream beside the fused kernel.  Follow with
        ``summary_quantiles(x, ..., brackets_ready=True)`` on the aligned matrix of the same series, steps,
        segments and levels, ordered behind this call."""
        t = self.torch
        val, off = self._dev(val, t.float64), self._dev(off, t.int64)
        d_perm, d_goff = self._dev(perm_goff[0], t.int32), self._dev(perm_goff[1], t.int64)
        S = off.numel() - 1
        _abi.check(self.lib.pmx_summary_brackets_csr(self.ctx, self._ptr(val), self._ptr(off), S, n_steps, n_seg, self._ptr(d_perm),
                                                     self._ptr(d_goff), self._ptr(self._quantile_levels(qs)), len(qs),
                                                     self._stream()))
        self.launches += 2 if S > 0 and n_steps > 0 else 1

    def summary_quantiles(self, x, seg, n_seg, qs, perm_goff, brackets_ready=False):
        """One GPU: ``group_reduce(x, ..., perm_goff)`` and ``quantiles(x, seg, n_seg, qs)`` from ONE sweep
        of ``x`` (pmx_summary_quantiles: sampled brackets, counted and verified on the device).  Returns
        ((sum, count, min, max, mean), quantiles dict, failed) -- ``failed`` is an int32 device tensor [1]:
        the outputs are exact iff it is 0 (else some quantiles are NaN and the per-pass selection has to
        run).  At most four quantile levels."""
        t = self.torch
        x = self._dev(x, t.float64)
        S, T = x.shape
        d_seg = self._dev(seg, t.int32)
        d_perm, d_goff = self._dev(perm_goff[0], t.int32), self._dev(perm_goff[1], t.int64)
        n_q = len(qs)
        d_q = self._quantile_levels(qs)
        sm, mn, mx, mean = (self._empty((n_seg,), t.float64) for _ in range(4))
        cnt = self._empty((n_seg,), t.int64)
        outs = {"n": self._empty((n_seg,), t.int64), "lo": self._empty((n_seg, n_q), t.int64),
                "hi": self._empty((n_seg, n_q), t.int64), "x_lo": self._empty((n_seg, n_q), t.float64),
                "x_hi": self._empty((n_seg, n_q), t.float64), "value": self._empty((n_seg, n_q), t.float64)}
        failed = self._empty((1,), t.int32)
        fn = self.lib.pmx_summary_sweep if brackets_ready else self.lib.pmx_summary_quantiles
        _abi.check(fn(
            self.ctx, self._ptr(x), S, T, self._ptr(d_seg), n_seg, self._ptr(d_perm), self._ptr(d_goff), self._ptr(d_q), n_q,
            self._ptr(sm), self._ptr(cnt), self._ptr(mn), self._ptr(mx), self._ptr(mean), self._ptr(outs["n"]),
            self._ptr(outs["lo"]), self._ptr(outs["hi"]), self._ptr(outs["x_lo"]), self._ptr(outs["x_hi"]),
            self._ptr(outs["value"]), self._ptr(failed), self._stream()))
        # sample, bounds, sweep, resolve (which also folds the summary)
        self.launches += (4 if S > 0 and T > 0 else 2) - (0 if not brackets_ready else (1 if S > 0 and T > 0 else 0))   # (a clear instead)
        return (sm, cnt, mn, mx, mean), outs, failed

    def _bsel_buffers(self, n_seg, n_q, world):
        import ctypes as C
        t = self.torch
        key = (n_seg, n_q, world)
        cache = self.__dict__.setdefault("_bsel", {})
        if key not in cache:
            w = [C.c_int(0) for _ in range(3)]
            _abi.check(self.lib.pmx_bsel_widths(C.byref(w[0]), C.byref(w[1]), C.byref(w[2])))
            bw, pw, rw = (int(v.value) for v in w)
            cache[key] = {
                "bounds": self._empty((n_seg * bw,), t.float64), "bounds_all": self._empty((world * n_seg * bw,), t.float64),
                # counters and partial summaries, then the first histograms (u32 pairs)
                "packed": self._empty((n_seg * (pw + n_q * _abi.QUANTILE_BINS // 2),), t.int64),
                "packed_all": self._empty((world * n_seg * (pw + n_q * _abi.QUANTILE_BINS // 2),), t.int64),
                "hist2": self._empty((n_seg * n_q * 2 * _abi.QUANTILE_BINS,), t.int32),
                "rows": self._empty((n_seg * n_q * rw,), t.int64), "rows_all": self._empty((world * n_seg * n_q * rw,), t.int64)}
        return cache[key]

    def summary_begin_csr_ranks(self, val, off, n_steps, n_seg, qs, perm_goff, group):
        """First call of ``summary_quantiles_ranks`` from the series' raw samples, on the current stream (a side
        stream beside the fused kernel); pass ``begun=True`` to ``summary_quantiles_ranks`` afterwards."""
        import torch.distributed as dist
        t = self.torch
        val, off = self._dev(val, t.float64), self._dev(off, t.int64)
        d_perm, d_goff = self._dev(perm_goff[0], t.int32), self._dev(perm_goff[1], t.int64)
        S = off.numel() - 1
        b = self._bsel_buffers(n_seg, len(qs), dist.get_world_size(group))
        _abi.check(self.lib.pmx_bsel_begin_csr(self.ctx, self._ptr(val), self._ptr(off), S, n_steps, n_seg, self._ptr(d_perm),
                                               self._ptr(d_goff), self._ptr(self._quantile_levels(qs)), len(qs),
                                               self._ptr(b["bounds"]), self._stream()))
        self.launches += 2 if S > 0 and n_steps > 0 else 1

    def bsel_region_bytes(self, n_seg, n_q):
        """Bytes per rank of the four exchanges of ``summary_quantiles_ranks`` (for ``sharding.PeerExchange``)."""
        b = self._bsel_buffers(n_seg, n_q, 1)
        return {"bounds": b["bounds"].numel() * 8, "packed": b["packed"].numel() * 8, "hist2": b["hist2"].numel() * 4,
                "rows": b["rows"].numel() * 8}

    def _stage(self, cache, key, fn):
        """Run ``fn`` (a fixed sequence of kernel launches on fixed buffers, no collective inside): eagerly
        twice, then captured into a CUDA graph and replayed -- the launches between two exchanges are
        launch-bound.  Stays eager if the capture is refused."""
        t = self.torch
        ent = cache.setdefault(key, [0, None])
        if ent[1] is not None:
            ent[1].replay()
            return
        if ent[0] != 2:
            fn()
            ent[0] += 1
            return
        try:
            g = t.cuda.CUDAGraph()
            with t.cuda.graph(g, capture_error_mode="thread_local"):
                fn()
            ent[1] = g
            g.replay()
        except Exception:  # noqa: BLE001
            t.cuda.synchronize()
            ent[0] = 3                                         # never again
            fn()

    def summary_quantiles_ranks(self, x, seg, n_seg, qs, perm_goff, group, begun=False, peer=None, stage_graphs=False):
        """``summary_quantiles`` over ALL ranks of ``group`` (every rank passes its own rows): five library
        calls with four small collectives between them (brackets; counters + partial summaries + first
        histogram; second histogram; the keys of the buckets that hold the wanted ranks), one sweep of
        ``x`` on every rank.  ``peer``: a ``sharding.PeerExchange`` with the regions of
        ``bsel_region_bytes`` -- the exchanges then are stores into peer memory + a device-side barrier
        instead of NCCL calls (and the second histograms are gathered and added by the next kernel).
        ``stage_graphs``: ``x``, ``seg`` and ``perm_goff`` are the same device tensors on every call; the
        launches between two exchanges are then replayed from CUDA graphs and the results live in buffers
        that the next call overwrites.
        Returns ((sum, count, min, max, mean) over all ranks, quantiles dict, failed) -- ``failed`` is the
        same on every rank."""
        import ctypes as C
        import torch.distributed as dist
        t = self.torch
        x = self._dev(x, t.float64)
        S, T = x.shape
        d_seg = self._dev(seg, t.int32)
        d_perm, d_goff = self._dev(perm_goff[0], t.int32), self._dev(perm_goff[1], t.int64)
        n_q = len(qs)
        d_q = self._quantile_levels(qs)
        world = dist.get_world_size(group)
        b = self._bsel_buffers(n_seg, n_q, world)

        def results():
            return {"sm": self._empty((n_seg,), t.float64), "mn": self._empty((n_seg,), t.float64),
                    "mx": self._empty((n_seg,), t.float64), "mean": self._empty((n_seg,), t.float64),
                    "cnt": self._empty((n_seg,), t.int64), "failed": self._empty((1,), t.int32),
                    "outs": {"n": self._empty((n_seg,), t.int64), "lo": self._empty((n_seg, n_q), t.int64),
                             "hi": self._empty((n_seg, n_q), t.int64), "x_lo": self._empty((n_seg, n_q), t.float64),
                             "x_hi": self._empty((n_seg, n_q), t.float64), "value": self._empty((n_seg, n_q), t.float64)}}
        gkey = None
        if stage_graphs:
            gkey = (x.data_ptr(), S, T, d_seg.data_ptr(), d_perm.data_ptr(), d_goff.data_ptr(), n_seg, tuple(qs), world, peer is not None)
            graphs = b.setdefault("graphs", {})
            if ("res", gkey) not in graphs:
                graphs[("res", gkey)] = results()
            R = graphs[("res", gkey)]
        else:
            R = results()
        sm, mn, mx, mean, cnt, failed, outs = R["sm"], R["mn"], R["mx"], R["mean"], R["cnt"], R["failed"], R["outs"]
        lib, p, st = self.lib, self._ptr, self._stream

        def stage(name, fn):
            if gkey is None:
                fn()
            else:
                self._stage(b["graphs"], (name, gkey), fn)
        if not begun:
            _abi.check(lib.pmx_bsel_begin(self.ctx, p(x), S, T, p(d_seg), n_seg, p(d_perm), p(d_goff), p(d_q), n_q, p(b["bounds"]), st()))
        if peer is not None:
            bounds_all = peer.all_gather("bounds", b["bounds"])
        else:
            dist.all_gather_into_tensor(b["bounds_all"], b["bounds"], group=group)
            bounds_all = b["bounds_all"]
        stage("sweep", lambda: _abi.check(lib.pmx_bsel_sweep(self.ctx, p(x), S, T, p(d_seg), n_seg, p(d_perm), p(d_goff), n_q,
                                                             p(bounds_all), world, p(b["packed"]), st())))
        if peer is not None:
            packed_all = peer.all_gather("packed", b["packed"])
        else:
            dist.all_gather_into_tensor(b["packed_all"], b["packed"], group=group)
            packed_all = b["packed_all"]
        stage("hist", lambda: _abi.check(lib.pmx_bsel_hist(self.ctx, S, T, n_seg, p(d_q), n_q, p(packed_all), world, p(sm), p(cnt),
                                                           p(mn), p(mx), p(mean), p(b["hist2"]), st())))
        if peer is not None:
            hist2, n_hist = peer.all_gather("hist2", b["hist2"]), world
        else:
            dist.all_reduce(b["hist2"], op=dist.ReduceOp.SUM, group=group)
            hist2, n_hist = b["hist2"], 1
        stage("rows", lambda: _abi.check(lib.pmx_bsel_rows(self.ctx, S, T, n_seg, n_q, p(hist2), n_hist, p(b["rows"]), st())))
        if peer is not None:
            rows_all = peer.all_gather("rows", b["rows"])
        else:
            dist.all_gather_into_tensor(b["rows_all"], b["rows"], group=group)
            rows_all = b["rows_all"]
        stage("finish", lambda: _abi.check(lib.pmx_bsel_finish(self.ctx, S, T, n_seg, n_q, p(rows_all), world, p(cnt), p(outs["n"]),
                                                               p(outs["lo"]), p(outs["hi"]), p(outs["x_lo"]), p(outs["x_hi"]),
                                                               p(outs["value"]), p(failed), st())))
        # sample, bounds | merge, sweep, histogram (+ fold and pack) | pick, histogram | pick | final
        self.launches += (9 if S > 0 and T > 0 else 6) - (0 if not begun else (2 if S > 0 and T > 0 else 1))
        return (sm, cnt, mn, mx, mean), outs, failed

    def defer_flag(self, counter):
        """Copies a device int32 counter to pinned host memory behind the stream; ``deferred_ok`` looks at it."""
        t = self.torch
        pool = self.__dict__.setdefault("_flag_pool", [])
        flag = pool.pop() if pool else t.empty((1,), dtype=t.int32).pin_memory()
        flag.copy_(counter, non_blocking=True)
        ev = t.cuda.Event(); ev.record()
        self.__dict__.setdefault("_deferred", []).append((ev, flag))

    def deferred_ok(self, wait=True):
        """Looks at the counters of earlier ``quantiles(..., defer_check=True)`` calls.  Returns True
        when every finished one was exact.  ``wait=False`` leaves calls whose copy has not arrived
        pending (nothing blocks); ``wait=True`` waits for all of them."""
        pend = self.__dict__.get("_deferred", [])
        ok, keep = True, []
        for ev, flag in pend:
            if not wait and not ev.query():
                keep.append((ev, flag))
                continue
            ev.synchronize()
            ok = ok and int(flag[0]) == 0
            self.__dict__.setdefault("_flag_pool", []).append(flag)
        self._deferred = keep
        return ok

    def _quantile_levels(self, qs):
        """The quantile levels as a device tensor, uploaded once per distinct tuple (a pageable
        host-to-device copy in every call would also forbid capturing the step in a CUDA graph)."""
        key = tuple(float(q) for q in qs)
        cache = self.__dict__.setdefault("_qlevels", {})
        if key not in cache:
            cache[key] = self._dev(np.asarray(key, dtype=np.float64), self.torch.float64)
        return cache[key]

    def quantile_buffers(self, n_seg, n_q):
        t = self.torch
        state = self.torch.zeros((n_seg, 3 + 4 * n_q), dtype=t.int64, device=self.device)
        hist = self._empty((n_seg * 2 * n_q, _abi.QUANTILE_BINS), t.int64)
        return state, hist

    def _tail_buffers(self, n_seg, world):
        t = self.torch
        key = (n_seg, world)
        cache = self.__dict__.setdefault("_tailbuf", {})
        if key not in cache:
            row = _abi.QUANTILE_GATHER_ROW
            cache[key] = (self._empty((n_seg * row,), t.int64), self._empty((world * n_seg * row,), t.int64),
                          self._empty((1,), t.int32))
        return cache[key]

    def quantiles(self, x, seg, n_seg, qs, group=None, buffers=None, row_order=None, summary=None, defer_check=False,
                  list_pass=None):
        """Exact quantiles per segment.  x: [S,T] device; seg: int per series (host or device).
        With ``group`` (a torch.distributed process group over NCCL) the histograms of every pass
        are all-reduced, so the result is the quantile of the union of all ranks' values.
        ``row_order``: optional int32 device tensor of row indices grouped by segment (the perm of
        ``group_index``); lets a CTA keep one set of histograms for long runs of rows.
        ``summary``: optional (count int64[n_seg], min f64[n_seg], max f64[n_seg]) device tensors of
        each segment over ALL ranks (what ``group_reduce`` / ``combine_group_partials`` return for the
        same values): the selection then starts behind the key bits that min and max share and
        saves the first full pass.
        ``defer_check`` (several ranks only): after pass 2 the ranks exchange the few keys that still
        match a target; whether every segment fitted that exchange is a counter on the device.  By
        default the host reads it (one synchronisation per call) and runs the per-pass protocol if it
        is not zero.  With ``defer_check=True`` the call finishes the selection at once, copies the
        counter to pinned host memory behind the stream and leaves looking at it to
        ``deferred_ok()``: the outputs are exact iff that returns True (nothing is recomputed;
        a caller that needs the slow path repeats the call without ``defer_check``).
        ``list_pass``: the pass that keeps the keys still matching a target (2 or 3; default: what
        pmx_quantile_list_pass recommends for this shape).  All ranks must use the same value."""
        t = self.torch
        x = self._dev(x, t.float64)
        S, T = x.shape
        d_seg = self._dev(seg, t.int32)
        d_ord = self._dev(row_order, t.int32) if row_order is not None else None
        n_rows = d_ord.numel() if d_ord is not None else 0
        n_q = len(qs)
        d_q = self._quantile_levels(qs)
        outs = {"n": self._empty((n_seg,), t.int64), "lo": self._empty((n_seg, n_q), t.int64),
                "hi": self._empty((n_seg, n_q), t.int64), "x_lo": self._empty((n_seg, n_q), t.float64),
                "x_hi": self._empty((n_seg, n_q), t.float64), "value": self._empty((n_seg, n_q), t.float64)}
        if group is None and summary is None:
            _abi.check(self.lib.pmx_quantile_select(
                self.ctx, self._ptr(x), S, T, self._ptr(d_seg), n_seg, self._ptr(d_ord), n_rows, self._ptr(d_q), n_q,
                self._ptr(outs["n"]),
                self._ptr(outs["lo"]), self._ptr(outs["hi"]), self._ptr(outs["x_lo"]), self._ptr(outs["x_hi"]),
                self._ptr(outs["value"]), self._stream()))
            self.launches += 10                                 # hist + pick (two passes), collect + list hist (two kernels) + pick, tail, finish
            return outs
        state, hist = buffers if buffers is not None else self.quantile_buffers(n_seg, n_q)
        if summary is None:
            state.zero_()
            passes = range(_abi.QUANTILE_PASSES)
        else:
            cnt, mn, mx = (self._dev(summary[0], t.int64), self._dev(summary[1], t.float64),
                           self._dev(summary[2], t.float64))
            if cnt.numel() != n_seg or mn.numel() != n_seg or mx.numel() != n_seg:
                raise ValueError("summary tensors must have n_seg elements")
            _abi.check(self.lib.pmx_quantile_init(self.ctx, self._ptr(cnt), self._ptr(mn), self._ptr(mx), n_seg,
                                                  self._ptr(d_q), n_q, self._ptr(state), self._stream()))
            self.launches += 1
            passes = range(1, _abi.QUANTILE_PASSES + 1)
        # Several ranks starting from count/min/max: one more digit (read from the candidate lists) before
        # the keys are exchanged, so that 33 key bits are settled by then and a rank's row is short
        # whatever the value range looks like
        if list_pass is None:
            list_pass = int(self.lib.pmx_quantile_list_pass(S, T, n_seg)) if summary is not None else 2
        tail_after = 3 if (group is not None and summary is not None) else list_pass
        for p in passes:
            # hist + pick; from the list pass on the digit comes from the lists (two kernels), and the list pass itself collects first
            self.launches += 2 if p < list_pass else (4 if p == list_pass else 3)
            _abi.check(self.lib.pmx_quantile_hist_lp(self.ctx, self._ptr(x), S, T, self._ptr(d_seg), n_seg,
                                                     self._ptr(d_ord), n_rows, n_q, p, list_pass,
                                                     self._ptr(state), self._ptr(hist), self._stream()))
            n_h = n_seg * (1 if p == 0 else 2 * n_q)
            if group is not None:
                import torch.distributed as dist
                dist.all_reduce(hist[:n_h], op=dist.ReduceOp.SUM, group=group)
            _abi.check(self.lib.pmx_quantile_pick(self.ctx, self._ptr(hist), n_seg, self._ptr(d_q), n_q, p,
                                                  self._ptr(state), self._stream()))
            if p == tail_after and group is None and S > 0 and T > 0:   # one rank: the candidate lists settle the rest
                _abi.check(self.lib.pmx_quantile_tail(self.ctx, self._ptr(x), S, T, self._ptr(d_seg), n_seg, n_q, p + 1,
                                                      self._ptr(state), self._stream()))
                self.launches += 1
                break
            elif p == tail_after and group is not None:
                # several ranks: gather the few keys that still match a target and settle every segment
                # on every rank alike; the later passes (and their all-reduces) run only if a segment
                # did not fit, which all ranks see in the same counter
                import torch.distributed as dist
                world = dist.get_world_size(group)
                rows, gathered, unresolved = self._tail_buffers(n_seg, world)
                if S > 0 and T > 0:
                    _abi.check(self.lib.pmx_quantile_tail_filter(self.ctx, n_seg, n_q, p + 1, self._ptr(state),
                                                                 self._ptr(rows), self._stream()))
                    self.launches += 1
                else:
                    rows.zero_()
                dist.all_gather_into_tensor(gathered, rows, group=group)
                _abi.check(self.lib.pmx_quantile_tail_merge(self.ctx, self._ptr(gathered), world, n_seg, n_q, p + 1,
                                                            self._ptr(state), self._ptr(unresolved), self._stream()))
                self.launches += 1
                if defer_check:
                    self.defer_flag(unresolved)
                    break
                if int(unresolved.item()) == 0:
                    break
        _abi.check(self.lib.pmx_quantile_finish(self.ctx, self._ptr(state), n_seg, self._ptr(d_q), n_q,
                                                self._ptr(outs["n"]), self._ptr(outs["lo"]), self._ptr(outs["hi"]),
                                                self._ptr(outs["x_lo"]), self._ptr(outs["x_hi"]),
                                                self._ptr(outs["value"]), self._stream()))
        self.launches += 1                                  # finish
        return outs
