"""The reference-facing call surface: same names, arguments and error behaviour as app.py.

* ``fetch_gpu_metrics()``            -- drop-in for ``/root/reference/app.py:153-227``: two HTTP GETs,
  then the engine instead of ``json.loads`` + the per-series Python loop + pandas reductions.
  Returns ``(df_pivot, stats)`` or ``(None, None)`` after reporting the error, like the reference.
* ``aggregate_instant(body)``        -- the inner seam (``app.py:178-221``) on a response body.
* ``selected_averages(df, selected)``-- ``app.py:335-345``.
* ``aggregate_range(body, ...)``     -- the range-query counterpart north_star asks for: matrix body in,
  tidy DataFrame (or dict of arrays) out.

All numeric work happens in libpmx.so on the GPU; pandas is only used to assemble the frames
the Streamlit charts consume (object dtype, lexicographic ``gpu_id`` index, as the reference).
"""
from __future__ import annotations

import os

import numpy as np
import pandas as pd

from .engine import DEFAULT_LOOKBACK_MS, Engine, tokenize

PROMETHEUS_METRICS_ENDPOINT = os.environ.get("PROMETHEUS_METRICS_ENDPOINT", "http://localhost:9090/api/v1/query")
PROMETHEUS_METRICS_PODNAME = os.environ.get("PROMETHEUS_METRICS_PODNAME", "prometheus")

import threading

_local = threading.local()
last_error = None          # message of the last failure (what the reference passes to st.error)


def engine() -> Engine:
    """The calling thread's Engine.  Streamlit runs every browser session's script on a thread of
    its own (app.py:326-331: each session loops on fetch_gpu_metrics), and a pmx_ctx serves one
    caller at a time (its scratch is reallocated under a concurrent user), so every thread gets
    its own ctx; they share the device."""
    eng = getattr(_local, "engine", None)
    if eng is None:
        eng = _local.engine = Engine()
    return eng


def _report(msg):
    global last_error
    last_error = msg
    try:
        import streamlit as st
        st.error(msg)
    except Exception:  # noqa: BLE001  (streamlit absent or not in a script run)
        pass


def aggregate_instant(body, eng: Engine | None = None):
    """app.py:178-221 on the GPU.  Raises what the reference's code would raise
    (KeyError for a missing label, ValueError for a malformed number or duplicate key,
    ZeroDivisionError for total_vram == 0); ``fetch_gpu_metrics`` catches them."""
    eng = eng or engine()
    t = eng.torch
    tok = tokenize(body)                                              # app.py:178
    labels = tok.labels()
    if tok.result_type != "vector":
        raise ValueError("instant path expects resultType vector")
    # app.py:183-194 walks the series in order and touches gpu_id, __name__, float(value), card_model:
    # the first failure in THAT order is the one the reference reports.  All values are decoded in
    # one launch; a malformed one is held back until the walk reaches its series.
    bad, val = None, None
    try:
        _, val = eng.parse(tok, want_ts=False) if tok.n_samples else (None, eng._empty((0,), t.float64))
    except ValueError as e:
        bad = e
    models, keys = {}, []
    for i, lab in enumerate(labels):
        gid, name = lab["gpu_id"], lab["__name__"]
        if bad is not None and getattr(bad, "sample_index", 0) == i:
            raise bad
        keys.append((gid, name))
        if gid not in models:
            models[gid] = lab["card_model"]
    if bad is not None:
        raise bad
    if not keys:
        raise KeyError("gpu_id")                                      # pivot on an empty frame (app.py:207)
    if len(set(keys)) != len(keys):
        raise ValueError("Index contains duplicate entries, cannot reshape")
    gids = sorted(models)                                             # pivot sorts the index
    names = sorted({n for _, n in keys})
    for need in ("amd_gpu_used_vram", "amd_gpu_total_vram"):
        if need not in names:
            raise KeyError(need)                                      # app.py:210-212
    cols = names + ["vram_usage_ratio"]
    G, Cn = len(gids), len(cols)
    row = {g: i for i, g in enumerate(gids)}
    col = {n: i for i, n in enumerate(cols)}
    src = np.full(G * Cn, -1, dtype=np.int32)
    for i, (g, n) in enumerate(keys):
        src[row[g] * Cn + col[n]] = i
    mat = eng.pivot_gather(val, src).view(G, Cn)                      # app.py:207
    if eng.ratio_pct(mat, col["amd_gpu_used_vram"], col["amd_gpu_total_vram"], col["vram_usage_ratio"]):
        raise ZeroDivisionError("float division by zero")             # app.py:210-212 on Python floats
    # app.py:216-221: column-wise mean/max/min over cards, NaN skipped, cards in index order
    sm, cnt, mn, mx, mean = eng.group_reduce(mat, np.zeros(G, dtype=np.int32), 1, per_step=True)
    host = mat.cpu().numpy()
    h_mean, h_max, h_min = mean.cpu().numpy()[0], mx.cpu().numpy()[0], mn.cpu().numpy()[0]
    # assemble the frame the charts consume (object dtype like the reference's pivot)
    data = {n: [float(host[r, col[n]]) for r in range(G)] for n in names}
    data["card_model"] = [models[g] for g in gids]
    ordered = sorted(names + ["card_model"])
    df = pd.DataFrame({n: pd.Series(data[n], index=gids, dtype=object) for n in ordered},
                      index=pd.Index(gids, name="gpu_id"))
    df.columns.name = "metric_name"
    df["vram_usage_ratio"] = pd.Series([float(host[r, col["vram_usage_ratio"]]) for r in range(G)],
                                       index=df.index, dtype=object)
    numeric = [c for c in df.columns if c != "card_model"]
    idx = pd.Index(numeric, name="metric_name")
    stats = {
        "mean": pd.Series([float(h_mean[col[c]]) for c in numeric], index=idx, dtype=object),
        "max": pd.Series([float(h_max[col[c]]) for c in numeric], index=idx, dtype=object),
        "min": pd.Series([float(h_min[col[c]]) for c in numeric], index=idx, dtype=object),
    }
    return df, stats


def selected_averages(df_pivot, selected, eng: Engine | None = None):
    """app.py:335-345: column means over the selected cards (selection order), with the power
    mean taken over strictly positive readings when there is any."""
    eng = eng or engine()
    t = eng.torch
    f = df_pivot.loc[selected]                                        # KeyError like the reference
    numeric = [c for c in f.columns if c != "card_model"]
    mat = t.as_tensor(f[numeric].to_numpy(dtype=np.float64), dtype=t.float64, device=eng.device)
    k = len(selected)
    _, _, _, _, mean = eng.group_reduce(mat, np.zeros(k, dtype=np.int32), 1, per_step=True)
    avg = pd.Series([float(v) for v in mean.cpu().numpy()[0]], index=pd.Index(numeric, name=f.columns.name),
                    dtype=object)
    pc = numeric.index("amd_gpu_average_package_power")
    pos = eng.where_gt(mat[:, pc].contiguous(), 0.0).view(k, 1)
    _, cnt, _, _, pmean = eng.group_reduce(pos, np.zeros(k, dtype=np.int32), 1, per_step=True)
    if int(cnt.item()) > 0:
        avg["amd_gpu_average_package_power"] = float(pmean.item())
    return avg


def fetch_gpu_metrics():
    """Drop-in for app.py:153-227 (same two queries, same return contract)."""
    try:
        import requests
        q1 = f"kube_pod_info{{pod=~\".*{PROMETHEUS_METRICS_PODNAME}.*\"}}"
        r1 = requests.get(url=PROMETHEUS_METRICS_ENDPOINT, params={"query": q1})
        r1.raise_for_status()
        import json
        host_ip = json.loads(r1.text)["data"]["result"][0]["metric"]["host_ip"]
        q2 = ("{__name__=~\"amd_gpu_edge_temperature|amd_gpu_gfx_activity|amd_gpu_average_package_power|"
              "amd_gpu_used_vram|amd_gpu_total_vram\", instance=~\"" + host_ip + ":.+\"}")
        r2 = requests.get(url=PROMETHEUS_METRICS_ENDPOINT, params={"query": q2})
        r2.raise_for_status()
        return aggregate_instant(r2.text)
    except Exception as e:  # noqa: BLE001  -- the reference swallows everything (app.py:225-227)
        _report(f"Error fetching GPU metrics: {str(e)}")
        return None, None


def _range_tokens(eng, body, tokenizer):
    """X1 for a range body.  ``tokenizer``: "device" (pmx_tokenize_device: Prometheus' compact
    encoding only), "host" (pmx_tokenize: general JSON) or "auto" (device for bodies of 1 MiB and
    more; a body the device scan refuses -- white space, warnings -- goes to the host walk)."""
    if isinstance(body, str):
        body = body.encode("utf-8")
    if tokenizer == "host" or (tokenizer == "auto" and len(body) < (1 << 20)):
        return tokenize(body)
    if tokenizer == "device":
        return eng.tokenize_device(body)
    try:
        return eng.tokenize_device(body)
    except ValueError:
        return tokenize(body)


def _tidy_arrow(labels, grid, host, n_steps):
    """The tidy frame as a pyarrow Table: the three label columns are
    dictionary-encoded (int32 code per row, one string per distinct value) and the numeric columns
    wrap the arrays that came off the device without a copy, so no per-row Python object or
    fixed-width string is ever made.  ``Table.to_pandas()`` gives categorical label columns."""
    import pyarrow as pa
    S = len(labels)
    cols = {}
    for col, key in (("instance", "instance"), ("gpu_id", "gpu_id"), ("metric_name", "__name__")):
        vals = _label_values(labels, key)
        uniq = sorted(set(vals))
        code = {v: i for i, v in enumerate(uniq)}
        idx = np.repeat(np.fromiter((code[v] for v in vals), dtype=np.int32, count=S), n_steps)
        cols[col] = pa.DictionaryArray.from_arrays(pa.array(idx), pa.array(uniq, type=pa.string()))
    cols["ts"] = pa.array(np.tile(grid, S))
    for col, key in (("value", "aligned"), ("roll_min", "roll_min"), ("roll_max", "roll_max"),
                     ("roll_mean", "roll_mean"), ("rate", "rate")):
        cols[col] = pa.array(host[key].ravel())                    # float64 numpy -> arrow: zero copy
    return pa.table(cols)


def _summary_arrow(summary):
    import pyarrow as pa
    return pa.table({k: pa.array(list(v)) if isinstance(v, list) else pa.array(np.asarray(v)) for k, v in summary.items()})


def aggregate_range(body, start_ms, step_ms, n_steps, window_ms, lookback_ms=DEFAULT_LOOKBACK_MS,
                    quantiles=(0.5, 0.95, 0.99), as_frame=True, eng: Engine | None = None, tokenizer="auto",
                    group_keys=("__name__",)):
    """Range path: ``query_range`` matrix body -> per (series, step) tidy frame + per-metric summary.

    Returns (tidy, summary).  tidy columns: instance, gpu_id, metric_name, ts, value, roll_min,
    roll_max, roll_mean, rate.  summary: one row per metric name with sum/count/min/max/mean of the
    aligned values and the requested quantiles.  With ``as_frame=False`` both are dicts of arrays,
    with ``as_frame="arrow"`` pyarrow Tables (label columns dictionary-encoded, numeric columns
    zero-copy).
    ``tokenizer``: where the body is cut into tokens, see ``_range_tokens``.
    ``group_keys``: the labels that form a summary group -- ``("__name__",)`` is one row per metric
    over the whole answer (what app.py:216-221 computes for one node's instant vector),
    ``("instance", "__name__")`` per node and metric, ``("instance", "gpu_id", "__name__")`` per card
    (SURVEY X6's ``groupby(['instance','gpu_id','__name__'])``); rows come in sorted key order.
    """
    eng = eng or engine()
    tok = _range_tokens(eng, body, tokenizer)
    if hasattr(tok, "d_text"):                                        # tokens on the device: the decode kernel runs
        ts, val, decoded = eng.parse_async(tok)                       # while the host interns the label sets
        labels = tok.labels()
        decoded()                                                     # raises for a malformed token, like parse
    else:
        labels = tok.labels()
        ts, val = eng.parse(tok)
    return _aggregate_series(eng, labels, ts, val, tok.series_off, start_ms, step_ms, n_steps, window_ms,
                             lookback_ms, quantiles, as_frame, group_keys)


def _label_values(labels, key):
    """Label ``key`` of every series as a list of str ("" where missing): from the interned table when
    there is one, from the dicts otherwise."""
    if hasattr(labels, "values") and hasattr(labels, "column"):
        return labels.values(key)
    return [l.get(key, "") for l in labels]


def _group_ids(labels, group_keys):
    """Sorted distinct key tuples and the group id of every series.  With an interned label table
    this is integer work (ids ranked by their strings, np.unique over rows); no per-series dict."""
    if hasattr(labels, "column"):
        order = np.argsort(np.array(labels.strings + [""], dtype=object), kind="stable")
        rank = np.empty(len(order), dtype=np.int64)
        rank[order] = np.arange(len(order))
        empty = len(labels.strings)                                   # stands for "" when a label is missing
        if "" in labels._id:
            empty = labels._id[""]
        cols = []
        for k in group_keys:
            ids = labels.column(k).astype(np.int64)
            ids[ids < 0] = empty
            cols.append(rank[ids])
        strings = labels.strings + [""]
        base = len(strings)
        if cols and base ** len(cols) < (1 << 62):                    # rows as one integer each: a 1-D unique
            key = cols[0]
            for c in cols[1:]:
                key = key * base + c                                  # the first key is the most significant
            uk, inv = np.unique(key, return_inverse=True)
            digits = []
            for _ in cols:
                digits.append(uk % base)
                uk = uk // base
            uniq = np.stack(digits[::-1], axis=1)
        else:
            mat = np.stack(cols, axis=1) if cols else np.zeros((len(labels), 0), dtype=np.int64)
            uniq, inv = np.unique(mat, axis=0, return_inverse=True)
        ranked = [strings[i] for i in order.tolist()]                 # the strings in rank order
        per_key = [[ranked[r] for r in uniq[:, j].tolist()] for j in range(uniq.shape[1])]
        names = list(zip(*per_key)) if per_key else [()] * len(uniq)
        return names, inv.reshape(-1).astype(np.int32)
    tuples = [tuple(l.get(k, "") for k in group_keys) for l in labels]
    names = sorted(set(tuples))
    where = {n: i for i, n in enumerate(names)}
    return names, np.fromiter((where[tp] for tp in tuples), dtype=np.int32, count=len(tuples))


def _summary_column(key):
    return "metric_name" if key == "__name__" else key


def _aggregate_series(eng, labels, ts, val, series_off, start_ms, step_ms, n_steps, window_ms, lookback_ms,
                      quantiles, as_frame, group_keys=("__name__",)):
    """The range path behind the tokens: CSR samples (device) -> tidy frame + per-group summary."""
    w = int(window_ms // step_ms)
    out = eng.fused(ts, val, series_off, start_ms, step_ms, n_steps, w, lookback_ms)
    group_keys = tuple(group_keys)
    names, gid = _group_ids(labels, group_keys)
    from .engine import group_index
    perm_goff = group_index(gid, len(names))
    sm, cnt, mn, mx, mean = eng.group_reduce(out["aligned"], None, len(names), perm_goff=perm_goff)
    qd = eng.quantiles(out["aligned"], gid, len(names), list(quantiles), row_order=perm_goff[0],
                       summary=(cnt, mn, mx))
    host = {k: v.cpu().numpy() for k, v in out.items()}
    summary = {_summary_column(k): [n[i] for n in names] for i, k in enumerate(group_keys)}
    summary.update({"sum": sm.cpu().numpy(), "count": cnt.cpu().numpy(),
                    "min": mn.cpu().numpy(), "max": mx.cpu().numpy(), "mean": mean.cpu().numpy()})
    qv = qd["value"].cpu().numpy()
    for i, q in enumerate(quantiles):
        summary[f"p{int(round(q * 100))}"] = qv[:, i]
    S = len(labels)
    grid = start_ms + step_ms * np.arange(n_steps, dtype=np.int64)
    if as_frame == "arrow":
        return _tidy_arrow(labels, grid, host, n_steps), _summary_arrow(summary)
    tidy = {"instance": np.repeat(_label_values(labels, "instance"), n_steps),
            "gpu_id": np.repeat(_label_values(labels, "gpu_id"), n_steps),
            "metric_name": np.repeat(_label_values(labels, "__name__"), n_steps),
            "ts": np.tile(grid, S), "value": host["aligned"].ravel(), "roll_min": host["roll_min"].ravel(),
            "roll_max": host["roll_max"].ravel(), "roll_mean": host["roll_mean"].ravel(),
            "rate": host["rate"].ravel()}
    if not as_frame:
        return tidy, summary
    return pd.DataFrame(tidy), pd.DataFrame(summary)


# ----------------------------------------------------------------------------- range client
PROMETHEUS_MAX_POINTS = 11_000     # Prometheus refuses query_range answers with more points per series


def _range_endpoint():
    ep = PROMETHEUS_METRICS_ENDPOINT
    return ep + "_range" if ep.endswith("/query") else ep.rstrip("/") + "/query_range"


def merge_range_chunks(eng, chunks):
    """Concatenate the answers to consecutive sub-ranges of one query_range into one CSR.

    ``chunks``: [(labels, ts, val, series_off)] in time order, ts/val device tensors.  A series is
    identified by its label set; it may be missing from some chunks.  Returns (labels, ts, val,
    series_off) with every series' samples in time order.  Sample order inside a series follows
    the chunk order, so the chunks must not overlap in time."""
    import json
    t = eng.torch
    index, labels = {}, []
    for lab, _, _, _ in chunks:
        for l in lab:
            k = json.dumps(l, sort_keys=True, separators=(",", ":"))
            if k not in index:
                index[k] = len(labels)
                labels.append(l)
    S = len(labels)
    total = np.zeros(S + 1, dtype=np.int64)
    gids = []
    for lab, _, _, off in chunks:
        g = np.fromiter((index[json.dumps(l, sort_keys=True, separators=(",", ":"))] for l in lab), dtype=np.int64,
                        count=len(lab))
        gids.append(g)
        np.add.at(total, g + 1, np.diff(np.asarray(off, dtype=np.int64)))
    series_off = np.cumsum(total)
    n = int(series_off[-1])
    ts_all = eng._empty((n,), t.int64)
    val_all = eng._empty((n,), t.float64)
    filled = np.zeros(S, dtype=np.int64)                 # samples of each series already placed
    from . import _abi
    for (lab, ts, val, off), g in zip(chunks, gids):
        off = np.asarray(off, dtype=np.int64)
        lens = np.diff(off)
        if len(lens) == 0 or int(off[-1]) == 0:
            continue
        # local series s goes to series_off[g] + filled[g]: one kernel copies every series to its place
        d = t.as_tensor(np.concatenate([off, series_off[g] + filled[g]]), device=eng.device)
        _abi.check(eng.lib.pmx_csr_copy(eng.ctx, eng._ptr(ts), eng._ptr(val), eng._ptr(d), eng._ptr(d[len(off):]), len(lens),
                                        eng._ptr(ts_all), eng._ptr(val_all), eng._stream()))
        eng.launches += 1
        filled[g] += lens
    return labels, ts_all, val_all, series_off


def fetch_gpu_metrics_range(start_s, end_s, step_s, window_s, lookback_s=DEFAULT_LOOKBACK_MS // 1000,
                            quantiles=(0.5, 0.95, 0.99), as_frame=True, session=None,
                            max_points=PROMETHEUS_MAX_POINTS, eng: Engine | None = None, tokenizer="auto",
                            group_keys=("__name__",)):
    """The range counterpart of ``fetch_gpu_metrics`` (app.py:153-227): the same host lookup and metric
    selector, sent to ``/api/v1/query_range`` in sub-ranges of at most ``max_points`` steps (Prometheus'
    limit per series), merged per label set and aggregated on the GPU.  Returns ``(tidy, summary)`` as
    ``aggregate_range`` does, or ``(None, None)`` after reporting the error, like the reference.
    ``session``: anything with ``get(url=..., params=...)`` (default: the ``requests`` module)."""
    try:
        import json
        if session is None:
            import requests as session
        q1 = f"kube_pod_info{{pod=~\".*{PROMETHEUS_METRICS_PODNAME}.*\"}}"
        r1 = session.get(url=PROMETHEUS_METRICS_ENDPOINT, params={"query": q1})
        r1.raise_for_status()
        host_ip = json.loads(r1.text)["data"]["result"][0]["metric"]["host_ip"]
        q2 = ("{__name__=~\"amd_gpu_edge_temperature|amd_gpu_gfx_activity|amd_gpu_average_package_power|"
              "amd_gpu_used_vram|amd_gpu_total_vram\", instance=~\"" + host_ip + ":.+\"}")
        eng = eng or engine()
        n_steps = int((end_s - start_s) // step_s) + 1
        chunks = []
        for k0 in range(0, n_steps, max_points):
            k1 = min(k0 + max_points, n_steps) - 1
            r = session.get(url=_range_endpoint(), params={"query": q2, "start": start_s + k0 * step_s,
                                                            "end": start_s + k1 * step_s, "step": step_s})
            r.raise_for_status()
            body = r.content if hasattr(r, "content") else r.text
            tok = _range_tokens(eng, body, tokenizer)
            if tok.result_type != "matrix":
                raise ValueError("query_range must answer with resultType matrix")
            ts, val = eng.parse(tok) if tok.n_samples else (eng._empty((0,), eng.torch.int64), eng._empty((0,), eng.torch.float64))
            chunks.append((tok.labels(), ts, val, tok.series_off))
        labels, ts, val, off = merge_range_chunks(eng, chunks) if len(chunks) != 1 else chunks[0]
        return _aggregate_series(eng, labels, ts, val, off, int(round(start_s * 1000)), int(round(step_s * 1000)), n_steps,
                                 int(round(window_s * 1000)), int(round(lookback_s * 1000)), quantiles, as_frame,
                                 group_keys)
    except Exception as e:  # noqa: BLE001  -- the reference swallows everything (app.py:225-227)
        _report(f"Error fetching GPU metrics: {str(e)}")
        return None, None


def range_traces(tidy, column="value", metric=None):
    """Plotly-style line traces from a tidy frame (pandas DataFrame, dict of arrays or pyarrow Table):
    one ``{"type": "scatter", "mode": "lines", "name", "x", "y"}`` per (instance, gpu_id, metric),
    sliced from the frame's arrays (rows of one series are contiguous); ``create_visualization``
    (app.py:234-245) stays as it is and these traces feed ``go.Figure(data=...)``."""
    if hasattr(tidy, "to_pandas"):
        tidy = tidy.to_pandas()
    cols = {k: np.asarray(tidy[k]) for k in ("instance", "gpu_id", "metric_name", "ts", column)}
    n = len(cols["ts"])
    if n == 0:
        return []
    key = np.char.add(np.char.add(cols["instance"].astype(str), "/gpu"), cols["gpu_id"].astype(str))
    full = np.char.add(np.char.add(key, " "), cols["metric_name"].astype(str))
    cut = np.flatnonzero(full[1:] != full[:-1]) + 1
    starts = np.concatenate([[0], cut]); ends = np.concatenate([cut, [n]])
    traces = []
    for a, b in zip(starts, ends):
        if metric is not None and cols["metric_name"][a] != metric:
            continue
        traces.append({"type": "scatter", "mode": "lines", "name": str(full[a]),
                       "x": cols["ts"][a:b], "y": cols[column][a:b]})
    return traces
