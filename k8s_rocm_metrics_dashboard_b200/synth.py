"""Seeded synthetic Prometheus data of the shapes BASELINE.json names.

Host-side numpy generators (tests, CPU baseline, small benches) and a torch device-side
generator for the bench shapes that are too large to build on the host.  Series are the
per-card node-exporter series the reference queries (``app.py:168-171``) extended with
further gauge/counter names; labels are ``instance``, ``gpu_id``, ``card_model``, ``__name__``.
"""
from __future__ import annotations

import json
import math

import numpy as np

START_MS = 1_700_000_000_000
STEP_MS = 15_000
TOTAL_VRAM = 68_702_699_520
STALE_NAN_BITS = 0x7FF0000000000002

# the five names the reference queries (app.py:169-170) first, then extensions
METRIC_NAMES = [
    "amd_gpu_edge_temperature", "amd_gpu_gfx_activity", "amd_gpu_average_package_power",
    "amd_gpu_used_vram", "amd_gpu_total_vram", "amd_gpu_energy_counter",
    "amd_gpu_junction_temperature", "amd_gpu_umc_activity", "amd_gpu_pcie_rx_bytes_total",
    "amd_gpu_memory_temperature", "amd_gpu_mm_activity", "amd_gpu_pcie_tx_bytes_total",
    "amd_gpu_clock_sclk", "amd_gpu_clock_mclk", "amd_gpu_ecc_correct_total", "amd_gpu_fan_speed",
]
# kind per name: t=temperature, a=activity, p=power, u=used_vram, c=const total, k=counter
METRIC_KINDS = "tapuck" + "tak" + "tak" + "ttka"
CARD_MODELS = ["102-D65209-00", "102-G30211-0C", "102-G30219-00"]     # app.py:26-30 keys

CONFIGS = {
    # name: (nodes, cards, metrics, steps)
    "config1": (4, 8, 6, 240),
    "config2": (64, 8, 12, 5760),
    "config3": (1024, 8, 16, 40320),
    "config4": (64, 8, 12, 5760),
}


def series_labels(n_nodes, n_cards, n_metrics):
    """Label sets in the order series are laid out: node-major, then card, then metric."""
    labels = []
    for n in range(n_nodes):
        inst = f"10.0.{n // 250}.{n % 250 + 1}:5000"
        for c in range(n_cards):
            for m in range(n_metrics):
                labels.append({"__name__": METRIC_NAMES[m % len(METRIC_NAMES)] if m < len(METRIC_NAMES)
                               else f"amd_gpu_metric_{m}",
                               "instance": inst, "gpu_id": str(c),
                               "card_model": CARD_MODELS[(n + c) % len(CARD_MODELS)]})
    return labels


def _kind_values(kind, rng, T, all_counters=False, reset_prob=0.0):
    if all_counters or kind == "k":
        inc = rng.uniform(0, 1e4, T)
        if reset_prob > 0:
            v = np.empty(T)
            resets = rng.random(T) < reset_prob
            resets[0] = False
            acc = rng.uniform(0, 1e6)
            fresh = rng.uniform(0, 10, T)
            for i in range(T):
                acc = fresh[i] if resets[i] else acc + inc[i]
                v[i] = acc
            return v
        return rng.uniform(0, 1e6) + np.cumsum(inc)
    if kind == "t":
        return np.round(rng.uniform(30, 95, T), 3)
    if kind == "a":
        return rng.integers(0, 101, T).astype(np.float64)
    if kind == "p":
        v = np.round(rng.uniform(0, 750, T), 3)
        v[rng.random(T) < 0.10] = 0.0
        return v
    if kind == "u":
        return rng.integers(0, TOTAL_VRAM + 1, T).astype(np.float64)
    if kind == "c":
        return np.full(T, float(TOTAL_VRAM))
    raise ValueError(kind)


def dense(n_nodes, n_cards, n_metrics, n_steps, seed=0, nan_frac=0.0, all_counters=False,
          reset_prob=0.0, phase=True, kinds=None):
    """Regularly scraped series (configs 1-4) as CSR arrays.

    Each series has one sample per step at ``START + k*STEP + phase_s`` where ``phase_s`` is a
    per-series scrape offset in [0, STEP) ms (0 when ``phase`` is False).
    Returns dict(labels, ts int64[N], val float64[N], off int64[S+1], kinds str[S]).
    """
    rng = np.random.default_rng(seed)
    S = n_nodes * n_cards * n_metrics
    kinds = kinds or METRIC_KINDS
    ts = np.empty((S, n_steps), dtype=np.int64)
    val = np.empty((S, n_steps), dtype=np.float64)
    klist = []
    base = START_MS + STEP_MS * np.arange(n_steps, dtype=np.int64)
    for s in range(S):
        m = s % n_metrics
        kind = kinds[m % len(kinds)]
        klist.append("k" if all_counters else kind)
        ph = int(rng.integers(0, STEP_MS)) if phase else 0
        ts[s] = base + ph
        val[s] = _kind_values(kind, rng, n_steps, all_counters, reset_prob)
    if nan_frac > 0:
        val[rng.random(val.shape) < nan_frac] = np.nan
    off = np.arange(S + 1, dtype=np.int64) * n_steps
    return {"labels": series_labels(n_nodes, n_cards, n_metrics), "ts": ts.ravel(),
            "val": val.ravel(), "off": off, "kinds": klist,
            "start_ms": START_MS, "step_ms": STEP_MS, "n_steps": n_steps}


def config(name, seed=0, scale=1.0):
    """A named BASELINE.json config, optionally scaled down in node count for CPU-sized tests."""
    nodes, cards, metrics, steps = CONFIGS[name]
    nodes = max(1, int(round(nodes * scale)))
    if name == "config1":
        return dense(nodes, cards, metrics, steps, seed)
    if name == "config2":
        return dense(nodes, cards, metrics, steps, seed, nan_frac=1e-3, kinds="ttk")
    if name == "config3":
        return dense(nodes, cards, metrics, steps, seed, nan_frac=1e-3, kinds="ttk")
    if name == "config4":
        return dense(nodes, cards, metrics, steps, seed, all_counters=True, reset_prob=1e-3)
    raise KeyError(name)


def ragged(n_series, seed=0, mean_len=400, max_len=5760, nan_frac=1e-3, stale_frac=5e-4,
           gap_frac=0.01):
    """Config 5: high-cardinality ragged series with irregular timestamps.

    length ~ 1 + Geometric(mean ``mean_len``) capped at ``max_len``; gaps 15 s +- U(0, 2 s);
    ``gap_frac`` of the gaps stretched to U(6, 30) min; NaN samples and stale-marker samples.
    """
    rng = np.random.default_rng(seed)
    lens = np.minimum(1 + rng.geometric(1.0 / mean_len, n_series), max_len).astype(np.int64)
    off = np.zeros(n_series + 1, dtype=np.int64)
    np.cumsum(lens, out=off[1:])
    N = int(off[-1])
    gaps = STEP_MS + rng.integers(-2000, 2001, N)
    stretch = rng.random(N) < gap_frac
    gaps[stretch] = rng.integers(6 * 60_000, 30 * 60_000 + 1, int(stretch.sum()))
    first = np.zeros(N, dtype=bool); first[off[:-1]] = True
    gaps[first] = 0
    csum = np.cumsum(gaps)
    base = csum - np.repeat(csum[off[:-1]], lens)
    start = START_MS + rng.integers(0, 3_600_000, n_series)
    ts = (base + np.repeat(start, lens)).astype(np.int64)
    val = np.round(rng.uniform(0, 1000, N), 3)
    val[rng.random(N) < nan_frac] = np.nan
    bits = val.view(np.uint64)
    bits[rng.random(N) < stale_frac] = STALE_NAN_BITS
    labels = [{"__name__": "amd_gpu_gfx_activity", "instance": f"n{s // 8}:5000", "gpu_id": str(s % 8),
               "card_model": CARD_MODELS[s % 3]} for s in range(min(n_series, 4096))]
    return {"labels": labels, "ts": ts, "val": val, "off": off}


def ragged_device(n_series, seed=0, mean_len=400, max_len=5760, nan_frac=1e-3, stale_frac=5e-4,
                  gap_frac=0.01, device="cuda"):
    """Config 5 at its stated size (1 000 000 series, ~400 M samples) generated on the GPU with
    torch (plumbing only; seeded): the distributions of ``ragged``.  Returns (ts int64[N],
    val float64[N], off int64[S+1]) on the device."""
    import math
    import torch
    g = torch.Generator(device=device); g.manual_seed(seed)
    u = torch.rand(n_series, generator=g, device=device, dtype=torch.float64).clamp_(1e-12, 1.0)
    lens = torch.floor(torch.log(u) / math.log1p(-1.0 / mean_len)).to(torch.int64) + 2      # 1 + Geometric(1/mean)
    lens.clamp_(1, max_len)
    off = torch.zeros(n_series + 1, dtype=torch.int64, device=device)
    torch.cumsum(lens, 0, out=off[1:])
    N = int(off[-1])
    gaps = STEP_MS + torch.randint(-2000, 2001, (N,), generator=g, device=device, dtype=torch.int64)
    stretch = torch.rand(N, generator=g, device=device) < gap_frac
    long = torch.randint(6 * 60_000, 30 * 60_000 + 1, (N,), generator=g, device=device, dtype=torch.int64)
    gaps = torch.where(stretch, long, gaps)
    del stretch, long
    gaps[off[:-1]] = 0
    csum = torch.cumsum(gaps, 0)
    del gaps
    start = START_MS + torch.randint(0, 3_600_000, (n_series,), generator=g, device=device, dtype=torch.int64)
    ts = csum + torch.repeat_interleave(start - csum[off[:-1]], lens)
    del csum
    val = torch.round(torch.rand(N, generator=g, device=device, dtype=torch.float64) * 1e6) / 1e3
    val = torch.where(torch.rand(N, generator=g, device=device) < nan_frac, torch.full_like(val, float("nan")), val)
    stale = torch.rand(N, generator=g, device=device) < stale_frac
    bits = val.view(torch.int64)
    bits[stale] = STALE_NAN_BITS
    return ts.contiguous(), val.contiguous(), off


# ----------------------------------------------------------------------------- text
def format_value(v):
    """Prometheus value text: shortest round-trip decimal, ``NaN``, ``+Inf``, ``-Inf``."""
    if v != v:
        return "NaN"
    if math.isinf(v):
        return "+Inf" if v > 0 else "-Inf"
    return repr(float(v))


def format_ts(ts_ms):
    """Prometheus timestamp text: seconds with up to 3 decimals (Go ``FormatFloat(-1)`` style)."""
    q, r = divmod(int(ts_ms), 1000)
    if r == 0:
        return str(q)
    return f"{q}.{r:03d}".rstrip("0")


def to_matrix_json(labels, ts, val, off):
    """A Prometheus ``query_range`` response body (resultType matrix) as bytes."""
    parts = []
    for s in range(len(off) - 1):
        a, b = int(off[s]), int(off[s + 1])
        vals = ",".join(f'[{format_ts(ts[i])},"{format_value(val[i])}"]' for i in range(a, b))
        parts.append('{"metric":' + json.dumps(labels[s], separators=(",", ":")) + ',"values":[' + vals + "]}")
    return ('{"status":"success","data":{"resultType":"matrix","result":[' + ",".join(parts) + "]}}").encode()


def to_vector_json(labels, ts_ms, values):
    """A Prometheus instant-query response body (resultType vector) as bytes."""
    parts = ['{"metric":' + json.dumps(l, separators=(",", ":")) + f',"value":[{format_ts(ts_ms)},"{format_value(v)}"]}}'
             for l, v in zip(labels, values)]
    return ('{"status":"success","data":{"resultType":"vector","result":[' + ",".join(parts) + "]}}").encode()


def instant_snapshot(node=0, n_cards=8, seed=0, metrics=5):
    """One node's instant vector for the true reference path (app.py:167-178): the five
    amd_gpu_* series per card.  Returns (labels, values, body bytes)."""
    d = dense(node + 1, n_cards, 6, 1, seed, phase=False)
    sel = [i for i, l in enumerate(d["labels"])
           if l["instance"] == d["labels"][-1]["instance"] and l["__name__"] in METRIC_NAMES[:metrics]]
    labels = [d["labels"][i] for i in sel]
    values = [float(d["val"][i]) for i in sel]
    return labels, values, to_vector_json(labels, START_MS, values)


# ----------------------------------------------------------------------------- device
def dense_device(n_series, n_steps, n_metrics=12, seed=0, nan_frac=1e-3, all_counters=False,
                 reset_prob=0.0, device="cuda"):
    """Config 2/3/4-shaped data generated on the GPU with torch (plumbing only; seeded).

    Same distributions as ``dense``: metric kind cycles temperature / temperature / counter
    (``all_counters`` for config 4).  Returns (ts int64[S*T], val float64[S*T], off int64[S+1]).
    """
    import torch
    g = torch.Generator(device=device); g.manual_seed(seed)
    S, T = n_series, n_steps
    phase = torch.randint(0, STEP_MS, (S, 1), generator=g, device=device, dtype=torch.int64)
    ts = START_MS + STEP_MS * torch.arange(T, device=device, dtype=torch.int64)[None, :] + phase
    u = torch.rand((S, T), generator=g, device=device, dtype=torch.float64)
    m = torch.arange(S, device=device) % n_metrics
    is_counter = torch.ones(S, dtype=torch.bool, device=device) if all_counters else (m % 3 == 2)
    gauge = torch.round((30 + 65 * u) * 1000) / 1000
    inc = u * 1e4
    if reset_prob > 0:
        r = torch.rand((S, T), generator=g, device=device) < reset_prob
        r[:, 0] = False
        seg = torch.cumsum(r.to(torch.int64), dim=1)
        cs = torch.cumsum(inc, dim=1)
        # value = cumulative sum since the last reset (the reset sample starts a fresh small value)
        base = torch.where(r, cs - inc * 1e-3, torch.zeros_like(cs))
        base = torch.cummax(base, dim=1).values
        counter = cs - base
        del seg
    else:
        counter = torch.cumsum(inc, dim=1) + 1e6 * torch.rand((S, 1), generator=g, device=device, dtype=torch.float64)
    val = torch.where(is_counter[:, None], counter, gauge)
    if nan_frac > 0:
        val = torch.where(torch.rand((S, T), generator=g, device=device) < nan_frac,
                          torch.full_like(val, float("nan")), val)
    off = torch.arange(S + 1, device=device, dtype=torch.int64) * T
    return ts.reshape(-1).contiguous(), val.reshape(-1).contiguous(), off
