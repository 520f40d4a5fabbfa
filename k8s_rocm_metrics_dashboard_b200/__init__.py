"""B200-native engine for the Prometheus aggregate path of k8s-rocm-metrics-dashboard.

Public surface (mirrors the reference's data path, /root/reference/app.py:153-227, :335-345):

    from k8s_rocm_metrics_dashboard_b200 import fetch_gpu_metrics, aggregate_instant, aggregate_range

Numeric work runs in ``libpmx.so`` (hand-written sm_100a CUDA behind the C ABI in
``include/pmx.h``).  There is no CPU fallback: without the library or without a GPU the
calls raise.
"""
from .engine import DEFAULT_LOOKBACK_MS, DeviceTokens, Engine, Tokens, group_index, tokenize  # noqa: F401
from .cache import SeriesCache  # noqa: F401
from .adapter import (aggregate_instant, aggregate_range, fetch_gpu_metrics,  # noqa: F401
                      fetch_gpu_metrics_range, merge_range_chunks, range_traces, selected_averages)

__all__ = ["Engine", "Tokens", "DeviceTokens", "tokenize", "group_index", "fetch_gpu_metrics", "aggregate_instant",
           "aggregate_range", "selected_averages", "fetch_gpu_metrics_range", "merge_range_chunks", "range_traces", "SeriesCache", "DEFAULT_LOOKBACK_MS"]
