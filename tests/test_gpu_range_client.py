"""fetch_gpu_metrics_range against a fake Prometheus: the range is fetched in
sub-ranges below the point limit, merged per label set on the device and must give exactly what
aggregate_range gives for the whole range in one body."""
import json

import numpy as np
import pytest

from k8s_rocm_metrics_dashboard_b200 import adapter, synth

pytestmark = pytest.mark.gpu


class _Resp:
    def __init__(self, body):
        self.content = body
        self.text = body.decode()

    def raise_for_status(self):
        pass


class FakePrometheus:
    """Answers the host lookup and query_range from one synthetic data set.  Series in ``late``
    have no samples before ``late_from_s`` (a pod that started later)."""

    def __init__(self, d, late=(), late_from_s=None):
        self.d, self.late, self.late_from_s = d, set(late), late_from_s
        self.calls = []

    def body(self, start_s, end_s):
        d = self.d
        lab, ts, val, off = [], [], [], [0]
        for s in range(len(d["off"]) - 1):
            a, b = int(d["off"][s]), int(d["off"][s + 1])
            t = d["ts"][a:b]
            keep = (t >= int(start_s * 1000)) & (t <= int(end_s * 1000))
            if s in self.late:
                keep &= t >= int(self.late_from_s * 1000)
            if not keep.any():
                continue                                   # Prometheus leaves out series without samples
            lab.append(d["labels"][s]); ts.append(t[keep]); val.append(d["val"][a:b][keep]); off.append(off[-1] + int(keep.sum()))
        ts = np.concatenate(ts) if ts else np.zeros(0, dtype=np.int64)
        val = np.concatenate(val) if val else np.zeros(0)
        return synth.to_matrix_json(lab, ts, val, np.array(off, dtype=np.int64))

    def get(self, url, params):
        self.calls.append((url, dict(params)))
        if url.endswith("/query"):
            ip = self.d["labels"][0]["instance"].split(":")[0]
            return _Resp(json.dumps({"status": "success", "data": {"resultType": "vector", "result": [
                {"metric": {"host_ip": ip}, "value": [0, "1"]}]}}).encode())
        assert url.endswith("/query_range")
        # Prometheus evaluates at start, start+step, ...: a sample belongs to the sub-range whose
        # evaluation times it can serve; cutting at (start - step, end] keeps every sample once
        return _Resp(self.body(params["start"] - params["step"] + 1e-3, params["end"]))


def _frames_equal(a, b, rtol=0.0):
    assert list(a.columns) == list(b.columns) and len(a) == len(b)
    for c in a.columns:
        if a[c].dtype.kind in "fi":
            if rtol and a[c].dtype.kind == "f":
                np.testing.assert_allclose(a[c].to_numpy(), b[c].to_numpy(), rtol=rtol, equal_nan=True)
            else:
                np.testing.assert_array_equal(a[c].to_numpy(), b[c].to_numpy())
        else:
            assert a[c].astype(str).tolist() == b[c].astype(str).tolist()


def test_chunked_range_equals_one_body(eng):
    d = synth.dense(1, 2, 5, 300, seed=14, nan_frac=0.01, phase=False)
    start_s, step_s = d["start_ms"] / 1000, d["step_ms"] / 1000
    end_s = start_s + 299 * step_s
    whole = adapter.aggregate_range(FakePrometheus(d).body(start_s - step_s, end_s), d["start_ms"], d["step_ms"], 300,
                                    900_000, eng=eng)
    fake = FakePrometheus(d, late={3}, late_from_s=start_s + 130 * step_s)
    ref = adapter.aggregate_range(fake.body(start_s - step_s, end_s), d["start_ms"], d["step_ms"], 300, 900_000, eng=eng)
    for max_points, n_calls in ((11_000, 1), (100, 3), (64, 5)):
        f = FakePrometheus(d, late={3}, late_from_s=start_s + 130 * step_s)
        tidy, summary = adapter.fetch_gpu_metrics_range(start_s, end_s, step_s, 900, session=f, max_points=max_points, eng=eng)
        assert adapter.last_error is None or tidy is not None, adapter.last_error
        assert sum(1 for u, _ in f.calls if u.endswith("/query_range")) == n_calls
        _frames_equal(tidy.sort_values(["instance", "gpu_id", "metric_name", "ts"], kind="stable").reset_index(drop=True),
                      ref[0].sort_values(["instance", "gpu_id", "metric_name", "ts"], kind="stable").reset_index(drop=True))
        _frames_equal(summary, ref[1], rtol=1e-12)     # a series first seen in a later chunk is summed last
    assert len(whole[0]) == len(ref[0])
    # the chunked answer against the oracle on the same samples (series 3 starts late)
    from oracle import prom_oracle as O
    from tests.util import assert_bits_equal, assert_close
    olabels, ots, oval, ooff = O.tokenize_matrix(fake.body(start_s - step_s, end_s))
    oref = O.aggregate_range(ots, oval, ooff, d["start_ms"], d["step_ms"], 300, 900_000)
    for s, lab in enumerate(olabels):
        rows = tidy[(tidy["instance"] == lab["instance"]) & (tidy["gpu_id"] == lab["gpu_id"]) & (tidy["metric_name"] == lab["__name__"])]
        assert_bits_equal(rows["value"].to_numpy(), oref["aligned"][s], "chunked aligned")
        assert_close(rows["roll_mean"].to_numpy(), oref["roll_mean"][s], scale=O.rolling_scale(oref["aligned"][s:s + 1], 60)[0],
                     what="chunked roll_mean")
        assert_close(rows["rate"].to_numpy(), oref["rate"][s], rtol=1e-12, what="chunked rate")


def test_range_traces_slice_the_frame_per_series(eng):
    d = synth.dense(1, 2, 5, 60, seed=15, phase=False)
    start_s, step_s = d["start_ms"] / 1000, d["step_ms"] / 1000
    f = FakePrometheus(d)
    tidy, _ = adapter.fetch_gpu_metrics_range(start_s, start_s + 59 * step_s, step_s, 300, session=f, eng=eng)
    traces = adapter.range_traces(tidy, "roll_mean", metric=d["labels"][0]["__name__"])
    assert len(traces) == 2 and all(len(t["x"]) == 60 and t["type"] == "scatter" for t in traces)
    one = tidy[(tidy["gpu_id"] == d["labels"][0]["gpu_id"]) & (tidy["metric_name"] == d["labels"][0]["__name__"])]
    np.testing.assert_array_equal(traces[0]["y"], one["roll_mean"].to_numpy())


def test_errors_are_reported_like_the_reference(eng):
    class Broken:
        def get(self, url, params):
            raise OSError("connection refused")
    assert adapter.fetch_gpu_metrics_range(0, 60, 15, 60, session=Broken(), eng=eng) == (None, None)
    assert "connection refused" in adapter.last_error
