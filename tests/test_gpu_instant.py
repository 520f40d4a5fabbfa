"""GPU: the reference's own path (app.py:153-227, :335-345) through the adapter, against what the
UNMODIFIED reference returned (tests/golden/instant_golden.json)."""
import json
import os
import struct
import types

import pytest

from k8s_rocm_metrics_dashboard_b200 import adapter, synth

pytestmark = pytest.mark.gpu


def fbits(x):
    return "%016x" % struct.unpack("<Q", struct.pack("<d", float(x)))[0]


def zero_eq(a, b):
    """bit-equal, with -0.0 == 0.0 (tie order of +-0 in object-dtype max/min)."""
    return a == b or {a, b} <= {"0000000000000000", "8000000000000000"}


@pytest.fixture(scope="module")
def golden(golden_dir):
    return json.load(open(os.path.join(golden_dir, "instant_golden.json")))


def test_aggregate_instant_matches_reference(eng, golden):
    for case in golden["cases"]:
        if "error" in case:
            with pytest.raises(Exception) as ei:
                adapter.aggregate_instant(case["body"].encode(), eng)
            assert str(ei.value) in case["error"], (case["name"], str(ei.value), case["error"])
            assert type(ei.value).__name__ in ("KeyError", "ValueError", "ZeroDivisionError")
            continue
        piv, stats = adapter.aggregate_instant(case["body"].encode(), eng)
        assert list(piv.index) == case["pivot"]["index"], case["name"]
        assert piv.index.name == "gpu_id" and piv.columns.name == "metric_name"
        assert list(piv.columns) == case["pivot"]["columns"], case["name"]
        assert {c: str(t) for c, t in piv.dtypes.items()} == case["pivot"]["dtypes"]
        for gid, rowd in case["pivot"]["cells"].items():
            for c, v in rowd.items():
                cell = piv.loc[gid, c]
                if isinstance(cell, float):
                    assert zero_eq(fbits(cell), v), (case["name"], gid, c)
                else:
                    assert cell == v
        for k in ("mean", "max", "min"):
            assert list(stats[k].index) == list(case["stats"][k]), case["name"]
            for c, v in case["stats"][k].items():
                assert zero_eq(fbits(stats[k][c]), v), (case["name"], k, c, stats[k][c])
        avg = adapter.selected_averages(piv, case["selected"], eng)
        for c, v in case["averages"].items():
            assert zero_eq(fbits(avg[c]), v), (case["name"], "avg", c)


def test_fetch_gpu_metrics_contract(eng, golden, monkeypatch):
    """Same two GETs, same return contract, every failure -> (None, None) + message."""
    import sys
    case = golden["cases"][0]
    bodies = {"kube": json.dumps({"status": "success", "data": {"resultType": "vector", "result": [
        {"metric": {"host_ip": "10.0.0.1"}, "value": [1, "1"]}]}}), "gpu": case["body"]}
    calls = []

    class Resp:
        def __init__(self, text): self.text = text
        def raise_for_status(self): pass

    def fake_get(url=None, params=None, **kw):
        calls.append(params["query"])
        return Resp(bodies["kube"] if params["query"].startswith("kube_pod_info") else bodies["gpu"])

    monkeypatch.setitem(sys.modules, "requests", types.SimpleNamespace(get=fake_get))
    monkeypatch.setattr(adapter._local, "engine", eng, raising=False)
    piv, stats = adapter.fetch_gpu_metrics()
    assert piv is not None and list(piv.index) == case["pivot"]["index"]
    assert len(calls) == 2 and 'instance=~"10.0.0.1:.+"' in calls[1] and "amd_gpu_total_vram" in calls[1]
    assert piv.loc["0", "card_model"] == case["pivot"]["cells"]["0"]["card_model"]
    assert set(stats) == {"mean", "max", "min"}
    import pandas as pd
    assert pd.DataFrame(stats).round(2).shape == (6, 3)                        # app.py:480
    bodies["gpu"] = [c for c in golden["cases"] if c["name"] == "total_vram_zero"][0]["body"]
    assert adapter.fetch_gpu_metrics() == (None, None)
    assert adapter.last_error == "Error fetching GPU metrics: float division by zero"


@pytest.mark.parametrize("tokenizer", ["host", "device", "auto"])
def test_aggregate_range_tidy_frame(eng, tokenizer):
    from oracle import prom_oracle as O
    import numpy as np
    from tests.util import assert_bits_equal, assert_close
    d = synth.dense(1, 2, 6, 120, seed=9, nan_frac=0.02)
    body = synth.to_matrix_json(d["labels"], d["ts"], d["val"], d["off"])
    tidy, summary = adapter.aggregate_range(body, d["start_ms"], d["step_ms"], 120, 900_000, eng=eng, tokenizer=tokenizer)
    assert list(tidy.columns) == ["instance", "gpu_id", "metric_name", "ts", "value", "roll_min", "roll_max",
                                  "roll_mean", "rate"]
    assert len(tidy) == 12 * 120
    ref = O.aggregate_range(d["ts"], d["val"], d["off"], d["start_ms"], d["step_ms"], 120, 900_000)
    assert_bits_equal(tidy["value"].to_numpy(), ref["aligned"].ravel())
    assert_close(tidy["roll_mean"].to_numpy(), ref["roll_mean"].ravel(), scale=O.rolling_scale(ref["aligned"], 60).ravel())
    names = sorted({l["__name__"] for l in d["labels"]})
    assert list(summary["metric_name"]) == names
    for i, n in enumerate(names):
        rows = [k for k, l in enumerate(d["labels"]) if l["__name__"] == n]
        q = O.quantiles(ref["aligned"][rows], [0.5, 0.95, 0.99])
        assert_bits_equal(summary.loc[i, ["p50", "p95", "p99"]].to_numpy(dtype=np.float64), np.array(q["value"]))
        assert summary.loc[i, "count"] == q["n"]


def test_aggregate_range_arrow_table_equals_the_frame(eng):
    import numpy as np
    d = synth.dense(2, 2, 6, 90, seed=11, nan_frac=0.02)
    body = synth.to_matrix_json(d["labels"], d["ts"], d["val"], d["off"])
    tidy, summary = adapter.aggregate_range(body, d["start_ms"], d["step_ms"], 90, 900_000, eng=eng)
    at, asum = adapter.aggregate_range(body, d["start_ms"], d["step_ms"], 90, 900_000, eng=eng, as_frame="arrow")
    assert at.column_names == list(tidy.columns) and at.num_rows == len(tidy)
    import pyarrow as pa
    assert all(pa.types.is_dictionary(at.schema.field(c).type) for c in ("instance", "gpu_id", "metric_name"))
    back = at.to_pandas()
    for c in ("instance", "gpu_id", "metric_name"):
        assert back[c].astype(str).tolist() == tidy[c].astype(str).tolist()
    for c in ("ts", "value", "roll_min", "roll_max", "roll_mean", "rate"):
        np.testing.assert_array_equal(back[c].to_numpy(), tidy[c].to_numpy())
    sb = asum.to_pandas()
    assert list(sb.columns) == list(summary.columns)
    for c in summary.columns:
        if c == "metric_name":
            assert sb[c].tolist() == summary[c].tolist()
        else:
            np.testing.assert_array_equal(sb[c].to_numpy(), summary[c].to_numpy())


def test_four_sessions_refresh_concurrently(golden):
    """Streamlit runs every browser session on its own script thread and each loops on
    fetch_gpu_metrics (app.py:326-331): four threads call the drop-in at the same time on different
    bodies, each through the engine of ITS thread (adapter.engine()), many times over; every
    result must equal what the unmodified reference returned for that body."""
    import threading
    cases = [c for c in golden["cases"] if "error" not in c][:4]
    errors, engines = [], set()

    def session(case):
        try:
            engines.add(id(adapter.engine()))
            for _ in range(25):
                piv, stats = adapter.aggregate_instant(case["body"].encode())
                assert list(piv.index) == case["pivot"]["index"]
                for gid, rowd in case["pivot"]["cells"].items():
                    for c, v in rowd.items():
                        cell = piv.loc[gid, c]
                        assert zero_eq(fbits(cell), v) if isinstance(cell, float) else cell == v, (case["name"], gid, c)
                for k in ("mean", "max", "min"):
                    for c, v in case["stats"][k].items():
                        assert zero_eq(fbits(stats[k][c]), v), (case["name"], k, c)
                avg = adapter.selected_averages(piv, case["selected"])
                for c, v in case["averages"].items():
                    assert zero_eq(fbits(avg[c]), v), (case["name"], "avg", c)
        except Exception as e:  # noqa: BLE001
            errors.append(f"{case['name']}: {type(e).__name__}: {e}")

    threads = [threading.Thread(target=session, args=(c,)) for c in cases]
    for th in threads:
        th.start()
    for th in threads:
        th.join()
    assert not errors, errors
    assert len(engines) == len(cases), "every session thread must have its own engine (pmx_ctx)"


def test_aggregate_range_group_keys_per_node_and_per_card(eng):
    """SURVEY X6 on the range path: groupby(['instance','gpu_id','__name__']) and per node -- what
    app.py:335-345 does for the selected cards of an instant vector -- against pandas' groupby."""
    import numpy as np
    import pandas as pd
    from oracle import prom_oracle as O
    from tests.util import assert_bits_equal
    d = synth.dense(3, 4, 6, 150, seed=21, nan_frac=0.03)
    body = synth.to_matrix_json(d["labels"], d["ts"], d["val"], d["off"])
    al, _ = O.align(d["ts"], d["val"], d["off"], d["start_ms"], d["step_ms"], 150)
    for keys in (("instance", "gpu_id", "__name__"), ("instance", "__name__"), ("__name__",)):
        _, summary = adapter.aggregate_range(body, d["start_ms"], d["step_ms"], 150, 900_000, eng=eng, group_keys=keys)
        cols = {k: np.repeat([l[k] for l in d["labels"]], 150) for k in keys}
        g = pd.DataFrame({**cols, "value": al.ravel()}).groupby(list(keys), sort=True)["value"]
        agg = g.agg(["sum", "count", "min", "max", "mean"])
        q = g.quantile([0.5, 0.95, 0.99]).unstack()
        assert len(summary) == len(agg)
        for i, k in enumerate(keys):
            exp = [idx[i] if isinstance(idx, tuple) else idx for idx in agg.index]
            assert list(summary["metric_name" if k == "__name__" else k]) == exp
        assert np.array_equal(summary["count"].to_numpy(), agg["count"].to_numpy())
        assert np.array_equal(summary["min"].to_numpy(), agg["min"].to_numpy())
        assert np.array_equal(summary["max"].to_numpy(), agg["max"].to_numpy())
        np.testing.assert_allclose(summary["mean"].to_numpy(), agg["mean"].to_numpy(), rtol=1e-12)
        assert_bits_equal(summary[["p50", "p95", "p99"]].to_numpy(dtype=np.float64), q.to_numpy(), f"quantiles by {keys}")
