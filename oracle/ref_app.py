"""Run the UNMODIFIED reference ``app.py`` against a mocked Prometheus.  TEST INFRASTRUCTURE.

``streamlit`` and ``plotly`` are not installed (and cannot be): stub modules are inserted into
``sys.modules`` so that ``app.py:7-19`` imports and ``st.set_page_config`` runs, ``st.error``
(``app.py:226``) collects messages, and ``app.requests.get`` (``app.py:158``, ``:173``) is replaced
by a fake that serves canned bodies.  Nothing of the reference is copied: the file is loaded
by path (default ``/root/reference/app.py``; ``baseline/_ref/app.py`` when present on a GPU box).
Only the golden-vector generator, ``tests/`` and ``bench.py``'s CPU-baseline leg use this.
"""
from __future__ import annotations

import importlib.util
import json
import os
import sys
import types

_CANDIDATES = ("/root/reference/app.py",
               os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                            "baseline", "_ref", "app.py"))


def reference_path():
    for p in _CANDIDATES:
        if os.path.exists(p):
            return p
    return None


class _Anything:
    """Permissive dummy: any attribute access or call returns another dummy."""
    def __getattr__(self, name):
        return _Anything()

    def __call__(self, *a, **k):
        return _Anything()


class _Resp:
    def __init__(self, text, status=200):
        self.text, self.status = text, status

    def raise_for_status(self):
        if self.status >= 400:
            raise RuntimeError(f"{self.status} Server Error")


def load_reference(path=None):
    """Import the reference module under stubs; returns (module, errors_list)."""
    path = path or reference_path()
    if path is None:
        raise FileNotFoundError("reference app.py not available on this machine")
    errors = []
    st = types.ModuleType("streamlit")
    st.set_page_config = lambda **k: None
    st.error = lambda msg: errors.append(msg)
    st.__getattr__ = lambda name: _Anything()
    plotly = types.ModuleType("plotly")
    go = types.ModuleType("plotly.graph_objects")
    go.__getattr__ = lambda name: _Anything()
    plotly.graph_objects = go
    saved = {k: sys.modules.get(k) for k in ("streamlit", "plotly", "plotly.graph_objects")}
    sys.modules.update({"streamlit": st, "plotly": plotly, "plotly.graph_objects": go})
    try:
        spec = importlib.util.spec_from_file_location("_reference_app", path)
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
    finally:
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v
    return mod, errors


def pod_info_body(host_ip):
    """The kube_pod_info answer app.py:157-164 needs (only ``host_ip`` is read)."""
    return json.dumps({"status": "success", "data": {"resultType": "vector", "result": [
        {"metric": {"__name__": "kube_pod_info", "pod": "prometheus-0", "host_ip": host_ip},
         "value": [1700000000.0, "1"]}]}})


def run_fetch(mod, gpu_body, host_ip="10.0.0.1"):
    """Call the reference's fetch_gpu_metrics() with the two GETs mocked."""
    bodies = [pod_info_body(host_ip), gpu_body]
    calls = []

    def fake_get(url=None, params=None, **kw):
        calls.append(params)
        return _Resp(bodies[min(len(calls) - 1, 1)])

    mod.requests = types.SimpleNamespace(get=fake_get)
    return mod.fetch_gpu_metrics()
