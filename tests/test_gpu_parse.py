"""GPU: K1 value/timestamp decode through the C ABI, bit-exact with Python float()."""
import json
import os
import random
import struct

import numpy as np
import pytest

from k8s_rocm_metrics_dashboard_b200 import synth, tokenize
from oracle import prom_oracle as O
from tests.util import assert_bits_equal

pytestmark = pytest.mark.gpu


def _body(tokens, ts="1700000000"):
    vals = ",".join(f'[{ts},{json.dumps(t)}]' for t in tokens)
    return ('{"status":"success","data":{"resultType":"matrix","result":[{"metric":{"__name__":"m"},"values":['
            + vals + "]}]}}").encode()


def _parse_each(eng, tokens):
    """Parse tokens one body at a time where needed so that bad tokens do not mask good ones."""
    tok = tokenize(_body(tokens))
    t = eng.torch
    text = t.from_numpy(tok.flat).to(eng.device)
    val = t.empty(len(tokens), dtype=t.float64, device=eng.device)
    bad = t.empty(1, dtype=t.int64, device=eng.device)
    voff = t.from_numpy(tok.val_off.view(np.int32)).to(eng.device)
    import ctypes as C
    from k8s_rocm_metrics_dashboard_b200 import _abi
    _abi.check(eng.lib.pmx_parse_f64(eng.ctx, C.c_void_p(text.data_ptr()), text.numel(), C.c_void_p(voff.data_ptr()),
                                     len(tokens), C.c_void_p(val.data_ptr()), C.c_void_p(bad.data_ptr()),
                                     C.c_void_p(t.cuda.current_stream().cuda_stream)))
    return val.cpu().numpy(), int(bad.item())


def test_float_golden_bit_exact(eng, golden_dir):
    g = [(t, e) for t, e in json.load(open(os.path.join(golden_dir, "float_golden.json")))
         if all(ord(c) < 128 for c in t)]
    good = [(t, e) for t, e in g if e != "ValueError"]
    got, first_bad = _parse_each(eng, [t for t, _ in good])
    assert first_bad >= len(good)                              # no malformed token
    exp = np.array([int(e, 16) for _, e in good], dtype=np.uint64)
    bad = np.flatnonzero(got.view(np.uint64) != exp)
    assert len(bad) == 0, [good[i][0][:60] for i in bad[:5]]
    # every malformed token is flagged, each on its own
    for tok, e in g:
        if e == "ValueError":
            _, fb = _parse_each(eng, ["1.5", tok, "2.5"])
            assert fb == 1, repr(tok)


def test_engine_parse_raises_value_error_like_float(eng):
    tok = tokenize(_body(["1.5", "12abc", "7"]))
    with pytest.raises(ValueError, match="could not convert string to float: '12abc'"):
        eng.parse(tok)


def test_random_reprs_and_long_mantissas(eng):
    rnd = random.Random(11)
    toks = []
    for _ in range(20000):
        x = struct.unpack("<d", struct.pack("<Q", rnd.getrandbits(64)))[0]
        toks.append(repr(x) if x == x else "NaN")
    for _ in range(5000):
        nd = rnd.choice([17, 19, 20, 30, 100, 400, 800])
        toks.append("".join(rnd.choice("0123456789") for _ in range(nd)) + "e%d" % rnd.randint(-340 - nd, 300 - nd))
    got, fb = _parse_each(eng, toks)
    assert fb >= len(toks)
    exp = np.array([struct.unpack("<Q", struct.pack("<d", float(t)))[0] for t in toks], dtype=np.uint64)
    bad = np.flatnonzero(got.view(np.uint64) != exp)
    assert len(bad) == 0, [toks[i][:60] for i in bad[:5]]


def test_matrix_body_end_to_end_matches_json_loads(eng):
    d = synth.config("config1", seed=2)
    d["val"][5] = np.nan; d["val"][9] = np.inf; d["val"][11] = -np.inf
    body = synth.to_matrix_json(d["labels"], d["ts"], d["val"], d["off"])
    tok = tokenize(body)
    ts, val = eng.parse(tok)
    labels, ots, oval, ooff = O.tokenize_matrix(body)
    assert tok.n_samples == 46080 and np.array_equal(tok.series_off, ooff)
    assert np.array_equal(ts.cpu().numpy(), ots)
    assert_bits_equal(val.cpu().numpy(), oval, "values")
    assert np.array_equal(ts.cpu().numpy(), d["ts"])
    assert_bits_equal(val.cpu().numpy(), d["val"], "round trip")


def test_empty_and_bad_timestamp(eng):
    tok = tokenize(b'{"status":"success","data":{"resultType":"matrix","result":[]}}')
    ts, val = eng.parse(tok)
    assert ts.numel() == 0 and val.numel() == 0
    tok = tokenize(_body(["1"], ts="1.23456"))
    with pytest.raises(ValueError, match="timestamp"):
        eng.parse(tok)


def test_pinned_tensor_body_equals_bytes_body(eng):
    """tokenize_device takes the body as a pinned uint8 tensor too (full-rate H2D): same tokens, same labels,
    and a malformed value is reported as for a bytes body."""
    import torch
    from k8s_rocm_metrics_dashboard_b200 import synth
    d = synth.dense(1, 2, 3, 50, seed=3)
    body = synth.to_matrix_json(d["labels"], d["ts"], d["val"], d["off"])
    pinned = torch.frombuffer(bytearray(body), dtype=torch.uint8).pin_memory()
    a, b = eng.tokenize_device(body), eng.tokenize_device(pinned)
    assert np.array_equal(a.series_off, b.series_off) and a.metric_spans == b.metric_spans and a.result_type == b.result_type
    assert list(a.labels()) == list(b.labels())
    ta, va = eng.parse(a); tb, vb = eng.parse(b)
    assert torch.equal(ta, tb) and torch.equal(va.view(torch.int64), vb.view(torch.int64))
    bad = bytearray(body)
    k = bad.find(b'"', bad.find(b'"values":[[') + 11) + 1
    bad[k] = ord("x")
    with pytest.raises(ValueError, match="could not convert string to float"):
        eng.parse(eng.tokenize_device(torch.frombuffer(bad, dtype=torch.uint8).pin_memory()))


def test_parse_async_gives_parse_s_numbers_and_errors(eng):
    """parse_async (decode kernel in flight while the host interns the labels): the same numbers as
    parse, the same ValueError at finish(), and aggregate_range over device tokens still refuses a
    malformed body."""
    import torch
    from k8s_rocm_metrics_dashboard_b200 import adapter, synth
    d = synth.dense(1, 2, 3, 50, seed=4)
    body = synth.to_matrix_json(d["labels"], d["ts"], d["val"], d["off"])
    tok = eng.tokenize_device(body)
    t0, v0 = eng.parse(tok)
    t1, v1, done = eng.parse_async(tok)
    labels = tok.labels()
    done()
    assert torch.equal(t0, t1) and torch.equal(v0.view(torch.int64), v1.view(torch.int64)) and len(labels) == 6
    _, v2, done = eng.parse_async(tok, want_ts=False)
    done()
    assert torch.equal(v0.view(torch.int64), v2.view(torch.int64))
    with pytest.raises(TypeError):
        from k8s_rocm_metrics_dashboard_b200 import tokenize
        eng.parse_async(tokenize(body))
    bad = bytearray(body)
    k = bad.find(b'"', bad.find(b'"values":[[') + 11) + 1
    bad[k] = ord("x")
    _, _, done = eng.parse_async(eng.tokenize_device(bytes(bad)))
    with pytest.raises(ValueError, match="could not convert string to float"):
        done()
    step = int(d["step_ms"])
    kw = dict(start_ms=int(d["start_ms"]), step_ms=step, n_steps=int(d["n_steps"]), window_ms=5 * step, as_frame=False, eng=eng)
    tidy_d, summ_d = adapter.aggregate_range(body, tokenizer="device", **kw)
    tidy_h, summ_h = adapter.aggregate_range(body, tokenizer="host", **kw)
    for k2 in ("value", "roll_mean", "rate"):
        np.testing.assert_array_equal(np.asarray(tidy_d[k2]).view(np.int64), np.asarray(tidy_h[k2]).view(np.int64))
    np.testing.assert_array_equal(summ_d["count"], summ_h["count"])
    with pytest.raises(ValueError, match="could not convert string to float"):
        adapter.aggregate_range(bytes(bad), tokenizer="device", **kw)
