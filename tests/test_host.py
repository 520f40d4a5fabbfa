"""CPU: the C-ABI library loads and exports what include/pmx.h declares; host-side logic
(tokenizer, group index) and the host build of the parser state machine."""
import ctypes as C
import json
import os
import re
import struct
import subprocess

import numpy as np
import pytest

from k8s_rocm_metrics_dashboard_b200 import _abi, synth, tokenize, group_index
from oracle import prom_oracle as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "pmx.h")).read()
    declared = set(re.findall(r"\b(pmx_[a-z0-9_]+)\s*\(", hdr))
    assert len(declared) >= 20
    lib = _abi.lib()
    for name in sorted(declared):
        assert hasattr(lib, name), f"{name} declared in pmx.h but not exported by libpmx.so"
    assert declared == set(_abi.SIGNATURES), declared ^ set(_abi.SIGNATURES)
    assert lib.pmx_version() == 100


def test_no_gpu_fails_loudly():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from k8s_rocm_metrics_dashboard_b200 import Engine
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        Engine()
    p = C.c_void_p()
    assert _abi.lib().pmx_ctx_create(0, C.byref(p)) == _abi.E_CUDA
    assert b"no CPU path" in _abi.lib().pmx_last_error()


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "k8s_rocm_metrics_dashboard_b200")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dp, f)).read()
                assert "from oracle" not in src and "import oracle" not in src and "prom_oracle" not in src, f


def _tok_text(tok, offs, term):
    out = []
    for o in offs:
        e = int(o)
        while tok.flat[e] != term:
            e += 1
        out.append(bytes(tok.flat[int(o):e]).decode())
    return out


def test_tokenizer_matches_json_loads_matrix():
    d = synth.dense(2, 2, 3, 20, seed=4, nan_frac=0.1)
    d["val"][3] = np.inf; d["val"][7] = -np.inf
    body = synth.to_matrix_json(d["labels"], d["ts"], d["val"], d["off"])
    tok = tokenize(body)
    ts_tok, val_tok, off = O.raw_tokens(body)
    assert tok.result_type == "matrix"
    assert tok.series_off.tolist() == off
    assert _tok_text(tok, tok.val_off, 0x22) == val_tok
    assert [t.strip() for t in _tok_text(tok, tok.ts_off, 0x2C)] == ts_tok
    assert tok.labels() == d["labels"]


def test_tokenizer_vector_whitespace_escapes_and_extra_keys():
    body = ('{ "status" : "success", "warnings":["x",{"a":[1,2,{"b":"]}"}]}], "data" : {"result" : [\n'
            ' {"metric":{"__name__":"m","gpu_id":"0","q":"a\\"b}"}, "value" : [ 1700000000.5 , "1.5" ] , "extra": null},\n'
            ' {"value":[1700000001,"\\u0031\\t2"],"metric":{}},'
            ' {"metric":{"k":"v"},"values":[]}, {"metric":{"k":"w"},"histograms":[[1,{"count":"3"}]]}\n'
            '], "resultType":"vector" } }').encode()
    tok = tokenize(body)
    assert tok.result_type == "vector" and tok.n_series == 4 and tok.n_samples == 2
    assert tok.series_off.tolist() == [0, 1, 2, 2, 2]
    assert _tok_text(tok, tok.val_off, 0x22) == ["1.5", "1\t2"]
    assert tok.val_off[1] >= len(body)                      # decoded copy lives in the extra region
    assert [t.strip() for t in _tok_text(tok, tok.ts_off, 0x2C)] == ["1700000000.5", "1700000001"]
    assert tok.labels()[0] == {"__name__": "m", "gpu_id": "0", "q": 'a"b}'}


@pytest.mark.parametrize("body, exc", [
    (b'{"status":"error","errorType":"bad_data","error":"x"}', ValueError),
    (b'{"status":"success","data":{"resultType":"matrix"}}', KeyError),
    (b'{"status":"success","data":{"resultType":"matrix","result":[{"metric":{},"values":[[1,"2"]}]}}', ValueError),
    (b'{"status":"success","data":{"resultType":"scalar","result":[]}}', ValueError),
    (b'[1,2]', ValueError), (b'', ValueError),
    (b'{"status":"success","data":{"resultType":"matrix","result":[]}} trailing', ValueError),
])
def test_tokenizer_rejects(body, exc):
    with pytest.raises(exc):
        tokenize(body)


def test_tokenizer_empty_result():
    tok = tokenize(b'{"status":"success","data":{"resultType":"vector","result":[]}}')
    assert tok.n_series == 0 and tok.n_samples == 0 and tok.series_off.tolist() == [0]


def test_group_index_stable_csr():
    gid = np.array([2, 0, -1, 2, 1, 0, 2], dtype=np.int32)
    perm, goff = group_index(gid, 4)
    assert goff.tolist() == [0, 2, 3, 6, 6]
    assert perm.tolist() == [1, 5, 4, 0, 3, 6]


@pytest.fixture(scope="module")
def host_lib():
    so = os.path.join(ROOT, "tests", "_build", "strtod_host.so")
    src = os.path.join(ROOT, "tests", "host", "strtod_host.cpp")
    os.makedirs(os.path.dirname(so), exist_ok=True)
    subprocess.check_call(["g++", "-O2", "-shared", "-fPIC", "-o", so, src])
    return C.CDLL(so)


def _host_parse(lib, toks):
    enc = [t.encode("utf-8") for t in toks]
    off = np.zeros(len(enc) + 1, dtype=np.int64)
    off[1:] = np.cumsum([len(e) for e in enc])
    buf = b"".join(enc) + b"\0"
    bits = np.zeros(len(enc), dtype=np.uint64); st = np.zeros(len(enc), dtype=np.int32)
    lib.host_parse_many(buf, C.c_void_p(off.ctypes.data), C.c_int64(len(enc)), C.c_void_p(bits.ctypes.data),
                        C.c_void_p(st.ctypes.data))
    return bits, st


def test_parser_state_machine_on_host_matches_float_golden(host_lib, golden_dir):
    """Same strtod_core.h the K1 kernel compiles, built for the host: every golden token
    (ASCII ones; CPython also accepts non-ASCII digits, which Prometheus never emits)."""
    g = [(t, e) for t, e in json.load(open(os.path.join(golden_dir, "float_golden.json")))
         if all(ord(c) < 128 for c in t)]
    bits, st = _host_parse(host_lib, [t for t, _ in g])
    for (tok, exp), b, s in zip(g, bits, st):
        got = "ValueError" if s else "%016x" % b
        assert got == exp, repr(tok[:80])


def test_parser_state_machine_fuzz_vs_python_float(host_lib):
    import random
    rnd = random.Random(7)
    toks = []
    for _ in range(60000):
        k = rnd.random()
        if k < 0.35:
            x = struct.unpack("<d", struct.pack("<Q", rnd.getrandbits(64)))[0]
            toks.append(repr(x) if x == x else "nan")
        elif k < 0.7:
            nd = rnd.randint(1, 45)
            s = "".join(rnd.choice("0123456789") for _ in range(nd))
            p = rnd.randint(0, nd)
            s = s[:p] + "." + s[p:] if rnd.random() < 0.7 else s
            if rnd.random() < 0.7:
                s += "e%d" % rnd.randint(-350, 320)
            toks.append(rnd.choice(["", "-", "+"]) + s)
        else:
            x = struct.unpack("<d", struct.pack("<Q", rnd.getrandbits(63)))[0]
            if x == x and x != float("inf"):
                toks.append("%.*e" % (rnd.randint(15, 40), x))
    bits, st = _host_parse(host_lib, toks)
    exp = np.array([struct.unpack("<Q", struct.pack("<d", float(t)))[0] for t in toks], dtype=np.uint64)
    assert not st.any()
    bad = np.flatnonzero(bits != exp)
    assert len(bad) == 0, [toks[i] for i in bad[:5]]


def test_timestamp_parser_host(host_lib):
    for tok, exp in [("1700000000", 1700000000000), ("1700000000.5", 1700000000500), ("1700000000.123", 1700000000123),
                     ("0", 0), ("-1.5", -1500), (" 12.250 ", 12250), ("1.0000", 1000)]:
        out = C.c_int64()
        b = tok.encode()
        assert host_lib.host_parse_ts_ms(b, C.c_int64(len(b)), C.byref(out)) == 0 and out.value == exp, tok
        assert out.value == O.parse_ts_ms(tok.strip())
    for tok in ["", ".", "1.", "1.2345", "1e9", "abc", "1.5x", "99999999999999999999"]:
        out = C.c_int64()
        b = tok.encode()
        assert host_lib.host_parse_ts_ms(b, C.c_int64(len(b)), C.byref(out)) != 0, tok


def test_tokenizer_multithreaded_equals_single_walk():
    """pmx_tokenize_mt (series-boundary chunks, count + fill passes) against the single walk on a
    body above its 1 MiB threshold, with escaped value tokens and a warnings tail."""
    d = synth.dense(4, 8, 6, 400, seed=8, nan_frac=0.01)
    body = synth.to_matrix_json(d["labels"], d["ts"], d["val"], d["off"])
    assert len(body) > (1 << 20)
    # a few escaped value tokens (decoded copies go to the extra region) and a tail after "data"
    body = body.replace(b'"NaN"', b'"N\\u0061N"', 7)
    body = body[:-1] + b',"warnings":["x]}","y"],"infos":[]}'
    one = tokenize(body, threads=1)
    for th in (2, 5, 16):
        mt = tokenize(body, threads=th)
        assert mt.result_type == one.result_type and mt.n_series == one.n_series and mt.n_samples == one.n_samples
        assert np.array_equal(mt.series_off, one.series_off)
        assert np.array_equal(mt.ts_off, one.ts_off)
        assert mt.metric_spans == one.metric_spans
        assert _tok_text(mt, mt.val_off, 0x22) == _tok_text(one, one.val_off, 0x22)
    assert sum(1 for o in one.val_off if o >= len(body)) == 7


def test_tokenizer_multithreaded_reports_errors_like_single_walk():
    d = synth.dense(4, 8, 6, 400, seed=9)
    body = synth.to_matrix_json(d["labels"], d["ts"], d["val"], d["off"])
    broken = body[:700_000] + b"]" + body[700_000:]            # stray bracket deep inside one chunk
    for th in (1, 8):
        with pytest.raises(ValueError):
            tokenize(broken, threads=th)
    not_canonical = b' ' + body                                  # leading space: plain walk serves it
    assert tokenize(not_canonical, threads=8).n_samples == 192 * 400


def test_fast_window_decode_on_host_matches_python_float(host_lib, golden_dir):
    """csrc/strtod_fast.h (what K1 runs on [-]digits[.digits] tokens of at most 32 bytes): every token
    it accepts must give float()'s bits / the exact milliseconds; everything else must be handed on."""
    import random
    from decimal import Decimal
    lib = host_lib
    lib.host_fast_f64.argtypes = [C.c_char_p, C.c_int64, C.POINTER(C.c_uint64)]
    lib.host_fast_ts_ms.argtypes = [C.c_char_p, C.c_int64, C.POINTER(C.c_int64)]
    rnd = random.Random(11)
    toks = [t for t, _ in json.load(open(os.path.join(golden_dir, "float_golden.json"))) if all(ord(c) < 128 for c in t)]
    for _ in range(60000):
        k = rnd.random()
        if k < 0.3:
            toks.append(repr(round(rnd.uniform(0, 1000), 3)))
        elif k < 0.5:
            toks.append(repr(rnd.uniform(0, 1e9)))
        elif k < 0.6:
            toks.append(str(rnd.randrange(0, 10 ** rnd.randrange(1, 21))))
        elif k < 0.75:
            x = struct.unpack("<d", struct.pack("<Q", rnd.getrandbits(64)))[0]
            toks.append(repr(x))
        else:
            nd = rnd.randrange(1, 22); d = "".join(rnd.choice("0123456789") for _ in range(nd)); p = rnd.randrange(0, nd + 1)
            toks.append(("-" if rnd.random() < 0.3 else "") + d[:p] + ("." + d[p:] if p < nd else ""))
    decoded = 0
    for tk in toks:
        raw = tk.encode()
        b = C.c_uint64()
        if lib.host_fast_f64(raw, len(raw), C.byref(b)) == 0:
            decoded += 1
            assert b.value == struct.unpack("<Q", struct.pack("<d", float(tk)))[0], repr(tk)
    assert decoded > 40000                                    # the common shapes do take this path
    for _ in range(20000):
        s = rnd.randrange(0, 4 * 10 ** 9); f = rnd.randrange(0, 1000); nf = rnd.randrange(0, 4)
        tk = f"{s}" if nf == 0 else f"{s}.{f:03d}"[:len(str(s)) + 1 + nf]
        tk = ("-" if rnd.random() < 0.1 else "") + tk
        o = C.c_int64()
        if lib.host_fast_ts_ms(tk.encode(), len(tk), C.byref(o)) == 0:
            assert o.value == int(Decimal(tk) * 1000), tk
    for tk in ("1.2345", "12345678901234567", "1e3", "", "-", "1.", ".5", " 1"):     # not this path's business
        assert lib.host_fast_ts_ms(tk.encode(), len(tk), C.byref(C.c_int64())) == 1 or tk == "12345678901234567"
    # the scanning form (what the kernel runs now): window = token + terminator + whatever follows
    lib.host_scan_f64.argtypes = [C.c_char_p, C.c_int64, C.c_uint8, C.POINTER(C.c_uint64)]
    lib.host_scan_ts_ms.argtypes = [C.c_char_p, C.c_int64, C.c_uint8, C.POINTER(C.c_int64)]
    decoded = short = 0
    for tk in toks:
        raw = tk.encode()
        b = C.c_uint64()
        if lib.host_scan_f64(raw, len(raw), ord('"'), C.byref(b)) == 0:
            decoded += 1
            short += len(raw) <= 8
            assert b.value == struct.unpack("<Q", struct.pack("<d", float(tk)))[0], repr(tk)
        elif len(raw) < 24 and raw.lstrip(b"-").replace(b".", b"", 1).isdigit() and raw[-1:] != b"." and raw.lstrip(b"-")[:1] != b"." \
                and len(raw.lstrip(b"-").replace(b".", b"")) <= 19 and raw.count(b"-") <= 1:
            # a plain decimal it may only hand on for the Eisel-Lemire "uncertain" cases
            assert len(raw.lstrip(b"-").replace(b".", b"")) > 15, repr(tk)
    assert decoded > 40000 and short > 5000
    for tk in ("1.2345", "0.001", "7", "-3.5", "99999999", "1234567.8", "-0", "0.0", "12.34567", "100000000"):
        b = C.c_uint64()
        assert lib.host_scan_f64(tk.encode(), len(tk), ord('"'), C.byref(b)) == 0, tk
        assert b.value == struct.unpack("<Q", struct.pack("<d", float(tk)))[0], tk
    for tk in ("1e3", "", "-", "1.", ".5", " 1", "NaN", "+Inf", "1.2.3", "1_0", "--1", "1-"):
        assert lib.host_scan_f64(tk.encode(), len(tk), ord('"'), C.byref(C.c_uint64())) == 1, tk
    for _ in range(20000):
        s_ = rnd.randrange(0, 4 * 10 ** 9); f = rnd.randrange(0, 1000); nf = rnd.randrange(0, 4)
        tk = f"{s_}" if nf == 0 else f"{s_}.{f:03d}"[:len(str(s_)) + 1 + nf]
        tk = ("-" if rnd.random() < 0.1 else "") + tk
        o = C.c_int64()
        assert lib.host_scan_ts_ms(tk.encode(), len(tk), ord(","), C.byref(o)) == 0, tk
        assert o.value == int(Decimal(tk) * 1000), tk
    for tk in ("1.2345", "1e3", "", "-", "1.", ".5", " 1"):
        assert lib.host_scan_ts_ms(tk.encode(), len(tk), ord(","), C.byref(C.c_int64())) == 1, tk


def test_interned_label_table_equals_json_loads():
    """pmx_labels_intern (X1's label table): the same dicts json.loads builds (app.py:178), escapes and
    non-ASCII included; a column is one integer array; group ids need no per-series dict."""
    from k8s_rocm_metrics_dashboard_b200.adapter import _group_ids
    d = synth.dense(3, 4, 6, 3, seed=1)
    d["labels"][2]["weird"] = 'a"b\\cé中\U0001F600\n\t/'
    del d["labels"][5]["gpu_id"]
    d["labels"][7][""] = ""
    body = synth.to_matrix_json(d["labels"], d["ts"], d["val"], d["off"])
    table = tokenize(body).labels()
    ref = [s["metric"] for s in json.loads(body)["data"]["result"]]
    assert len(table) == len(ref) and list(table) == ref and table == ref
    assert table[-1] == ref[-1] and table[1:3] == ref[1:3]
    assert len(table.strings) == len(set(table.strings))                      # every string once
    assert table.values("gpu_id", missing="?") == [l.get("gpu_id", "?") for l in ref]
    assert (table.column("no_such_label") == -1).all()
    for keys in (("__name__",), ("instance", "gpu_id", "__name__"), ("instance", "__name__")):
        n1, g1 = _group_ids(table, keys)
        n2, g2 = _group_ids(ref, keys)
        assert n1 == n2 and np.array_equal(g1, g2), keys


def test_byte_class_masks_of_the_device_tokenizer_are_exact(host_lib):
    """csrc/byte_masks.h builds the masks of a 16-byte segment four bytes at a time.  Against a
    byte-by-byte restatement: random bytes, bytes drawn from the neighbours of every class byte
    (one code point away, bit 5 / bit 7 flipped, 0x00, 0x01, 0x7f, 0x80, 0xff), and every byte value
    at every position of a word beside a matching byte (where a carry or borrow would leak)."""
    lib = host_lib
    lib.host_segment_masks_many.argtypes = [C.c_void_p, C.c_int64, C.c_void_p]
    rng = np.random.default_rng(3)
    cls = [0x22, 0x5b, 0x7b, 0x2c, 0x5c]
    near = sorted({(c + d) & 0xff for c in cls for d in (-1, 0, 1)} | {c ^ 0x20 for c in cls} | {c ^ 0x80 for c in cls}
                  | {0x00, 0x01, 0x7f, 0x80, 0xff, 0x30, 0x2d})
    segs = [rng.integers(0, 256, size=(20000, 16), dtype=np.uint8),
            np.array(near, dtype=np.uint8)[rng.integers(0, len(near), size=(60000, 16))]]
    pairs = np.zeros((len(cls) * 256 * 4 * 2, 16), dtype=np.uint8)
    i = 0
    for c in cls:
        for v in range(256):
            for pos in range(4):
                for side in (0, 1):                                 # the matching byte below / above the probe
                    row = np.full(16, 0x41, dtype=np.uint8)
                    row[4 + pos] = v
                    other = 4 + (pos + (1 if side else -1)) % 4
                    row[other] = c
                    pairs[i] = row; i += 1
    segs = np.ascontiguousarray(np.concatenate(segs + [pairs]))
    out = np.zeros((len(segs), 22), dtype=np.uint32)
    lib.host_segment_masks_many(segs.ctypes.data, len(segs), out.ctypes.data)
    bit = (1 << np.arange(16, dtype=np.uint32))
    mask = lambda cond: (cond.astype(np.uint32) * bit).sum(axis=1).astype(np.uint32)
    np.testing.assert_array_equal(out[:, 0], mask(segs == 0x22))
    np.testing.assert_array_equal(out[:, 1], mask((segs == 0x5b) | (segs == 0x7b)))
    np.testing.assert_array_equal(out[:, 2], mask((segs & 0x20) != 0))
    np.testing.assert_array_equal(out[:, 3], mask(segs == 0x2c))
    # inside a string at position k (segment begins outside): odd number of quotes before k; bit 16: parity of all
    q = (segs == 0x22).astype(np.uint32)
    before = np.concatenate([np.zeros((len(segs), 1), dtype=np.uint32), np.cumsum(q, axis=1)], axis=1) & 1
    want_inside = (before * (1 << np.arange(17, dtype=np.uint32))).sum(axis=1).astype(np.uint32)
    np.testing.assert_array_equal(out[:, 4], want_inside)
    # maybe4: never misses a backslash (false positives are allowed)
    has_bs = mask(np.repeat((segs.reshape(-1, 4, 4) == 0x5c).any(axis=2), 4, axis=1))
    assert np.all((out[:, 5] & has_bs) == has_bs)
    np.testing.assert_array_equal(out[:, 6:], segs.astype(np.uint32))


def test_group_ids_on_the_interned_table_equal_the_dict_walk():
    """adapter._group_ids: the integer path (one 1-D unique over rows folded into one key) must give
    the names and ids of the plain walk over label dicts, for one and several keys, missing labels
    and an empty-string value."""
    import random
    from k8s_rocm_metrics_dashboard_b200 import adapter
    rnd = random.Random(4)
    labels = []
    for i in range(300):
        l = {"__name__": rnd.choice(["amd_gpu_power", "amd_gpu_edge_temperature", "amd_gpu_gfx_activity", ""]),
             "instance": f"node{rnd.randrange(7)}:9400", "gpu_id": str(rnd.randrange(8))}
        if rnd.random() < 0.2:
            del l["gpu_id"]
        if rnd.random() < 0.1:
            del l["__name__"]
        labels.append(l)
    ts = np.arange(len(labels), dtype=np.int64) * 1000 + 1_700_000_000_000
    body = synth.to_matrix_json(labels, ts, np.ones(len(labels)), np.arange(len(labels) + 1, dtype=np.int64))
    tok = tokenize(body)
    table = tok.labels()
    assert table == labels and tok.span_arrays is not None
    assert tok.metric_spans == list(zip(tok.span_arrays[0].tolist(), tok.span_arrays[1].tolist()))
    # the device tokenizer's result reads its labels the same way (body + the two span arrays)
    from k8s_rocm_metrics_dashboard_b200.engine import DeviceTokens
    dt = DeviceTokens(body, None, None, None, tok.series_off, tok.metric_spans, "matrix",
                      (tok.span_arrays[0].view(np.int32).view(np.uint32), tok.span_arrays[1].copy()))
    assert dt.labels() == labels and dt.n_series == len(labels)
    assert DeviceTokens(np.frombuffer(body, dtype=np.uint8), None, None, None, tok.series_off, tok.metric_spans,
                        "matrix").labels() == labels                  # a uint8 view of the body, spans as a list
    for keys in (("__name__",), ("instance", "gpu_id"), ("__name__", "instance", "gpu_id"), ("absent",), ()):
        n1, g1 = adapter._group_ids(table, keys)
        n2, g2 = adapter._group_ids(labels, keys)
        assert n1 == n2, keys
        np.testing.assert_array_equal(g1, g2)
