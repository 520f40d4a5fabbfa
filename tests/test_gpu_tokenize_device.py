"""pmx_tokenize_device against the host tokenizer (pmx_tokenize) on the same bodies: identical
offsets, CSR and label spans; then the parsed numbers against the oracle.  SURVEY 8(f) rank 1."""
import numpy as np
import pytest

from k8s_rocm_metrics_dashboard_b200 import synth, tokenize
from tests.util import assert_bits_equal

pytestmark = pytest.mark.gpu


def _same_as_host(eng, body):
    h = tokenize(body, threads=1)
    d = eng.tokenize_device(body)
    assert d.n_samples == h.n_samples and d.n_series == h.n_series
    assert d.result_type == h.result_type
    np.testing.assert_array_equal(d.series_off, h.series_off)
    np.testing.assert_array_equal(d.d_ts_off.cpu().numpy().view(np.uint32), h.ts_off)
    np.testing.assert_array_equal(d.d_val_off.cpu().numpy().view(np.uint32), h.val_off)
    assert d.metric_spans == h.metric_spans
    assert d.labels() == h.labels()
    return h, d


@pytest.mark.parametrize("cfg,scale", [("config1", 1 / 8), ("config2", 1 / 64)])
def test_device_tokens_equal_host_tokens_and_parse_to_the_same_numbers(eng, cfg, scale):
    d = synth.config(cfg, seed=3, scale=scale)
    body = synth.to_matrix_json(d["labels"], d["ts"], d["val"], d["off"])
    assert len(body) > 3 * 65536                       # several chunks: states and offsets cross chunk borders
    h, dv = _same_as_host(eng, body)
    ts, val = eng.parse(dv)
    assert_bits_equal(val.cpu().numpy(), d["val"], "values")
    np.testing.assert_array_equal(ts.cpu().numpy(), d["ts"])


def test_strings_that_look_like_structure_and_escapes(eng):
    """Label values holding '[1', '{"metric":{', escaped quotes and backslashes must not be taken
    for samples or series; an escaped backslash before the closing quote must end the string."""
    labels = [{"__name__": "m", "a": '[1700000000,"7"]', "b": 'x{"metric":{"y":"z"}}'},
              {"__name__": "m", "q": 'say "hi" [2,"3"]', "path": "C:\\dir\\"},
              {}]
    ts = np.array([1_700_000_000_000, 1_700_000_015_000, 1_700_000_000_123, 1500, 0], dtype=np.int64)
    val = np.array([1.5, np.nan, np.inf, 2.5e-5, 1e-300])
    off = np.array([0, 2, 2, 5], dtype=np.int64)       # the second series has no samples
    body = synth.to_matrix_json(labels, ts, val, off)
    h, dv = _same_as_host(eng, body)
    assert dv.labels() == labels
    t2, v2 = eng.parse(dv)
    assert_bits_equal(v2.cpu().numpy(), val, "values")
    np.testing.assert_array_equal(t2.cpu().numpy(), ts)


def test_vector_body_empty_result_and_tail_shorter_than_a_segment(eng):
    labels, values, body = synth.instant_snapshot(node=1, n_cards=8, seed=2)
    h, dv = _same_as_host(eng, body)
    _, v = eng.parse(dv, want_ts=False)
    assert_bits_equal(v.cpu().numpy(), np.array(values), "values")
    empty = b'{"status":"success","data":{"resultType":"matrix","result":[]}}'
    _same_as_host(eng, empty)
    for pad in range(1, 17):                               # every body length modulo 16
        l2 = [dict(labels[0], pad="p" * pad)]
        _same_as_host(eng, synth.to_vector_json(l2, synth.START_MS, values[:1]))


def test_bodies_outside_the_compact_encoding_are_refused_not_guessed(eng):
    good = b'{"status":"success","data":{"resultType":"matrix","result":[{"metric":{"__name__":"m"},"values":[[1700000000,"1"],[1700000015.5,"2"]]}]}}'
    d = eng.tokenize_device(good)
    assert d.n_samples == 2 and d.n_series == 1
    for bad in (good.replace(b'"success"', b'"error"'),
                good.replace(b'[1700000000', b'[ 1700000000'),           # white space after '['
                good.replace(b',"1"]', b',1]'),                           # value not a string
                b'{"status":"success","data":{"resultType":"scalar","result":[1700000000,"1"]}}'):
        with pytest.raises(ValueError):
            eng.tokenize_device(bad)
    assert tokenize(good.replace(b'[1700000000', b'[ 1700000000')).n_samples == 2   # the host tokenizer takes general JSON


def test_byte_classes_next_to_each_other_and_tokens_across_segments_and_chunks(eng):
    """The masks of the device tokenizer are built four bytes at a time: bytes one code point away
    from '"', '[', '{', ',' and the backslash must not be taken for them, wherever they fall in a word;
    a timestamp of 44 bytes (sub-millisecond zeros) takes the byte-wise search for its ','; and with a label of
    every length from 0 to 63 each structural byte meets each position of a segment, a warp's round
    (512 bytes) and a chunk (8 KiB)."""
    rng = np.random.default_rng(5)
    near = '"#![Z\\\\]{z|},+-\\\\"0:'
    series = []
    for n in range(64):
        junk = "".join(near[i] for i in rng.integers(0, len(near), size=n))
        lab = {"__name__": "m", "k": junk, "n": "x" * n}
        t0 = 1_700_000_000 + 15 * n
        samples = [(f"{t0}", "1"), (f"{t0 + 15}.5", "-2.5"), (f"{t0 + 30}.125{'0' * 30}", "NaN"),
                   (f"{t0 + 45}.250", f"{n}.125")] * (1 + n % 5)
        series.append((lab, samples))
    import json
    parts = []
    for lab, samples in series:
        vals = ",".join(f'[{t},"{v}"]' for t, v in samples)
        parts.append('{"metric":' + json.dumps(lab, separators=(",", ":")) + ',"values":[' + vals + "]}")
    body = ('{"status":"success","data":{"resultType":"matrix","result":[' + ",".join(parts) + "]}}").encode()
    assert len(body) > 3 * 8192
    h, d = _same_as_host(eng, body)
    assert d.n_series == 64 and d.n_samples == sum(len(s) for _, s in series)
    assert d.labels() == [lab for lab, _ in series]
    t_dev, v_dev = eng.parse(d)
    want_v = np.array([float(v) for _, s in series for _, v in s])
    assert_bits_equal(v_dev.cpu().numpy(), want_v, "values")
    want_t = np.array([int(round(float(t) * 1000)) for _, s in series for t, _ in s], dtype=np.int64)
    np.testing.assert_array_equal(t_dev.cpu().numpy(), want_t)
