"""Generate the committed golden fixtures.  Run in the build container (needs /root/reference):

    python tests/golden/make_golden.py

* ``float_golden.json``   -- value tokens and the bits Python's own ``float()`` returns
  (``app.py:188`` is ``float(metric["value"][1])``), or "ValueError".
* ``instant_golden.json`` -- bodies served to the UNMODIFIED reference ``fetch_gpu_metrics()``
  (``app.py:153-227``, loaded by ``oracle/ref_app.py`` under import stubs) and what it returned:
  the pivot frame cell by cell (float bits), the mean/max/min stats, the selected-card
  averages of ``app.py:335-345``, or the ``st.error`` text when it returned ``(None, None)``.
"""
import json
import os
import random
import struct
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

import numpy as np                                             # noqa: E402
import pandas as pd                                            # noqa: E402

from k8s_rocm_metrics_dashboard_b200 import synth              # noqa: E402
from oracle import ref_app                                     # noqa: E402


def bits(x):
    return "%016x" % struct.unpack("<Q", struct.pack("<d", float(x)))[0]


ADVERSARIAL = [
    "0", "-0", "0.0", "-0.0", "+0", "1", "-1", "1.0", "1.5", "100", "1e0", "1E0", "1e+0", "1e-0",
    "0.1", "0.2", "0.3", "123456789012345678", "1234567890123456789", "12345678901234567890",
    "9007199254740992", "9007199254740993", "9007199254740994", "9007199254740995",
    "4503599627370496.5", "4503599627370497.5", "9007199254740993.0000000000000000000001",
    "1e22", "1e23", "1e-22", "1e-23", "8.5e-5", "123456789e-5", "1.7976931348623157e308",
    "1.7976931348623158e308", "1.7976931348623159e308", "1.797693134862315807e308",
    "179769313486231580793728971405303415079934132710037826936173778980444968292764750946649017977587207096330286416692887910946555547851940402630657488671505820681908902000708383676273854845817711531764475730270069855571366959622842914819860834936475292719074168444365510704342711559699508093042880177904174497791",
    "179769313486231580793728971405303415079934132710037826936173778980444968292764750946649017977587207096330286416692887910946555547851940402630657488671505820681908902000708383676273854845817711531764475730270069855571366959622842914819860834936475292719074168444365510704342711559699508093042880177904174497792",
    "1e308", "1e309", "1e400", "-1e400", "1e-400", "2e308", "4.9e-324", "5e-324", "4e-324", "3e-324",
    "2.5e-324", "2.4703282292062327e-324", "2.4703282292062328e-324", "2.47032822920623272088284396434110686182e-324",
    "2.4703282292062327208828439643411068618252990130716238221279284125033775363510437593264991818081799618989828234772285886546332835517796989819938739800539093906315035659515570226392290858392449105184435931802849936536152500319370457678249219365623669863658480757001585769269903706311928279558551332927834338409351978015531246597263579574622766465272827220056374006485499977096599470454020828166226237857393450736339007967761930577506740176324673600968951340535537458516661134223766678604162159680461914467291840300530057530849048765391711386591646239524912623653881879636239373280423891018672348497668235089863388587925628302755995657524455507255189313690836254779186948667994968324049705821028513185451396213837722826145437693412532098591327667236328125e-324",
    "2.4703282292062327208828439643411068618252990130716238221279284125033775363510437593264991818081799618989828234772285886546332835517796989819938739800539093906315035659515570226392290858392449105184435931802849936536152500319370457678249219365623669863658480757001585769269903706311928279558551332927834338409351978015531246597263579574622766465272827220056374006485499977096599470454020828166226237857393450736339007967761930577506740176324673600968951340535537458516661134223766678604162159680461914467291840300530057530849048765391711386591646239524912623653881879636239373280423891018672348497668235089863388587925628302755995657524455507255189313690836254779186948667994968324049705821028513185451396213837722826145437693412532098591327667236328126e-324",
    "2.2250738585072014e-308", "2.2250738585072011e-308", "2.2250738585072009e-308", "2.225073858507201e-308",
    "2.2250738585072012e-308", "1.00000000000000011102230246251565404236316680908203125",
    "1.00000000000000011102230246251565404236316680908203124", "1.00000000000000011102230246251565404236316680908203126",
    "1.0000000000000002", "0.500000000000000166533453693773481063544750213623046875",
    "3.14159265358979323846264338327950288419716939937510582097494459230781640628620899862803482534211706798214808651",
    "0." + "0" * 400 + "1", "1" + "0" * 400, "1" + "0" * 308, "0" * 500 + "1", "0" * 500 + ".5",
    "1" + "0" * 400 + "e-400", "0." + "0" * 400 + "1e400", "123" + "0" * 30 + "e-30",
    "9" * 19, "9" * 20, "9" * 21, "1" + "9" * 18, "18446744073709551615", "18446744073709551616",
    "1e18446744073709551616", "1e-18446744073709551616", "0e18446744073709551616", "1e99999999999999999999",
    "6.0221409e+23", "1.602176634e-19", "68702699520", "749.999", "30.001", "94.123",
    "NaN", "nan", "NAN", "-nan", "+nan", "Inf", "+Inf", "-Inf", "inf", "INF", "infinity", "-Infinity",
    "+INFINITY", " 1.5", "1.5 ", "\t1.5\n", " nan ", "1_000", "1_0.0_1", "1e1_0", "1__0", "_1", "1_",
    "1._5", "1_.5", "1e_5", "1_e5", ".5", "5.", ".", "e5", "1e", "1e+", "1e-", "+", "-", "", " ", "+-1",
    "--1", "0x10", "0x1p3", "1.5.2", "1,5", "infinit", "infinityy", "in", "na", "nann", "1d5", "1f",
    "١٢٣", "1e5.0", "1 e5", "1e 5", "- 1", "١", "1\x00", "12345678901234567890123456789012345678901234567890",
    "0.000001", "1e-7", "1.7e-310", "4.4501477170144023e-308", "8.98846567431158e307",
    "0.1e1", "00001.5000", "-00.0e-00", "1e+00005", "000e500",
]


def float_golden():
    rnd = random.Random(0)
    toks = list(ADVERSARIAL)
    for _ in range(400):                                        # shortest round-trip reprs
        b = rnd.getrandbits(64)
        x = struct.unpack("<d", struct.pack("<Q", b))[0]
        if x == x and abs(x) != float("inf"):
            toks.append(repr(x))
    for _ in range(300):                                        # long mantissas, near-halfway
        e = rnd.randint(-340, 308)
        nd = rnd.choice([17, 18, 19, 20, 21, 25, 40, 80, 200, 770, 800])
        toks.append("".join(rnd.choice("0123456789") for _ in range(nd)).lstrip("0") + "5" + f"e{e - nd}")
    for _ in range(200):                                        # exact halfway points + neighbours
        m = rnd.getrandbits(52) | (1 << 52)
        e = rnd.randint(-60, 60)
        from fractions import Fraction
        half = Fraction(2 * m + 1, 2) * (Fraction(2) ** e)
        # halfway values have finite decimal expansions
        num, den = half.numerator, half.denominator
        k = 0
        while den % 2 == 0:
            den //= 2; k += 1
        digits = str(num * 5 ** k)
        base = digits[:-k] + "." + digits[-k:] if k else digits
        if k and len(digits) <= k:
            base = "0." + "0" * (k - len(digits)) + digits
        toks += [base, base + "0" * 5 + "1", base[:-1] + str((int(base[-1]) - 1) % 10) + "9" * 6]
    out = []
    for t in toks:
        try:
            out.append([t, bits(float(t))])
        except ValueError:
            out.append([t, "ValueError"])
    return out


def frame_to_json(piv):
    cells = {}
    for gid, row in piv.iterrows():
        cells[gid] = {c: (bits(v) if isinstance(v, float) else v) for c, v in row.items()}
    return {"index": list(piv.index), "columns": list(piv.columns), "cells": cells,
            "dtypes": {c: str(t) for c, t in piv.dtypes.items()}}


def instant_case(mod, errors, name, labels, values, body=None, selected=None):
    body = body if body is not None else synth.to_vector_json(labels, synth.START_MS, values)
    del errors[:]
    piv, stats = ref_app.run_fetch(mod, body.decode())
    case = {"name": name, "body": body.decode()}
    if piv is None:
        case["error"] = errors[-1]
        return case
    case["pivot"] = frame_to_json(piv)
    case["stats"] = {k: {c: bits(v) for c, v in s.items()} for k, s in stats.items()}
    sel = selected or list(piv.index)[:3]
    # app.py:335-345 on the reference's own frame
    f = piv.loc[sel]
    cols = [c for c in f.columns if c != "card_model"]
    avg = f[cols].mean()
    p = f["amd_gpu_average_package_power"]
    nz = p[p > 0]
    if not nz.empty:
        avg["amd_gpu_average_package_power"] = nz.mean()
    case["selected"] = sel
    case["averages"] = {c: bits(v) for c, v in avg.items()}
    return case


def instant_golden():
    mod, errors = ref_app.load_reference()
    cases = []
    for node in range(4):                                       # config 1: 4 nodes x 8 cards x 5 raw series
        labels, values, _ = synth.instant_snapshot(node, 8, seed=node)
        cases.append(instant_case(mod, errors, f"config1_node{node}", labels, values))
    labels, values, _ = synth.instant_snapshot(0, 12, seed=7)   # gpu_id '10','11' sort before '2'
    cases.append(instant_case(mod, errors, "lexicographic_12_cards", labels, values, selected=["10", "2", "0"]))
    lab, val = list(labels), list(values)
    val[2] = float("nan"); val[7] = 0.0; val[12] = 0.0          # NaN power, zero power readings
    cases.append(instant_case(mod, errors, "nan_and_zero_power", lab, val))
    val = list(values)
    for i, l in enumerate(lab):
        if l["__name__"] == "amd_gpu_total_vram" and l["gpu_id"] == "3":
            val[i] = 0.0
    cases.append(instant_case(mod, errors, "total_vram_zero", lab, val))
    val = list(values); val[5] = float("inf"); val[6] = float("-inf")
    cases.append(instant_case(mod, errors, "infinities", lab, val))
    cases.append(instant_case(mod, errors, "duplicate_key", lab + [lab[0]], values + [1.0]))
    lab2 = [dict(l) for l in lab]; del lab2[3]["gpu_id"]
    cases.append(instant_case(mod, errors, "missing_gpu_id", lab2, values))
    lab2 = [dict(l) for l in lab]; del lab2[0]["card_model"]
    cases.append(instant_case(mod, errors, "missing_card_model", lab2, values))
    lab2 = [l for l in lab if not (l["__name__"] == "amd_gpu_gfx_activity" and l["gpu_id"] == "1")]
    val2 = [v for l, v in zip(lab, values) if not (l["__name__"] == "amd_gpu_gfx_activity" and l["gpu_id"] == "1")]
    cases.append(instant_case(mod, errors, "missing_cell", lab2, val2))
    body = synth.to_vector_json(lab, synth.START_MS, values).replace(b'"' + synth.format_value(values[4]).encode() + b'"', b'"12abc"', 1)
    cases.append(instant_case(mod, errors, "malformed_value", lab, values, body=body))
    # two failures in one body: the reference reports the one its series walk meets first
    lab2 = [dict(l) for l in lab]; del lab2[2]["gpu_id"]
    body = synth.to_vector_json(lab2, synth.START_MS, values).replace(b'"' + synth.format_value(values[9]).encode() + b'"', b'"9x"', 1)
    cases.append(instant_case(mod, errors, "missing_label_before_malformed_value", lab2, values, body=body))
    lab2 = [dict(l) for l in lab]; del lab2[9]["__name__"]
    body = synth.to_vector_json(lab2, synth.START_MS, values).replace(b'"' + synth.format_value(values[2]).encode() + b'"', b'"2x"', 1)
    cases.append(instant_case(mod, errors, "malformed_value_before_missing_label", lab2, values, body=body))
    cases.append(instant_case(mod, errors, "empty_result", [], []))
    body = synth.to_vector_json(lab, synth.START_MS, values).replace(b'"' + synth.format_value(values[0]).encode() + b'"', b'" 4_2.5e0\\t"', 1)
    cases.append(instant_case(mod, errors, "python_float_grammar", lab, values, body=body))
    return {"versions": {"pandas": pd.__version__, "numpy": np.__version__, "python": sys.version.split()[0]},
            "cases": cases}


if __name__ == "__main__":
    with open(os.path.join(HERE, "float_golden.json"), "w") as f:
        json.dump(float_golden(), f, indent=0)
    with open(os.path.join(HERE, "instant_golden.json"), "w") as f:
        json.dump(instant_golden(), f, indent=0)
    print("wrote goldens")
