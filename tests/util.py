"""Comparison rules of SURVEY.md section 8(a): bit-exact vs 1e-12 relative."""
import numpy as np


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float64).view(np.uint64)


def assert_bits_equal(got, exp, what=""):
    g, e = bits(got), bits(exp)
    bad = np.flatnonzero(g.ravel() != e.ravel())
    assert len(bad) == 0, f"{what}: {len(bad)} of {g.size} differ, first at {bad[:5]}: " \
                          f"{np.asarray(got).ravel()[bad[:5]]} vs {np.asarray(exp).ravel()[bad[:5]]}"


def assert_exact_or_nan(got, exp, what=""):
    """Bit-exact for numbers (with -0.0 == 0.0), NaN where the oracle has NaN."""
    got, exp = np.asarray(got, dtype=np.float64), np.asarray(exp, dtype=np.float64)
    assert got.shape == exp.shape, f"{what}: shape {got.shape} vs {exp.shape}"
    gn, en = np.isnan(got), np.isnan(exp)
    bad = np.flatnonzero((gn != en).ravel() | (~en & (got != exp)).ravel())
    assert len(bad) == 0, f"{what}: {len(bad)} of {got.size} differ, first at {bad[:5]}: " \
                          f"{got.ravel()[bad[:5]]} vs {exp.ravel()[bad[:5]]}"


WORST = {}     # what -> worst observed |got - exp| / max(|exp|, scale); printed at the end of the run (conftest.py)


def assert_close(got, exp, scale=None, rtol=1e-12, what=""):
    """|got - exp| <= rtol * max(|exp|, scale); NaN pattern identical; inf must match exactly.
    ``scale=None``: relative to the result itself."""
    got, exp = np.asarray(got, dtype=np.float64), np.asarray(exp, dtype=np.float64)
    assert got.shape == exp.shape, f"{what}: shape {got.shape} vs {exp.shape}"
    gn, en = np.isnan(got), np.isnan(exp)
    assert np.array_equal(gn, en), f"{what}: NaN pattern differs at {np.flatnonzero((gn != en).ravel())[:5]}"
    fin = np.isfinite(exp)
    assert np.array_equal(got[~fin & ~en], exp[~fin & ~en]), f"{what}: infinities differ"
    tol = np.abs(exp) if scale is None else np.maximum(np.abs(exp), np.asarray(scale, dtype=np.float64))
    err = np.abs(got - exp)
    if what and fin.any():
        with np.errstate(divide="ignore", invalid="ignore"):
            rel = np.where(fin & (tol > 0), err / tol, 0.0)
        WORST[what] = max(WORST.get(what, 0.0), float(np.nanmax(rel)))
    bad = np.flatnonzero((fin & (err > rtol * tol)).ravel())
    assert len(bad) == 0, f"{what}: {len(bad)} of {got.size} beyond {rtol} rel, first at {bad[:5]}: " \
                          f"{got.ravel()[bad[:5]]} vs {exp.ravel()[bad[:5]]}"
