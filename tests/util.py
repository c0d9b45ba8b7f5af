"""Error measures of the parity tests, and a recorder of the errors actually observed.

``relerr``      norm-wise: max |a-b| / max |b| over the whole array (one scale).
``relerr_cols`` per column: every column of a [T, K] array (factor returns: one factor each) is held
                to its OWN largest magnitude, so a factor whose returns are 100x smaller than the
                country factor's is pinned just as tightly.
``relerr_cov``  per matrix and element: |a_ij - b_ij| / sqrt(b_ii b_jj), the scale a covariance element
                lives on (a small covariance between two large variances is not a small number).
NaN patterns must agree in all of them.  ``record`` keeps the observed value of every named check; the
session writes them to ``gpurun_out/parity_observed.json`` (tests/conftest.py) so DESIGN.md §2 can quote
measured errors next to the asserted bounds.
"""
import numpy as np

OBSERVED = {}


def record(name, value):
    """Remember the largest observed error of a named check."""
    v = float(value)
    if name not in OBSERVED or not (v <= OBSERVED[name]):
        OBSERVED[name] = v
    return v


def _prep(a, b):
    a = np.asarray(a, dtype=float)
    b = np.asarray(b, dtype=float)
    assert a.shape == b.shape, (a.shape, b.shape)
    na, nb = np.isnan(a), np.isnan(b)
    assert np.array_equal(na, nb), "NaN pattern differs"
    return a, b, na


def relerr(a, b, scale=None):
    """max |a-b| / max |b| (norm-wise), NaN patterns must agree."""
    a, b, na = _prep(a, b)
    if a.size == 0 or na.all():
        return 0.0
    if scale is None:
        scale = np.nanmax(np.abs(b))
    if scale == 0:
        return float(np.nanmax(np.abs(a - b)))
    return float(np.nanmax(np.abs(a - b)) / scale)


def relerr_cols(a, b):
    """max over columns of (max |a-b| in the column / max |b| in the column); a, b [rows, cols]."""
    a, b, na = _prep(a, b)
    if a.size == 0 or na.all():
        return 0.0
    a = a.reshape(-1, a.shape[-1])
    b = b.reshape(-1, b.shape[-1])
    with np.errstate(invalid="ignore"):
        diff = np.nanmax(np.abs(a - b), axis=0)
        scale = np.nanmax(np.abs(b), axis=0)
    ok = np.isfinite(scale)
    diff, scale = diff[ok], scale[ok]
    scale = np.where(scale == 0, 1.0, scale)
    return float(np.max(diff / scale)) if diff.size else 0.0


def relerr_cov(a, b):
    """max |a_ij - b_ij| / sqrt(b_ii b_jj) over the elements of every K x K matrix of a [..., K, K] stack
    (matrices that are all-NaN in both are skipped)."""
    a, b, na = _prep(a, b)
    K = a.shape[-1]
    a = a.reshape(-1, K, K)
    b = b.reshape(-1, K, K)
    worst = 0.0
    for x, y in zip(a, b):
        if np.isnan(y).all():
            continue
        d = np.sqrt(np.abs(np.diag(y)))
        s = np.outer(d, d)
        s = np.where(s == 0, np.nanmax(np.abs(y)) or 1.0, s)
        worst = max(worst, float(np.nanmax(np.abs(x - y) / s)))
    return worst
