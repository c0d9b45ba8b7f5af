import numpy as np


def relerr(a, b, scale=None):
    """max |a-b| / max |b| (norm-wise), NaN patterns must agree."""
    a = np.asarray(a, dtype=float)
    b = np.asarray(b, dtype=float)
    assert a.shape == b.shape, (a.shape, b.shape)
    na, nb = np.isnan(a), np.isnan(b)
    assert np.array_equal(na, nb), "NaN pattern differs"
    if a.size == 0 or na.all():
        return 0.0
    if scale is None:
        scale = np.nanmax(np.abs(b))
    if scale == 0:
        return float(np.nanmax(np.abs(a - b)))
    return float(np.nanmax(np.abs(a - b)) / scale)
