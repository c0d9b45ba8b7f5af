"""Array-level entry points over the C-ABI (device memory owned by torch).

Everything here runs on the GPU through libbmfm.so; numpy arrays in, numpy
arrays out.  The ``mfm`` drop-in package, the tests and ``bench.py`` call these.
"""
from __future__ import annotations

import numpy as np

from . import _cabi
from .panel import pack_arrays


def _torch():
    import torch
    return torch


def _dev(a, dtype=None):
    torch = _torch()
    t = torch.from_numpy(np.ascontiguousarray(a))
    if dtype is not None:
        t = t.to(dtype)
    return t.cuda(non_blocking=False)


class LinAlgError(np.linalg.LinAlgError):
    pass


def reg_by_time_device(ctx, T, P, Q, max_rows, d_off, d_cap, d_ret, d_ind, d_style, want_exposure=False):
    """Device tensors in, device tensors out (factor_ret [T,K], spec [rows], r2 [T], status [T], exposure|None)."""
    torch = _torch()
    K = 1 + P + Q
    dev = d_cap.device
    fr = torch.empty((T, K), dtype=torch.float64, device=dev)
    spec = torch.empty(d_cap.shape[0], dtype=torch.float64, device=dev)
    r2 = torch.empty(T, dtype=torch.float64, device=dev)
    status = torch.empty(T, dtype=torch.int32, device=dev)
    expo = torch.empty((T, K, K), dtype=torch.float64, device=dev) if want_exposure else None
    ctx.reg_by_time(T, P, Q, max_rows, d_off, d_cap, d_ret, d_ind, d_style, fr, spec, r2, status, expo)
    return fr, spec, r2, status, expo


def reg_by_time(date_off, cap, ret, ind, style, P, want_exposure=False, group_by_industry=True, ctx=None):
    """MFM.reg_by_time on date-major arrays (Barra-master/mfm/MFM.py:48-76).

    Inputs are in stock order inside each date; they are stored grouped by
    industry on the device and the specific returns are mapped back.
    Returns (factor_ret [T,K], specific_ret [rows], R2 [T], status [T], exposure or None).
    Raises numpy.linalg.LinAlgError where the reference's pinv would (non-finite Gram)."""
    ctx = ctx or _cabi.default_context()
    date_off = np.asarray(date_off, dtype=np.int64)
    T = len(date_off) - 1
    style = np.asarray(style, dtype=np.float64)
    Q = style.shape[1] if style.ndim == 2 else 0
    (c, r, i, s), pos = pack_arrays(date_off, cap, ret, ind, style.reshape(len(cap), Q), P, group_by_industry)
    max_rows = int(np.diff(date_off).max()) if T else 0
    fr, spec, r2, status, expo = reg_by_time_device(
        ctx, T, P, Q, max_rows, _dev(date_off), _dev(c), _dev(r), _dev(i), _dev(s), want_exposure)
    status = status.cpu().numpy()
    if (status == _cabi.ST_NONFINITE).any():
        raise LinAlgError("SVD did not converge")     # what numpy's pinv raises on a non-finite Gram
    spec = spec.cpu().numpy()[pos]
    return (fr.cpu().numpy(), spec, r2.cpu().numpy(), status,
            expo.cpu().numpy() if expo is not None else None)
