"""DataFrame -> date-major packed panel (the HBM layout of the regression path).

Mirrors what ``MFM.__init__`` / ``MFM.reg_by_time`` do to the long table
(Barra-master/mfm/MFM.py:28-36, 58-61): dates parsed with ``pd.to_datetime``,
sorted unique; inside a date rows ordered by ``stocknames``; positional columns
0 date, 1 stocknames, 2 capital, 3 ret, 4..4+P one-hot industries, last Q
styles.

Device layout (all contiguous, fp64 unless stated):
    date_off int64[T+1] | cap[rows] | ret[rows] | ind int32[rows] | style[rows, Q]
Rows of one date are stored grouped by industry index (then by stock name):
the kernel's per-industry sums then walk contiguous runs.  ``name_pos`` maps the
reference's (date, stockname) order back onto that storage order.
"""
from __future__ import annotations

import numpy as np
import pandas as pd


class PackedPanel:
    __slots__ = ("T", "P", "Q", "K", "rows", "sorted_dates", "date_off", "cap", "ret", "ind",
                 "style", "name_pos", "names_sorted", "max_rows", "last_capital")


def onehot_to_index(block):
    """Bit-exact one-hot -> industry index.  Rows must hold only 0/1 with at most
    one 1 (the ``1*(industry == code)`` construction of demo.py:33); a row of
    zeros becomes -1."""
    block = np.asarray(block)
    if block.shape[1] == 0:
        return np.full(block.shape[0], -1, dtype=np.int32)
    ok = np.logical_or(block == 0, block == 1)
    if not ok.all():
        raise NotImplementedError("industry block must be a 0/1 one-hot matrix")
    cnt = (block != 0).sum(axis=1)
    if (cnt > 1).any():
        raise NotImplementedError("a stock belongs to more than one industry column")
    idx = np.argmax(block != 0, axis=1).astype(np.int32)
    idx[cnt == 0] = -1
    return idx


def pack_frame(data, P, Q):
    """Pack the reference's long table.  Host-side, vectorised (no per-date loop)."""
    out = PackedPanel()
    dates = pd.to_datetime(data.iloc[:, 0].values)
    uniq, date_code = np.unique(dates.values, return_inverse=True)
    out.sorted_dates = pd.to_datetime(uniq)
    T = len(uniq)
    names = data.iloc[:, 1].values
    name_code, name_uniques = pd.factorize(names, sort=True)
    ind = onehot_to_index(data.iloc[:, 4:4 + P].values)
    ncols = data.shape[1]
    style = np.ascontiguousarray(data.iloc[:, ncols - Q:].values, dtype=np.float64) if Q > 0 \
        else np.zeros((len(data), 0))
    cap = np.asarray(data.iloc[:, 2].values, dtype=np.float64)
    ret = np.asarray(data.iloc[:, 3].values, dtype=np.float64)

    # storage order: (date, industry, stockname); rows without industry last inside a date
    ind_key = np.where(ind < 0, P, ind)
    dev_order = np.lexsort((name_code, ind_key, date_code))
    # the reference's order: (date, stockname), stable like DataFrame.sort_values (MFM.py:59)
    ref_order = np.lexsort((name_code, date_code))
    pos = np.empty(len(data), dtype=np.int64)
    pos[dev_order] = np.arange(len(data), dtype=np.int64)

    counts = np.bincount(date_code, minlength=T)
    date_off = np.zeros(T + 1, dtype=np.int64)
    np.cumsum(counts, out=date_off[1:])

    out.T, out.P, out.Q, out.K = T, int(P), int(Q), 1 + int(P) + int(Q)
    out.rows = len(data)
    out.date_off = date_off
    out.cap = np.ascontiguousarray(cap[dev_order])
    out.ret = np.ascontiguousarray(ret[dev_order])
    out.ind = np.ascontiguousarray(ind[dev_order])
    out.style = np.ascontiguousarray(style[dev_order])
    out.name_pos = pos[ref_order]                       # storage index of the j-th row in reference order
    out.names_sorted = np.asarray(name_uniques)[name_code[ref_order]]
    out.max_rows = int(counts.max()) if T else 0
    a = date_off[T - 1] if T else 0
    out.last_capital = cap[ref_order][a:] if T else cap[:0]     # MFM.py:68
    return out


def pack_arrays(date_off, cap, ret, ind, style, P, group_by_industry=True):
    """Pack already date-major arrays (stock order inside a date), e.g. from
    ``synth.Panel.to_csr``.  Returns (arrays in storage order, name_pos)."""
    date_off = np.asarray(date_off, dtype=np.int64)
    rows = int(date_off[-1])
    T = len(date_off) - 1
    if not group_by_industry:
        order = np.arange(rows, dtype=np.int64)
    else:
        date_code = np.repeat(np.arange(T), np.diff(date_off))
        key = np.where(np.asarray(ind) < 0, P, ind)
        order = np.lexsort((np.arange(rows), key, date_code))
    pos = np.empty(rows, dtype=np.int64)
    pos[order] = np.arange(rows, dtype=np.int64)
    return (np.ascontiguousarray(np.asarray(cap, dtype=np.float64)[order]),
            np.ascontiguousarray(np.asarray(ret, dtype=np.float64)[order]),
            np.ascontiguousarray(np.asarray(ind, dtype=np.int32)[order]),
            np.ascontiguousarray(np.asarray(style, dtype=np.float64)[order])), pos
