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


class DevicePanel:
    """The packed panel resident on the GPU (torch tensors) plus the small host-side bookkeeping the
    DataFrame wrappers need.  ``name_pos`` (device, int64) maps the reference's (date, stockname) order onto
    the storage order; ``ref_order`` / ``name_code`` stay on the device until a DataFrame view asks for names."""
    __slots__ = ("T", "P", "Q", "K", "rows", "sorted_dates", "date_off", "date_off_host", "cap", "ret", "ind",
                 "style", "name_pos", "ref_order", "name_uniques", "name_code", "max_rows", "last_capital")


def pack_frame_device(data, P, Q, ctx=None, dates=None):
    """``pack_frame`` with the row work on the GPU: the raw columns are uploaded as pandas stores them, the
    one-hot block becomes an industry index in a kernel (bmfm_onehot_to_index, same checks as the host packer),
    the (date, industry, stock name) order comes from one radix sort of a 64-bit key per row, and
    bmfm_pack_gather writes the date-major layout the regression reads.  Only the strings are handled on the host:
    dates are parsed once (``dates``: the DatetimeIndex MFM.__init__ already built) and stock names factorised."""
    import torch
    from . import _cabi
    ctx = ctx or _cabi.default_context()
    dev = torch.device("cuda", ctx.device)
    rows = len(data)
    out = DevicePanel()
    if dates is None:
        dates = pd.to_datetime(data.iloc[:, 0].values)
    date_ns = torch.from_numpy(np.asarray(dates.values).astype("datetime64[ns]").view(np.int64)).to(dev)
    name_code, name_uniques = pd.factorize(data.iloc[:, 1].values, sort=True)
    n_names = max(1, len(name_uniques))
    d_name = torch.from_numpy(np.asarray(name_code, dtype=np.int64)).to(dev)
    uniq_ns, date_code = torch.unique(date_ns, sorted=True, return_inverse=True)
    T = int(uniq_ns.numel())
    out.sorted_dates = pd.to_datetime(uniq_ns.cpu().numpy().view("datetime64[ns]"))
    ncols = data.shape[1]

    def block(lo, hi, dtype):
        """columns lo..hi as a [rows, n] tensor on the device without a host transpose (pandas keeps a block
        column-major; the strides travel with the tensor)"""
        a = data.iloc[:, lo:hi].to_numpy()
        if a.dtype != dtype:
            a = a.astype(dtype)
        t = torch.from_numpy(a)
        return t.to(dev) if t.is_contiguous() or t.t().is_contiguous() else t.contiguous().to(dev)

    ind = torch.empty(rows, dtype=torch.int32, device=dev)
    flags = torch.zeros(2, dtype=torch.int32, device=dev)
    if P > 0:
        oh = block(4, 4 + P, np.int64)
        ctx.onehot_to_index(rows, P, oh, oh.stride(0), oh.stride(1), ind, flags)
    else:
        ind.fill_(-1)
    cap = torch.from_numpy(np.asarray(data.iloc[:, 2].values, dtype=np.float64)).to(dev)
    ret = torch.from_numpy(np.asarray(data.iloc[:, 3].values, dtype=np.float64)).to(dev)
    sty = block(ncols - Q, ncols, np.float64) if Q > 0 else None
    ind_key = torch.where(ind < 0, torch.full_like(ind, P), ind).to(torch.int64)
    # storage order (date, industry, stockname) and the reference's order (date, stockname), both stable
    key = (date_code * (P + 1) + ind_key) * n_names + d_name
    dev_order = torch.sort(key, stable=True).indices
    ref_order = torch.sort(date_code * n_names + d_name, stable=True).indices
    pos = torch.empty(rows, dtype=torch.int64, device=dev)
    pos[dev_order] = torch.arange(rows, dtype=torch.int64, device=dev)
    counts = torch.bincount(date_code, minlength=T)
    date_off = torch.zeros(T + 1, dtype=torch.int64, device=dev)
    date_off[1:] = torch.cumsum(counts, 0)
    out.cap = torch.empty(rows, dtype=torch.float64, device=dev)
    out.ret = torch.empty(rows, dtype=torch.float64, device=dev)
    out.ind = torch.empty(rows, dtype=torch.int32, device=dev)
    out.style = torch.empty((rows, Q), dtype=torch.float64, device=dev)
    ctx.pack_gather(rows, Q, dev_order, cap, ret, ind, sty, sty.stride(0) if Q > 0 else 0, sty.stride(1) if Q > 0 else 0,
                    out.cap, out.ret, out.ind, out.style)
    bad = flags.cpu().numpy()
    if bad[0]:
        raise NotImplementedError("industry block must be a 0/1 one-hot matrix")
    if bad[1]:
        raise NotImplementedError("a stock belongs to more than one industry column")
    out.T, out.P, out.Q, out.K, out.rows = T, int(P), int(Q), 1 + int(P) + int(Q), rows
    out.date_off = date_off
    out.date_off_host = date_off.cpu().numpy()
    out.name_pos = pos[ref_order]
    out.ref_order = ref_order
    out.name_uniques = np.asarray(name_uniques)
    out.name_code = d_name
    out.max_rows = int(counts.max().item()) if T else 0
    a = int(out.date_off_host[T - 1]) if T else 0
    out.last_capital = cap[ref_order[a:]].cpu().numpy() if T else np.zeros(0)      # MFM.py:68
    return out


class LazyFrameList(list):
    """``MFM.reg_by_time``'s list of T one-row DataFrames (MFM.py:66), built on demand.

    The reference constructs T DataFrames in its date loop; for a 15-year panel that construction costs far more
    than the regressions.  Here the specific returns stay one array (fetched from the device on first use) and
    element t becomes a DataFrame — same values, columns (that date's stock names, sorted) and index (the date) —
    when somebody asks for it: indexing, iteration, ``pd.concat(specific_ret)`` (demo.py:78) all work as on a list.
    """

    def __init__(self, T, fetch):
        super().__init__([None] * T)
        self._fetch = fetch            # () -> (values [rows] in reference order, names [rows], date_off [T+1], dates)
        self._data = None

    def _frame(self, t):
        cur = list.__getitem__(self, t)
        if cur is None:
            if self._data is None:
                self._data = self._fetch()
            vals, names, off, dates = self._data
            a, b = off[t], off[t + 1]
            cur = pd.DataFrame(vals[None, a:b], columns=pd.Index(names[a:b]), index=dates[t:t + 1])
            list.__setitem__(self, t, cur)
        return cur

    def __getitem__(self, i):
        if isinstance(i, slice):
            return [self._frame(t) for t in range(*i.indices(len(self)))]
        return self._frame(i if i >= 0 else len(self) + i)

    def __iter__(self):
        return (self._frame(t) for t in range(len(self)))

    def __reversed__(self):
        return (self._frame(t) for t in range(len(self) - 1, -1, -1))
