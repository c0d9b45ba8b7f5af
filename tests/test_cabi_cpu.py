"""CPU-side checks: the C-ABI library loads and exports every symbol that
include/bmfm.h declares; host-side packing logic."""
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_header_symbols():
    from bmfm import _cabi
    L = _cabi.lib()
    hdr = open(os.path.join(ROOT, "include", "bmfm.h")).read()
    declared = sorted(set(re.findall(r"\b(bmfm_[a-z0-9_]+)\s*\(", hdr)))
    assert declared, "no declarations found"
    for name in declared:
        assert hasattr(L, name), f"{name} declared in bmfm.h but not exported"
    assert declared == _cabi.exported_symbols(), "ctypes binding out of sync with bmfm.h"


def test_no_cpu_fallback():
    import torch
    from bmfm import _cabi
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(RuntimeError):
        _cabi.Context()


def test_pack_frame_matches_reference_ordering():
    from bmfm.panel import pack_frame
    from bmfm.synth import make_panel
    pan = make_panel(5, 40, 6, 3, seed=9, ragged=0.2)
    df = pan.to_frame(shuffle_seed=11)
    pk = pack_frame(df, 6, 3)
    csr = pan.to_csr()
    assert np.array_equal(pk.date_off, csr["date_off"])
    # reference order (date, stockname) recovered through name_pos
    assert np.array_equal(pk.cap[pk.name_pos], csr["cap"])
    assert np.array_equal(pk.ret[pk.name_pos], csr["ret"])
    assert np.array_equal(pk.ind[pk.name_pos], csr["ind"])
    assert np.array_equal(pk.style[pk.name_pos], csr["style"])
    # storage order is grouped by industry inside each date
    for t in range(pk.T):
        seg = pk.ind[pk.date_off[t]:pk.date_off[t + 1]]
        assert (np.diff(seg) >= 0).all()


def test_onehot_to_index_bit_exact():
    from bmfm.panel import onehot_to_index
    rng = np.random.default_rng(0)
    idx = rng.integers(-1, 5, size=200)
    oh = (idx[:, None] == np.arange(5)[None, :]).astype(np.int64)
    assert np.array_equal(onehot_to_index(oh), idx.astype(np.int32))
    oh[3, :] = 1
    with pytest.raises(NotImplementedError):
        onehot_to_index(oh)
