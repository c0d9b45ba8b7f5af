"""The timed CPU arm of bench.py: the reference's own ``mfm`` package from oracle/_ref (byte-compiled by
oracle/build_ref.py where /root/reference exists).  When it is there it must load with the stand-ins and agree with
the numpy port on a small cross-section — which also pins the port against the reference at run time."""
import contextlib
import io
import os

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_loads_and_matches_port():
    from oracle import barra_oracle as orc
    from oracle import ref_loader
    if os.path.isdir("/root/reference"):
        import importlib.util
        spec = importlib.util.spec_from_file_location("build_ref", os.path.join(ROOT, "oracle", "build_ref.py"))
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        assert mod.build(quiet=True)
    ref = ref_loader.load()
    if ref is None:
        pytest.skip("oracle/_ref has not been built (no /root/reference on this machine)")
    ref_mfm, ref_cs, ref_utils = ref
    assert ref_mfm.__file__.startswith(os.path.join(ROOT, "oracle", "_ref"))
    import mfm.MFM as dropin                    # the drop-in keeps the name `mfm` after the loader ran
    assert dropin.__file__.startswith(os.path.join(ROOT, "llm-driven-multi-factor-model_b200"))
    from bmfm.synth import make_panel, make_factor_returns
    import pandas as pd
    pan = make_panel(T=3, N=80, P=5, Q=4, seed=5)
    with contextlib.redirect_stdout(io.StringIO()):
        model = ref_mfm.MFM(pan.to_frame(), pan.P, pan.Q)
        fr, spec, r2 = model.reg_by_time()
    csr = pan.to_csr()
    wfr, wspec, wr2 = orc.reg_by_time(csr["date_off"], csr["cap"], csr["ret"], csr["ind"], csr["style"], pan.P)
    assert np.max(np.abs(fr.values - wfr)) < 1e-12 * np.max(np.abs(wfr))
    x = pd.DataFrame(make_factor_returns(60, 6, seed=2))
    assert np.max(np.abs(ref_utils.Newey_West(x, 2, 20).values - orc.newey_west(x.values, 2, 20))) < 1e-18
