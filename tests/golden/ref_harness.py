"""Import the UNMODIFIED reference source from /root/reference (container only).

Used by ``make_golden.py`` to produce the committed fixtures, and by nothing
that runs on the GPU box (``/root/reference`` does not exist there).

The reference needs three third-party modules that are not installed here
(SURVEY.md §8(c)); they are stubbed with restatements of the published
behaviour of the calls the reference makes:

* ``matplotlib.pyplot``          — ``plot`` no-op (utils.py:11,116); no arithmetic.
* ``statsmodels.stats.weightstats.DescrStatsW(data, weights).mean``
                                 — ``data.T @ w / w.sum()`` (CrossSection.py:16-18,
                                   utils.py:25,36-37).
* ``statsmodels.api``            — ``add_constant`` (prepend, has_constant='skip'),
                                   ``WLS`` / ``OLS`` ``.fit()`` with the default
                                   ``method='pinv'``: ``params``, ``resid``, ``scale``
                                   (factor_calculator.py:96-102, post_processing.py:59-61).
  statsmodels' version is pinned nowhere in the reference (no requirements file).
"""
from __future__ import annotations

import contextlib
import io
import os
import sys
import types
import warnings

import numpy as np
import pandas as pd

REF_ROOT = "/root/reference"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))


from oracle.ref_stubs import install_stubs as _install_stubs  # noqa: E402


def load_reference():
    """Returns (mfm_MFM_module, mfm_CrossSection_module, mfm_utils_module,
    factor_calculator_module, post_processing_module, config_module)."""
    if not os.path.isdir(REF_ROOT):
        raise RuntimeError("reference tree not present (only available in the build container)")
    _install_stubs()
    warnings.filterwarnings("ignore", category=PendingDeprecationWarning)
    # make sure the *reference* mfm wins over this repo's drop-in of the same name
    for name in [m for m in sys.modules if m == "mfm" or m.startswith("mfm.")]:
        del sys.modules[name]
    for name in ("factor_calculator", "post_processing", "config"):
        sys.modules.pop(name, None)
    sys.path.insert(0, os.path.join(REF_ROOT, "Barra_factor_cal"))
    sys.path.insert(0, os.path.join(REF_ROOT, "Barra-master"))
    try:
        import mfm.MFM as ref_mfm
        import mfm.CrossSection as ref_cs
        import mfm.utils as ref_utils
        import factor_calculator as ref_fc
        import post_processing as ref_pp
        import config as ref_cfg
    finally:
        sys.path.remove(os.path.join(REF_ROOT, "Barra_factor_cal"))
        sys.path.remove(os.path.join(REF_ROOT, "Barra-master"))
    assert ref_mfm.__file__.startswith(REF_ROOT), ref_mfm.__file__
    # the reference modules stay reachable through the returned handles only
    for name in [m for m in sys.modules if m == "mfm" or m.startswith("mfm.")]:
        del sys.modules[name]
    for name in ("factor_calculator", "post_processing", "config"):
        sys.modules.pop(name, None)
    return ref_mfm, ref_cs, ref_utils, ref_fc, ref_pp, ref_cfg


@contextlib.contextmanager
def quiet():
    """The reference prints a status line per date (CrossSection.py:52-53)."""
    buf = io.StringIO()
    with contextlib.redirect_stdout(buf), warnings.catch_warnings():
        warnings.simplefilter("ignore")
        yield


def canonical_eig(a):
    """SURVEY.md §8(c) mode (ii): eigh of the symmetrised matrix, eigenvalues
    DESCENDING, each eigenvector sign-fixed so that its largest-magnitude
    component is positive.  Patched over ``mfm.utils.np.linalg.eig``."""
    a = np.asarray(a, dtype=float)
    w, v = np.linalg.eigh(0.5 * (a + a.T))
    order = np.argsort(-w, kind="stable")
    w = w[order]
    v = v[:, order]
    idx = np.argmax(np.abs(v), axis=0)
    sgn = np.sign(v[idx, np.arange(v.shape[1])])
    sgn[sgn == 0] = 1.0
    return w, v * sgn[None, :]


@contextlib.contextmanager
def patched_eig(ref_utils, fn=canonical_eig):
    """Run reference code with ``np.linalg.eig`` replaced *inside mfm.utils only*."""

    class _LinalgProxy:
        def __init__(self, real):
            self._real = real

        def __getattr__(self, k):
            if k == "eig":
                return fn
            return getattr(self._real, k)

    class _NpProxy:
        def __init__(self, real):
            self._real = real
            self.linalg = _LinalgProxy(real.linalg)

        def __getattr__(self, k):
            return getattr(self._real, k)

    real_np = ref_utils.np
    ref_utils.np = _NpProxy(real_np)
    try:
        yield
    finally:
        ref_utils.np = real_np
