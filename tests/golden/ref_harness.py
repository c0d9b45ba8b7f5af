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


def _install_stubs():
    if "matplotlib" not in sys.modules:
        mpl = types.ModuleType("matplotlib")
        plt = types.ModuleType("matplotlib.pyplot")
        plt.plot = lambda *a, **k: None
        plt.show = lambda *a, **k: None
        mpl.pyplot = plt
        sys.modules["matplotlib"] = mpl
        sys.modules["matplotlib.pyplot"] = plt

    if "statsmodels" in sys.modules:
        return
    sm_pkg = types.ModuleType("statsmodels")
    sm_stats = types.ModuleType("statsmodels.stats")
    sm_ws = types.ModuleType("statsmodels.stats.weightstats")
    sm_api = types.ModuleType("statsmodels.api")

    class DescrStatsW:
        def __init__(self, data, weights=None, ddof=0):
            self.data = np.asarray(data, dtype=float)
            self.weights = np.asarray(weights, dtype=float)

        @property
        def sum_weights(self):
            return self.weights.sum(0)

        @property
        def mean(self):
            # statsmodels: np.dot(self.data.T, self.weights) / self.sum_weights
            return np.dot(self.data.T, self.weights) / self.sum_weights

    sm_ws.DescrStatsW = DescrStatsW

    def add_constant(data, prepend=True, has_constant="skip"):
        if isinstance(data, pd.Series):
            data = data.to_frame()
        if isinstance(data, pd.DataFrame):
            vals = data.values.astype(float)
            # a column that is already a non-zero constant -> returned unchanged
            ptp = np.ptp(vals, axis=0) if len(vals) else np.zeros(vals.shape[1])
            is_const = (ptp == 0) & np.all(vals != 0, axis=0) if len(vals) else np.zeros(vals.shape[1], bool)
            if is_const.any():
                return data
            out = data.copy()
            out.insert(0, "const", 1.0)
            return out
        x = np.asarray(data, dtype=float)
        if x.ndim == 1:
            x = x[:, None]
        return np.column_stack([np.ones(len(x)), x])

    class _Results:
        def __init__(self, params, resid, scale):
            self.params = params
            self.resid = resid
            self.scale = scale

    class WLS:
        def __init__(self, endog, exog, weights=1.0):
            self.endog = endog
            self.exog = exog
            self.weights = weights

        def fit(self):
            y = np.asarray(self.endog, dtype=float)
            X = np.asarray(self.exog, dtype=float)
            w = np.broadcast_to(np.asarray(self.weights, dtype=float), y.shape)
            sw = np.sqrt(w)
            wx = X * sw[:, None]
            wy = y * sw
            pinv_wx = np.linalg.pinv(wx)
            beta = pinv_wx @ wy
            # statsmodels: rank = matrix_rank(diag(singular values of wexog))
            rank = np.linalg.matrix_rank(np.diag(np.linalg.svd(wx, compute_uv=False)))
            resid = y - X @ beta
            wresid = wy - wx @ beta
            df_resid = len(y) - rank
            scale = float(wresid @ wresid) / df_resid
            if isinstance(self.exog, pd.DataFrame):
                params = pd.Series(beta, index=self.exog.columns)
            else:
                params = beta
            if isinstance(self.endog, pd.Series):
                resid = pd.Series(resid, index=self.endog.index)
            return _Results(params, resid, scale)

    class OLS(WLS):
        def __init__(self, endog, exog):
            super().__init__(endog, exog, 1.0)

    sm_api.add_constant = add_constant
    sm_api.WLS = WLS
    sm_api.OLS = OLS
    sm_pkg.stats = sm_stats
    sm_pkg.api = sm_api
    sm_stats.weightstats = sm_ws
    sys.modules["statsmodels"] = sm_pkg
    sys.modules["statsmodels.stats"] = sm_stats
    sys.modules["statsmodels.stats.weightstats"] = sm_ws
    sys.modules["statsmodels.api"] = sm_api


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
