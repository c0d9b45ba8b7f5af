"""Stand-ins for the three third-party modules the reference imports and this image lacks
(SURVEY.md §8(c)): restatements of the published behaviour of exactly the calls the reference makes.

* ``matplotlib.pyplot``          — ``plot`` no-op (utils.py:11,116); no arithmetic.
* ``statsmodels.stats.weightstats.DescrStatsW(data, weights).mean``
                                 — ``data.T @ w / w.sum()`` (CrossSection.py:16-18, utils.py:25,36-37).
* ``statsmodels.api``            — ``add_constant`` (prepend, has_constant='skip'), ``WLS`` / ``OLS``
                                   ``.fit()`` with the default ``method='pinv'``: ``params``, ``resid``,
                                   ``scale`` (factor_calculator.py:96-102, post_processing.py:59-61).
  statsmodels' version is pinned nowhere in the reference (no requirements file).

Test / baseline infrastructure only (used by tests/golden/ref_harness.py and by the timed reference arm of
bench.py through oracle/ref_loader.py); nothing in the product package imports it.
"""
from __future__ import annotations

import sys
import types

import numpy as np
import pandas as pd


def install_stubs():
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
