"""Hand-off between the descriptor pipeline and the risk model, and the result writers of demo.py.

In the reference the two halves meet through files: ``Barra_factor_cal/main.py:150`` stores the Barra table
in MongoDB, somebody exports ``barra_data_csi.csv`` / ``industry_info.csv``, and ``Barra-master/demo.py:22-35``
reads them back, drops rows with missing values and expands the industry code into one-hot columns.
``barra_frame_for_mfm`` does demo.py's part on the in-memory tables, so
``MFM(*barra_frame_for_mfm(final_barra_df, ind_info))`` continues without a round trip;
``write_results`` is demo.py:59-94 as a function over the model's return values.
"""
import os

import numpy as np
import pandas as pd


def barra_frame_for_mfm(data, industry_info):
    """demo.py:25-35 -> (frame in MFM's positional schema, P, Q): rows with any NaN dropped, index reset,
    one 0/1 column per row of ``industry_info`` (named by ``industry_names``, in that order) in place of the
    ``industry`` code column, style columns last."""
    naidx = 1 * np.sum(pd.isna(data), axis=1) > 0
    data = data[~naidx]
    data.index = range(len(data))
    industry = np.array([1 * (data.industry.values == x) for x in industry_info.code.values]).T
    industry = industry.reshape(len(data), len(industry_info))
    industry = pd.DataFrame(industry, columns=list(industry_info.industry_names.values))
    frame = pd.concat([data.iloc[:, :4], industry, data.iloc[:, 5:]], axis=1)
    return frame, len(industry_info), data.shape[1] - 5


def write_results(output_dir, model, factor_ret, specific_ret, R2, vr_cov_ls, lamb):
    """demo.py:59-94: the five result files.  Returns the list of paths written."""
    os.makedirs(output_dir, exist_ok=True)
    written = []

    def path(name):
        written.append(os.path.join(output_dir, name))
        return written[-1]

    factor_ret.to_csv(path('factor_returns.csv'))
    R2.to_csv(path('r_squared.csv'))
    if specific_ret:
        pd.concat(specific_ret).to_csv(path('specific_returns.csv'))
    if vr_cov_ls:
        vr_cov_ls[-1].to_csv(path('final_vol_regime_adj_covariance.csv'))
    if lamb:
        pd.Series(lamb, index=model.sorted_dates, name='lambda').to_csv(path('volatility_multiplier_lambda.csv'))
    return written


def read_results(results_dir):
    """The five CSVs demo.py:59-94 (or ``write_results``) writes -> dict of arrays (labels as unicode arrays)."""
    out = {}
    fr = pd.read_csv(os.path.join(results_dir, "factor_returns.csv"), index_col=0)
    out["factor_ret"] = fr.values.astype(float)
    out["factor_ret_index"] = np.asarray(fr.index.astype(str)).astype("U")
    out["factor_ret_columns"] = np.asarray(fr.columns.astype(str)).astype("U")
    r2 = pd.read_csv(os.path.join(results_dir, "r_squared.csv"), index_col=0)
    out["R2"] = r2.values[:, 0].astype(float)
    sp = pd.read_csv(os.path.join(results_dir, "specific_returns.csv"), index_col=0)
    out["specific_columns"] = np.asarray(sp.columns.astype(str)).astype("U")
    out["specific_index"] = np.asarray(sp.index.astype(str)).astype("U")
    out["specific_full"] = sp.values.astype(float)
    cov = pd.read_csv(os.path.join(results_dir, "final_vol_regime_adj_covariance.csv"), index_col=0)
    out["final_cov"] = cov.values.astype(float)
    out["final_cov_labels"] = np.asarray(cov.columns.astype(str)).astype("U")
    lam = pd.read_csv(os.path.join(results_dir, "volatility_multiplier_lambda.csv"), index_col=0)
    out["lamb"] = lam.values[:, 0].astype(float)
    return out
