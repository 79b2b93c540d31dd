"""Drop-in mirror of the hot-path part of General_functions.py (same names, same argument meaning).

get_settings / get_features / pfml_feat_fun are configuration and column-name contracts of the reference's artefacts
(the database columns and pickle labels a drop-in has to read and write); create_cov, create_lambda and m_func run
on the GPU through libpfml_b200.so.  pandas in, pandas / numpy out, like the reference.
"""
from __future__ import annotations

import numpy as np
import pandas as pd

_FEATURES = (
    "age aliq_at aliq_mat ami_126d at_be at_gr1 at_me at_turnover be_gr1a be_me beta_60m beta_dimson_21d "
    "betabab_1260d betadown_252d bev_mev bidaskhl_21d capex_abn capx_gr1 capx_gr2 capx_gr3 cash_at "
    "chcsho_12m coa_gr1a col_gr1a cop_at cop_atl1 corr_1260d coskew_21d cowc_gr1a dbnetis_at debt_gr3 "
    "debt_me dgp_dsale div12m_me dolvol_126d dolvol_var_126d dsale_dinv dsale_drec dsale_dsga "
    "earnings_variability ebit_bev ebit_sale ebitda_mev emp_gr1 eq_dur eqnetis_at eqnpo_12m eqnpo_me "
    "eqpo_me f_score fcf_me fnl_gr1a gp_at gp_atl1 ival_me inv_gr1 inv_gr1a iskew_capm_21d iskew_ff3_21d "
    "iskew_hxz4_21d ivol_capm_21d ivol_capm_252d ivol_ff3_21d ivol_hxz4_21d kz_index lnoa_gr1a lti_gr1a "
    "market_equity mispricing_mgmt mispricing_perf ncoa_gr1a ncol_gr1a netdebt_me netis_at nfna_gr1a "
    "ni_ar1 ni_be ni_inc8q ni_ivol ni_me niq_at niq_at_chg1 niq_be niq_be_chg1 niq_su nncoa_gr1a noa_at "
    "noa_gr1a o_score oaccruals_at oaccruals_ni ocf_at ocf_at_chg1 ocf_me ocfq_saleq_std op_at op_atl1 "
    "ope_be ope_bel1 opex_at pi_nix ppeinv_gr1a prc prc_highprc_252d qmj qmj_growth qmj_prof qmj_safety "
    "rd_me rd_sale rd5_at resff3_12_1 resff3_6_1 ret_1_0 ret_12_1 ret_12_7 ret_3_1 ret_6_1 ret_60_12 "
    "ret_9_1 rmax1_21d rmax5_21d rmax5_rvol_21d rskew_21d rvol_21d sale_bev sale_emp_gr1 sale_gr1 "
    "sale_gr3 sale_me saleq_gr1 saleq_su seas_1_1an seas_1_1na seas_11_15an seas_11_15na seas_16_20an "
    "seas_16_20na seas_2_5an seas_2_5na seas_6_10an seas_6_10na sti_gr1a taccruals_at taccruals_ni "
    "tangibility tax_gr1a turnover_126d turnover_var_126d z_score zero_trades_126d zero_trades_21d "
    "zero_trades_252d rvol_252d "
).split()

_FEAT_EXCL = (
    "capex_abn capx_gr2 capx_gr3 debt_gr3 dgp_dsale dsale_dinv dsale_drec dsale_dsga earnings_variability "
    "eqnetis_at eqnpo_me eqpo_me f_score iskew_hxz4_21d ivol_hxz4_21d netis_at ni_ar1 ni_inc8q ni_ivol "
    "niq_at niq_at_chg1 niq_be niq_be_chg1 niq_su ocfq_saleq_std qmj qmj_growth rd_me rd_sale rd5_at "
    "resff3_12_1 resff3_6_1 sale_gr3 saleq_gr1 saleq_su seas_16_20an seas_16_20na sti_gr1a z_score "
).split()


def get_features(exclude_poor_coverage=True):
    """Column names of the JKP characteristics (General_functions.py:113-170): 154 names, 115 after the exclusion."""
    if exclude_poor_coverage:
        return [f for f in _FEATURES if f not in _FEAT_EXCL]
    return list(_FEATURES)


def get_settings():
    """The keys of General_functions.py:26-109 that the hot path reads, with the reference's defaults."""
    settings = {
        "parallel": True, "seed_no": 1, "months": False, "Transaction_Costs": True,
        "split": {"train_end": pd.to_datetime("1970-12-31"), "test_end": pd.to_datetime("2023-12-31"), "val_years": 10,
                  "model_update_freq": "yearly", "train_lookback": 1000, "retrain_lookback": 1000},
        "pf": {"dates": {"start_year": 1971, "end_yr": 2023, "split_years": 10}},
        "pf_ml": {"g_vec": list(np.exp(np.arange(-3, -1))), "p_vec": [2 ** i for i in range(6, 10)],
                  "l_vec": [0.0] + list(np.exp(np.linspace(-10, 10, 100))), "orig_feat": False, "scale": True},
        "cov_set": {"industries": True, "obs": 252 * 10, "hl_cor": int(252 * 3 / 2), "hl_var": int(252 / 2),
                    "hl_stock_var": int(252 / 2), "min_stock_obs": 252, "initial_var_obs": 21 * 3},
    }
    pf_set = {"wealth": 1e10, "gamma_rel": 10, "mu": 0.007, "lb_hor": 11}
    return settings, pf_set


def pfml_feat_fun(p):
    """General_functions.py:837-844."""
    return (["constant"] + ["rff" + str(x) + "_cos" for x in range(1, p // 2 + 1)]
            + ["rff" + str(x) + "_sin" for x in range(1, p // 2 + 1)])


def create_cov(x, ids=None):
    """General_functions.py:847-897: Sigma = load @ fct_cov @ load.T + diag(ivol) (DataFrame, labelled by id)."""
    from .device import dgemm
    if ids is None:
        load, ivol = x["fct_load"], x["ivol_vec"]
    else:
        load, ivol = x["fct_load"].loc[[i for i in ids]], x["ivol_vec"].loc[[i for i in ids]]
    L = np.ascontiguousarray(load.to_numpy(), dtype=np.float64)
    LF = dgemm(L[None], np.ascontiguousarray(x["fct_cov"].to_numpy(), dtype=np.float64)[None])[0]
    sigma = dgemm(LF[None], L[None], trans_b=True)[0] + np.diag(np.asarray(ivol, dtype=np.float64))
    if sigma.diagonal().min() < 0:
        print("Warning: Negative Variances:", np.arange(len(sigma)))
    return pd.DataFrame(sigma, index=load.index, columns=load.index)


def create_lambda(x, ids):
    """General_functions.py:900-916."""
    selected = [x[i] for i in ids] if isinstance(x, dict) else x[ids]
    return np.diag(selected)


def m_func(w, mu, rf, sigma_gam, gam, K_Lambda, iterations):
    """General_functions.py:919-962 on the GPU (pfml_mfunc_batched); K_Lambda is the diagonal matrix, as there."""
    from .device import m_func_batched
    sg = np.ascontiguousarray(np.asarray(sigma_gam, dtype=np.float64))[None]
    lam = np.ascontiguousarray(np.diag(np.asarray(K_Lambda, dtype=np.float64)))[None]
    m, status = m_func_batched(float(w), float(mu), float(rf), sg, float(gam), lam, int(iterations))
    if status[0]:
        raise np.linalg.LinAlgError(f"m_func: numerical failure on the device (status {int(status[0])})")
    return m[0]
