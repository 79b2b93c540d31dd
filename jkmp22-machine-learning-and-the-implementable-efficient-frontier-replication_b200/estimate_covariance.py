"""The numerical pieces of "Estimate Covariance Matrix.py" on the GPU.

Same names as the reference where it has them (weighted_cov_wt, weighted_cor_wt, ewma_vol); the per-day OLS and
the per-calc-date factor covariance are the loop bodies at :214-241 and :297-335.  Date logic (calc dates,
month-ends, the 200-of-252 coverage rule) stays in host code.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib
from ._lib import check


def _ptr(t):
    return C.c_void_p(t.data_ptr())


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _dev(a, device, dtype=torch.float64):
    return a.to(device) if torch.is_tensor(a) else torch.as_tensor(np.ascontiguousarray(a), dtype=dtype).to(device)


def daily_ols(X, y, day_ptr, device="cuda:0"):
    """Per-day cross-sectional OLS.  X [rows,K], y [rows] sorted by day, day_ptr [D+1] row offsets.
    Returns (fct_ret [D,K], residual [rows], status [D])."""
    lib = _lib.load()
    dev = torch.device(device)
    Xd, yd = _dev(X, dev), _dev(y, dev)
    pd_ = _dev(np.asarray(day_ptr, dtype=np.int64), dev, torch.int64)
    D, K = len(day_ptr) - 1, Xd.shape[1]
    coef = torch.empty((D, K), dtype=torch.float64, device=dev)
    resid = torch.empty_like(yd)
    status = torch.zeros(D, dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        check(lib.pfml_xs_ols_daily(_ptr(Xd), _ptr(yd), _ptr(pd_), D, K, _ptr(coef), _ptr(resid), _ptr(status), _stream()),
              "pfml_xs_ols_daily")
        torch.cuda.synchronize(dev)
    return coef.cpu().numpy(), resid.cpu().numpy(), status.cpu().numpy()


def ewma_weights(obs, half_life):
    """"Estimate Covariance Matrix.py":280-282."""
    return (0.5 ** (1 / half_life)) ** np.arange(obs, 0, -1)


def factor_cov(fct_ret, end_idx, obs=2520, hl_cor=378, hl_var=126, scale=1.0, device="cuda:0", w_cor=None, w_var=None):
    """sd_var @ weighted_cor_wt @ sd_var per calc date (:297-335); window = last min(obs, end+1) rows up to end_idx."""
    lib = _lib.load()
    dev = torch.device(device)
    F = _dev(fct_ret, dev)
    K = F.shape[1]
    ends = _dev(np.asarray(end_idx, dtype=np.int32), dev, torch.int32)
    wc = _dev(ewma_weights(obs, hl_cor) if w_cor is None else w_cor, dev)
    wv = _dev(ewma_weights(obs, hl_var) if w_var is None else w_var, dev)
    out = torch.empty((len(end_idx), K, K), dtype=torch.float64, device=dev)
    with torch.cuda.device(dev):
        check(lib.pfml_wcov_wcor(_ptr(F), _ptr(ends), len(end_idx), K, int(obs), _ptr(wc), _ptr(wv), float(scale), _ptr(out),
                                 _stream()), "pfml_wcov_wcor")
        torch.cuda.synchronize(dev)
    return out.cpu().numpy()


def weighted_cov_wt(df, weights, device="cuda:0"):
    """General_functions.py:745-784 (values of a DataFrame or an array in, array out)."""
    X = np.asarray(getattr(df, "to_numpy", lambda: df)(), dtype=np.float64)
    t = len(X)
    w = np.asarray(weights, dtype=np.float64)
    cov = _full_wcov(X, w, device)
    return cov


def _full_wcov(X, w, device):
    # the device kernel returns sd_v cor_c sd_v; with w_cor == w_var that is exactly the weighted covariance
    return factor_cov(X, [len(X) - 1], obs=len(X), w_cor=w, w_var=w, device=device)[0]


def weighted_cor_wt(df, weights, device="cuda:0"):
    """General_functions.py:786-835: unit-diagonal weighted correlation."""
    cov = weighted_cov_wt(df, weights, device)
    sd = np.sqrt(np.diag(cov))
    cor = cov / np.outer(sd, sd)
    np.fill_diagonal(cor, 1.0)
    return cor


def ewma_vol_batch(series, lam, start, device="cuda:0"):
    """ewma_vol for a list of 1-D residual series (one per stock); returns a list of vol arrays."""
    lib = _lib.load()
    dev = torch.device(device)
    lens = np.array([len(s) for s in series], dtype=np.int64)
    ptr = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    x = _dev(np.concatenate([np.asarray(s, dtype=np.float64) for s in series]) if len(series) else np.zeros(0), dev)
    pt = _dev(ptr, dev, torch.int64)
    vol = torch.empty_like(x)
    with torch.cuda.device(dev):
        check(lib.pfml_ewma_vol(_ptr(x), _ptr(pt), len(series), float(lam), int(start), _ptr(vol), _stream()), "pfml_ewma_vol")
        torch.cuda.synchronize(dev)
    v = vol.cpu().numpy()
    return [v[ptr[i]:ptr[i + 1]] for i in range(len(series))]


def ewma_vol(x, lam, start, device="cuda:0"):
    """"Estimate Covariance Matrix.py":345-386 for one series."""
    return ewma_vol_batch([x], lam, start, device)[0]


def ewma_wrapper(series, settings=None, device="cuda:0"):
    """"Estimate Covariance Matrix.py":389-397: Series -> ewma_vol with lam = 0.5 ** (1 / hl_stock_var) and
    start = initial_var_obs from ``settings['cov_set']`` (the reference reads the global ``settings``; here it is an
    argument that defaults to ``get_settings()``) -> Series on the same index.  For many stocks use ewma_vol_batch: one launch."""
    import pandas as pd

    from . import general_functions as GF
    cs = (settings or GF.get_settings()[0])["cov_set"]
    out = ewma_vol(series.values.astype(float), 0.5 ** (1 / cs["hl_stock_var"]), cs["initial_var_obs"], device)
    return pd.Series(out, index=series.index)


# ------------------------------------------------------------------------------------------------------
# script-level drop-in
# ------------------------------------------------------------------------------------------------------
def run(path, settings=None, device="cuda:0"):
    """Drop-in for "Estimate Covariance Matrix.py": reads Data/JKP_US_SP500.db:Factors_processed,
    Data/crsp_daily_SP500.db:d_ret_ex and Data/cluster_labels_processed.csv (:40-106), writes Data/Barra_Cov.pkl
    = {eom: {'fct_load', 'fct_cov', 'ivol_vec'}} (:490-497).

    pandas does the bookkeeping the reference does with pandas (joins, calendars, group medians); the three
    numerical loops -- daily OLS (:214-241), factor covariance (:297-335), EWMA volatility (:345-405) -- are
    single kernel launches over all days / calc dates / stocks."""
    import os
    import pickle
    import sqlite3

    import pandas as pd

    from . import general_functions as GF
    settings, _pf = settings or GF.get_settings()
    cs = settings["cov_set"]
    features = GF.get_features(exclude_poor_coverage=True)
    data = os.path.join(path, "Data")
    con = sqlite3.connect(os.path.join(data, "JKP_US_SP500.db"))
    chars = pd.read_sql_query("SELECT id, eom, size_grp, ff12, valid, " + ", ".join(features) + " FROM Factors_processed",
                              con, parse_dates=["eom"])
    con.close()
    chars["valid"] = chars["valid"].astype(bool)
    con = sqlite3.connect(os.path.join(data, "crsp_daily_SP500.db"))
    daily = pd.read_sql_query("SELECT permno as id, date, ret_excess as ret_exc FROM d_ret_ex", con, parse_dates={"date"})
    con.close()
    labels = pd.read_csv(os.path.join(data, "cluster_labels_processed.csv"))

    valid_ids = chars.loc[chars["valid"], "id"].unique()
    daily = daily[daily["ret_exc"].notna() & daily["id"].isin(valid_ids)].copy()                     # :94-95
    daily["eom_ret"] = daily["date"] + pd.offsets.MonthEnd(0)                                       # :98, :171

    # ---- a14: cluster ranks, industry dummies, cross-sectional z-scores (:115-158) ---------------------
    m = chars.loc[chars["valid"], ["id", "eom", "size_grp", "ff12"] + features].reset_index(drop=True)
    clusters = labels["cluster"].unique().tolist()
    ranks = {}
    for cl in clusters:
        sub = labels[(labels["cluster"] == cl) & labels["characteristic"].isin(features)]
        sign = sub["direction"].to_numpy()
        v = m[sub["characteristic"].tolist()].to_numpy(dtype=np.float64)
        v = np.where(sign[None, :] == -1, 1.0 - v, v)
        ranks[cl] = np.nanmean(v, axis=1)
    cm = m[["id", "eom", "size_grp", "ff12"]].copy()
    cm["eom_ret"] = cm["eom"] + pd.offsets.MonthEnd(1)
    for cl in clusters:
        cm[cl] = ranks[cl]
    industries = sorted(cm["ff12"].dropna().unique())
    for ind in industries:
        cm[str(ind)] = (cm["ff12"] == ind).astype(int)
    grp = cm.groupby("eom")[clusters]
    cm[clusters] = (cm[clusters] - grp.transform("mean")) / grp.transform("std")
    factor_cols = [str(i) for i in industries] + clusters
    K = len(factor_cols)

    # ---- regression table: month-m characteristics explain the days of month m+1 (:168-183) -------------
    daily = daily[daily["date"] >= cm["eom"].min()]
    reg = pd.merge(cm, daily[["id", "date", "ret_exc", "eom_ret"]], how="inner", on=["id", "eom_ret"]).dropna()
    reg = reg.sort_values(["date", "id"], kind="stable").reset_index(drop=True)
    dates = np.sort(reg["date"].unique())
    day_ptr = np.concatenate([[0], np.cumsum(reg.groupby("date", sort=True).size().to_numpy())]).astype(np.int64)

    # ---- a15: daily cross-sectional OLS -----------------------------------------------------------------
    fct_ret, resid, _status = daily_ols(reg[factor_cols].to_numpy(dtype=np.float64), reg["ret_exc"].to_numpy(dtype=np.float64),
                                        day_ptr, device)
    spec = pd.DataFrame({"id": reg["id"].to_numpy(), "date": reg["date"].to_numpy(), "residual": resid})
    spec = spec.sort_values(["id", "date"], kind="stable").reset_index(drop=True)
    fct_dates = pd.Series(dates)

    # ---- a16: EWMA factor covariance per calc date (:280-335) -----------------------------------------
    obs = int(cs["obs"])
    ref_date = fct_dates.iloc[obs]
    min_date = ref_date.to_period("M").to_timestamp() - pd.offsets.MonthEnd(1)
    calc_dates = sorted(cm.loc[cm["eom"] >= min_date, "eom"].unique())
    ends = np.searchsorted(dates, np.array(calc_dates, dtype=dates.dtype), side="right") - 1
    fcov = factor_cov(fct_ret, ends, obs=obs, hl_cor=cs["hl_cor"], hl_var=cs["hl_var"], scale=21.0, device=device)

    # ---- a17: EWMA idiosyncratic volatility per stock (:345-405) ---------------------------------------
    sizes = spec.groupby("id", sort=True).size().to_numpy()
    bounds = np.concatenate([[0], np.cumsum(sizes)])
    r = spec["residual"].to_numpy()
    vols = ewma_vol_batch([r[bounds[i]:bounds[i + 1]] for i in range(len(sizes))], 0.5 ** (1 / cs["hl_stock_var"]),
                          int(cs["initial_var_obs"]), device)
    spec["res_vol"] = np.concatenate(vols) if len(vols) else np.zeros(0)

    # ---- a18: 200-of-252 coverage rule, month-end pick (:409-439) --------------------------------------
    td = pd.DataFrame({"date": fct_dates, "td_252d": fct_dates.shift(252)})
    spec = spec.merge(td, on="date", how="left")
    spec["date_200d"] = spec.groupby("id")["date"].shift(200)
    spec = spec[(spec["date_200d"] >= spec["td_252d"]) & spec["res_vol"].notna()][["id", "date", "res_vol"]]
    spec["eom"] = spec["date"] + pd.offsets.MonthEnd(0)
    spec_m = spec[spec["date"] == spec.groupby(["id", "eom"])["date"].transform("max")][["id", "eom", "res_vol"]]

    # ---- assembly with size-group median imputation, x21 monthly scaling (:453-494) -----------------
    barra = {}
    for c, d in enumerate(calc_dates):
        cd = cm[cm["eom"] == d].merge(spec_m, on=["id", "eom"], how="left")
        med = cd.groupby("size_grp")["res_vol"].transform("median")
        med = med.fillna(cd["res_vol"].median())
        cd["res_vol"] = cd["res_vol"].fillna(med)
        cd = cd.sort_values("id")
        ids = cd["id"].astype(int).to_numpy()
        barra[pd.Timestamp(d)] = {
            "fct_load": pd.DataFrame(cd[factor_cols].to_numpy(), index=pd.Index(ids, name="id"), columns=factor_cols),
            "fct_cov": pd.DataFrame(fcov[c], index=factor_cols, columns=factor_cols),
            "ivol_vec": pd.Series((cd["res_vol"] ** 2 * 21).to_numpy(), index=pd.Index(ids, name="id"), name="res_vol"),
        }
    with open(os.path.join(data, "Barra_Cov.pkl"), "wb") as f:
        pickle.dump(barra, f)
    return barra
