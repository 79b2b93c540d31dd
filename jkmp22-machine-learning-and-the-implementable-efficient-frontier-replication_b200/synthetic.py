"""Synthetic S&P-500-like panels for the PFML hot path (host code, numpy only).

The reference ships no data (CRSP/JKP are proprietary), so every test, the oracle and
``bench.py`` run on panels drawn here.  Value ranges follow SURVEY.md section 8(d) "C1..C5":
characteristics are cross-sectional percentile ranks in [0, 1] (Prepare_Data.py:338-345,368),
Kyle's lambda = 0.2 / dolvol (Prepare_Data.py:180), wealth is 1e10 at the end of the sample and
back-cumulated (General_functions.py:205), and the Barra triple is
(fct_load [one-hot FF12 | 13 z-scored cluster ranks], fct_cov, ivol_vec) as written by
"Estimate Covariance Matrix.py":490-494.

Everything is dense: month axis ``t`` (0 .. T_all-1, calendar month-ends), stock axis ``i``
(ids ascending).  A ragged universe is expressed through ``valid[t, i]``; the
reference needs a row for every id of date d in all 13 look-back months
(PFML_Input_Data.py:357-361), which a dense panel gives by construction.
"""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np

N_FEATURES = 115          # len(get_features(True)), General_functions.py:113-170
N_IND = 12                # FF12 one-hot industries
N_CLUSTER = 13            # characteristic clusters
LB = 11                   # pf_set['lb_hor'], General_functions.py:107


@dataclass
class Panel:
    """Dense synthetic panel.  Shapes: T = months in panel, N = stocks, F = characteristics, K = factors."""
    ids: np.ndarray            # [N] int64 ascending
    month_ends: np.ndarray     # [T] datetime64[D] calendar month-ends
    feats: np.ndarray          # [T, N, F] in [0, 1]
    valid: np.ndarray          # [T, N] bool
    ret_ld1: np.ndarray        # [T, N]
    tr_ld0: np.ndarray         # [T, N]
    mu_ld0: np.ndarray         # [T, N] (same value for every stock of a month)
    lam: np.ndarray            # [T, N] Kyle's lambda
    wealth: np.ndarray         # [T]
    rf: np.ndarray             # [T] decimal monthly risk-free rate
    fct_load: np.ndarray       # [T, N, K]
    fct_cov: np.ndarray        # [T, K, K]
    ivol: np.ndarray           # [T, N]
    barra_has: np.ndarray      # [T, N] bool: id present in Barra_Cov[d]
    W: np.ndarray              # [F, p/2] random-Fourier-feature frequencies (an INPUT, quirk Q1)
    gamma_rel: float = 10.0
    mu: float = 0.007
    lb: int = LB
    iterations: int = 10
    meta: dict = field(default_factory=dict)
    me: np.ndarray = None      # [T, N] market equity (initial value weights of the back-test, PFML_best_hps.py:143-147)
    tr_ld1: np.ndarray = None  # [T, N] lead total return (weight drift, PFML_best_hps.py:199)
    mu_ld1: np.ndarray = None  # [T]    lead market return (wealth_processed.csv 'mu_ld1')

    def __post_init__(self):
        T, N = self.valid.shape
        if self.me is None:
            self.me = np.ones((T, N))
        if self.tr_ld1 is None:
            self.tr_ld1 = np.zeros((T, N))
        if self.mu_ld1 is None:
            self.mu_ld1 = np.zeros(T)

    @property
    def T(self):
        return self.feats.shape[0]

    @property
    def N(self):
        return self.feats.shape[1]

    @property
    def p(self):
        return 2 * self.W.shape[1]

    def hot_months(self):
        """Month indices of dates_m2 (PFML_Input_Data.py:138-142): first_cov + lb+1 months .. last-1."""
        return np.arange(self.lb + 1, self.T - 1)


def month_end_range(start="2010-01", periods=24):
    m = np.datetime64(start, "M") + np.arange(periods)
    return (m + 1).astype("datetime64[D]") - np.timedelta64(1, "D")


def make_panel(N=500, T=24, p=64, g=np.exp(-3.0), seed=1, F=N_FEATURES, ragged=0.0,
               barra_missing=0.0, start="2010-01", nan_tr=0.0, quantize=False) -> Panel:
    """Draw a panel.  ``ragged``: probability that a stock is invalid in a month (the universe then
    differs by date); ``barra_missing``: probability that an id is absent from the Barra object of a
    month (exercises the eom-median fill of vol_scale, PFML_Input_Data.py:300-305; only applied to
    rows that are not in the date-d universe because create_cov(ids) needs them, General_functions.py:865);
    ``nan_tr``: probability of a missing tr_ld0 in look-back months (quirk Q8)."""
    rng = np.random.default_rng(seed)
    K = N_IND + N_CLUSTER
    ids = 10001 + np.arange(N, dtype=np.int64)
    feats = rng.random((T, N, F))
    if quantize:                     # 3-decimal ranks: golden files can store them as uint16
        feats = np.round(feats * 1000.0) / 1000.0
    valid = np.ones((T, N), dtype=bool)
    if ragged > 0:
        # a stock drops out for a stretch; validity needs all 13 look-back rows, which the dense panel has
        valid = rng.random((T, N)) >= ragged
    dolvol = np.exp(np.log(1e8) + rng.standard_normal(N))[None, :] * np.exp(0.1 * rng.standard_normal((T, N)))
    lam = 0.2 / dolvol
    tr = np.clip(0.01 + 0.08 * rng.standard_normal((T, N)), -0.9, None)
    mu_m = 0.008 + 0.04 * rng.standard_normal(T)
    mu_ld0 = np.repeat(mu_m[:, None], N, axis=1)
    if nan_tr > 0:
        mask = rng.random((T, N)) < nan_tr
        tr = np.where(mask, np.nan, tr)
    ret_ld1 = 0.006 + 0.08 * rng.standard_normal((T, N))
    rf = rng.uniform(0.1, 0.4, T) / 100.0
    # wealth: 1e10 at the end, back-cumulated with the market's total return (General_functions.py:205)
    wealth = np.empty(T)
    wealth[-1] = 1e10
    for t in range(T - 2, -1, -1):
        wealth[t] = wealth[t + 1] / (1.0 + mu_m[t + 1] + rf[t + 1])
    # Barra triple
    ind = rng.integers(0, N_IND, N)
    onehot = np.zeros((N, N_IND))
    onehot[np.arange(N), ind] = 1.0
    fct_load = np.empty((T, N, K))
    fct_cov = np.empty((T, K, K))
    for t in range(T):
        z = rng.standard_normal((N, N_CLUSTER))
        z = (z - z.mean(0)) / z.std(0, ddof=1)
        fct_load[t] = np.concatenate([onehot, z], axis=1)
        A = rng.standard_normal((K, K))
        fct_cov[t] = 21.0 * 0.005 ** 2 * (A @ A.T) / K
    ivol = 21.0 * rng.uniform(0.008, 0.03, (T, N)) ** 2
    barra_has = np.ones((T, N), dtype=bool)
    if barra_missing > 0:
        barra_has = rng.random((T, N)) >= barra_missing
        barra_has |= valid          # create_cov(cov_list[d], ids) needs every universe id
        # look-back rows of later universes may still be missing -> median fill path
    W = np.sqrt(g) * rng.standard_normal((F, p // 2))
    rng2 = np.random.default_rng(seed + 100003)          # extra fields drawn from a separate stream
    me = np.exp(np.log(1e9) + 1.5 * rng2.standard_normal(N))[None, :] * np.exp(0.05 * rng2.standard_normal((T, N)))
    tr_ld1 = np.clip(0.008 + 0.08 * rng2.standard_normal((T, N)), -0.9, None)
    mu_ld1 = 0.008 + 0.04 * rng2.standard_normal(T)
    return Panel(me=me, tr_ld1=tr_ld1, mu_ld1=mu_ld1, ids=ids, month_ends=month_end_range(start, T), feats=feats, valid=valid, ret_ld1=ret_ld1,
                 tr_ld0=tr, mu_ld0=mu_ld0, lam=lam, wealth=wealth, rf=rf, fct_load=fct_load, fct_cov=fct_cov,
                 ivol=ivol, barra_has=barra_has, W=W, meta=dict(seed=seed, g=float(g)))


def make_daily(N=60, years=3.2, seed=3, start="2015-01-01", F=N_FEATURES, miss=0.02):
    """Daily panel for the Barra/EWMA covariance estimator ("Estimate Covariance Matrix.py"), SURVEY config C2.

    Returns a dict of dense arrays: business days [D], month-ends [Tm] (one lagged month in front), ids [N],
    monthly characteristics feats [Tm,N,F] in [0,1], valid [Tm,N], direction [F] (+-1), cluster_of [F] (0..12),
    ff12 [N] (1..12), size_grp [N], daily excess returns ret [D,N] = X f + e with NaN where unobserved."""
    rng = np.random.default_rng(seed)
    days = np.arange(np.datetime64(start, "D"), np.datetime64(start, "D") + np.timedelta64(int(365.25 * years), "D"))
    dow = (days.view("int64") - 4) % 7                                # 1970-01-01 is a Thursday; Monday = 0
    days = days[dow < 5]
    months = np.unique(days.astype("datetime64[M]"))
    mlist = np.concatenate([[months[0] - 1], months])
    month_ends = (mlist + 1).astype("datetime64[D]") - np.timedelta64(1, "D")
    Tm, D = len(month_ends), len(days)
    ids = 10001 + np.arange(N, dtype=np.int64)
    feats = np.round(rng.random((Tm, N, F)) * 1000.0) / 1000.0      # 3-decimal ranks: golden files store them as uint16
    valid = rng.random((Tm, N)) >= 0.03
    direction = rng.choice([-1, 1], F)
    cluster_of = np.arange(F) % N_CLUSTER
    rng.shuffle(cluster_of)
    ff12 = (np.arange(N) % N_IND) + 1
    rng.shuffle(ff12)
    size_grp = rng.choice(np.array(["mega", "large", "small"]), N)
    K = N_IND + N_CLUSTER
    A = rng.standard_normal((K, K))
    f = 0.005 * rng.standard_normal((D, K)) @ (A / np.sqrt(K)).T
    sig = rng.uniform(0.008, 0.03, N)
    load = np.concatenate([np.eye(N_IND)[ff12 - 1], rng.standard_normal((N, N_CLUSTER))], axis=1)
    ret = f @ load.T + sig[None, :] * rng.standard_normal((D, N))
    ret[rng.random((D, N)) < miss] = np.nan
    late = rng.integers(0, N)                                          # one stock lists late: coverage filter + median fill
    ret[: D // 2, late] = np.nan
    return dict(days=days, month_ends=month_ends, ids=ids, feats=feats, valid=valid, direction=direction,
                cluster_of=cluster_of, ff12=ff12, size_grp=size_grp, ret=ret, seed=seed)


def write_covariance_artefacts(daily, path, features=None):
    """Files "Estimate Covariance Matrix.py" reads: Factors_processed (JKP_US_SP500.db), d_ret_ex
    (crsp_daily_SP500.db), cluster_labels_processed.csv (:40-106)."""
    import os
    import sqlite3

    import pandas as pd
    if features is None:
        from .general_functions import get_features
        features = get_features(True)
    data = os.path.join(path, "Data")
    os.makedirs(data, exist_ok=True)
    Tm, N, F = daily["feats"].shape
    me = pd.to_datetime(daily["month_ends"])
    rows = {
        "id": np.tile(daily["ids"], Tm), "eom": np.repeat(me.strftime("%Y-%m-%d 00:00:00").values, N),
        "sic": 0, "ff49": 0, "size_grp": np.tile(daily["size_grp"], Tm), "me": 1.0, "crsp_exchcd": 1, "dolvol": 1.0,
        "lambda": 1e-8, "rvol_m": 0.0, "tr_ld0": 0.0,
        "eom_ret": np.repeat((me + pd.offsets.MonthEnd(1)).strftime("%Y-%m-%d 00:00:00").values, N),
        "ret_ld1": 0.0, "tr_ld1": 0.0, "mu_ld0": 0.0, "ff12": np.tile(daily["ff12"].astype(str), Tm),
        "valid": daily["valid"].reshape(-1).astype(int),
    }
    df = pd.concat([pd.DataFrame(rows), pd.DataFrame(daily["feats"].reshape(Tm * N, F), columns=features)], axis=1)
    con = sqlite3.connect(os.path.join(data, "JKP_US_SP500.db"))
    df.to_sql("Factors_processed", con, if_exists="replace", index=False)
    con.close()
    D = len(daily["days"])
    dd = pd.DataFrame({"permno": np.tile(daily["ids"], D),
                       "date": np.repeat(pd.to_datetime(daily["days"]).strftime("%Y-%m-%d 00:00:00").values, N),
                       "ret_excess": daily["ret"].reshape(-1)})
    con = sqlite3.connect(os.path.join(data, "crsp_daily_SP500.db"))
    dd.to_sql("d_ret_ex", con, if_exists="replace", index=False)
    con.close()
    pd.DataFrame({"characteristic": features, "direction": daily["direction"],
                  "cluster": [f"cl{c:02d}" for c in daily["cluster_of"]]}).to_csv(
        os.path.join(data, "cluster_labels_processed.csv"), index=False)


def write_reference_artefacts(panel, path, features=None):
    """Write the files the reference scripts read under ``path + "Data/"``: Factors_processed
    (SQLite), FF_RF_monthly.csv, wealth_processed.csv, Barra_Cov.pkl, rff_w.csv."""
    import os
    import pickle
    import sqlite3

    import pandas as pd
    if features is None:
        from .general_functions import get_features
        features = get_features(True)
    data = os.path.join(path, "Data")
    os.makedirs(data, exist_ok=True)
    T, N, F = panel.feats.shape
    assert F == len(features)
    me = pd.to_datetime(panel.month_ends)
    rows = {
        "id": np.tile(panel.ids, T),
        "eom": np.repeat(me.strftime("%Y-%m-%d 00:00:00").values, N),
        "sic": 0, "ff49": 0, "size_grp": "large", "me": panel.me.reshape(-1), "crsp_exchcd": 1, "dolvol": 1.0,
        "lambda": panel.lam.reshape(-1), "rvol_m": 0.0, "tr_ld0": panel.tr_ld0.reshape(-1),
        "eom_ret": np.repeat((me + pd.offsets.MonthEnd(1)).strftime("%Y-%m-%d 00:00:00").values, N),
        "ret_ld1": panel.ret_ld1.reshape(-1), "tr_ld1": panel.tr_ld1.reshape(-1), "mu_ld0": panel.mu_ld0.reshape(-1),
        "ff12": "1", "valid": panel.valid.reshape(-1).astype(int),
    }
    df = pd.concat([pd.DataFrame(rows), pd.DataFrame(panel.feats.reshape(T * N, F), columns=features)], axis=1)
    con = sqlite3.connect(os.path.join(data, "JKP_US_SP500.db"))
    df.to_sql("Factors_processed", con, if_exists="replace", index=False)
    con.close()
    pd.DataFrame({"yyyymm": me.strftime("%Y%m").astype(int), "RF": panel.rf * 100.0}).to_csv(
        os.path.join(data, "FF_RF_monthly.csv"), index=False)
    pd.DataFrame({"eom": me, "wealth": panel.wealth, "mu_ld1": panel.mu_ld1}).to_csv(
        os.path.join(data, "wealth_processed.csv"), index=False)
    names = [f"f{k}" for k in range(panel.fct_cov.shape[1])]
    barra = {}
    for t in range(T):
        has = panel.barra_has[t]
        ids = panel.ids[has].astype(int)
        barra[me[t]] = {"fct_load": pd.DataFrame(panel.fct_load[t][has], index=ids, columns=names),
                        "fct_cov": pd.DataFrame(panel.fct_cov[t], index=names, columns=names),
                        "ivol_vec": pd.Series(panel.ivol[t][has], index=ids)}
    with open(os.path.join(data, "Barra_Cov.pkl"), "wb") as f:
        pickle.dump(barra, f)
    pd.DataFrame(panel.W).to_csv(os.path.join(data, "rff_w.csv"))      # index column -> 'Unnamed: 0'
