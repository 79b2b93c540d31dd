"""CPU ORACLE for the PFML hot path -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

A plain numpy/scipy restatement of the reference's algorithm for the path SURVEY.md section 8 scopes
(rows a1..a13), operation by operation and in the reference's own evaluation order (dense Sigma,
dense diag(Lambda), the 22+24 GEMM chain as written, LU solves).  Each function cites the reference
file:line it follows.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
reference legs may import this module; the product package never does.

Parity status: the reference ships no tests, fixtures or golden vectors, so the
oracle is pinned against the reference ITSELF: oracle/make_golden.py exec's the unmodified reference
scripts from /root/reference on a synthetic panel in the build container (numpy 2.3.5, scipy 1.18.1,
pandas 3.0.2) and freezes their outputs under tests/golden/; tests/test_oracle_golden.py checks this
restatement against those vectors.

Third-party arithmetic the reference delegates to (not under /root/reference, versions unpinned
there): numpy/OpenBLAS (dgemm, dgesv, dgetri), scipy.linalg.sqrtm (Schur method).
"""
from __future__ import annotations

import numpy as np
from scipy.linalg import sqrtm


# ----------------------------------------------------------------------------------------------
# a1: random Fourier features
# ----------------------------------------------------------------------------------------------
def rff(X, W):
    """PFML_Input_Data.py:159-185 with W given (quirk Q1: W is always read from rff_w.csv, :245).
    Returns [cos | sin] concatenated as at :254."""
    X_new = np.dot(X, W)                                     # :179
    return np.concatenate([np.cos(X_new), np.sin(X_new)], axis=1)   # :183-184, :254


def pfml_feat_idx(p, p_max):
    """Integer form of pfml_feat_fun (General_functions.py:837-844) inside the feat_all = pfml_feat_fun(p_max)
    layout (PFML_Input_Data.py:221): constant, cos 1..p/2, sin 1..p/2."""
    h, hm = p // 2, p_max // 2
    return np.concatenate([[0], 1 + np.arange(h), 1 + hm + np.arange(h)]).astype(np.int64)


# ----------------------------------------------------------------------------------------------
# a3/a4/a5: Sigma, Lambda, m
# ----------------------------------------------------------------------------------------------
def create_cov(load, fct_cov, ivol):
    """General_functions.py:869: sigma = load @ fct_cov @ load.T + diag(ivol)."""
    return load @ fct_cov @ load.T + np.diag(ivol)


def create_lambda(lam):
    """General_functions.py:900-916."""
    return np.diag(lam)


def m_func(w, mu, rf, sigma_gam, gam, K_Lambda, iterations):
    """General_functions.py:919-962, line by line."""
    sigma_gam = np.asarray(sigma_gam)
    K_Lambda = np.asarray(K_Lambda)
    n = K_Lambda.shape[0]
    g_bar = np.ones(n)
    mu_bar_vec = np.full(n, 1 + rf + mu)
    sigma_gr = (1 / (1 + rf + mu) ** 2) * (np.outer(mu_bar_vec, mu_bar_vec) + sigma_gam / gam)   # :944
    K_Lambda_neg05 = np.diag(np.diag(K_Lambda) ** (-0.5))                                         # :948
    x = (w ** -1) * K_Lambda_neg05 @ sigma_gam @ K_Lambda_neg05                                   # :951
    y = np.diag(1 + np.diag(sigma_gr))                                                            # :952
    sigma_hat = x + np.diag(1 + g_bar)                                                            # :955
    m_tilde = 0.5 * (sigma_hat - np.real(sqrtm(sigma_hat @ sigma_hat - 4 * np.eye(n))))           # :956
    for _ in range(iterations):                                                                   # :959-960
        m_tilde = np.linalg.inv(x + y - m_tilde * sigma_gr)
    return K_Lambda_neg05 @ m_tilde @ np.sqrt(K_Lambda)                                           # :962


# ----------------------------------------------------------------------------------------------
# a2: vol scale with eom-median fill
# ----------------------------------------------------------------------------------------------
def vol_scale_panel(panel):
    """PFML_Input_Data.py:274-305.  sqrt(diag(create_cov(cov_list[d]))) for every id present in the Barra
    object of month d (:278-281), left-merged on (id, eom) (:300) and NaN-filled by the per-eom median
    over all rows, skipna (:303-305)."""
    T, N = panel.valid.shape
    vs = np.full((T, N), np.nan)
    for t in range(T):
        has = panel.barra_has[t]
        load = panel.fct_load[t][has]
        sigma = create_cov(load, panel.fct_cov[t], panel.ivol[t][has])
        vs[t, has] = np.sqrt(np.diag(sigma))
        if (~has).any():
            vs[t, ~has] = np.nanmedian(vs[t])
    return vs


# ----------------------------------------------------------------------------------------------
# a6..a9: one date-step
# ----------------------------------------------------------------------------------------------
def date_step(panel, t, rff_x=None, vol_scale=None, keep=False):
    """One iteration of the loop PFML_Input_Data.py:318-491 for month index ``t``.

    ``rff_x`` [T, N, p] and ``vol_scale`` [T, N] are the panel-level precomputations (:245-256, :274-305);
    they are computed here if omitted.  Output columns are in feat_all order (constant first, :484,:489);
    the computation itself runs in feat_cons order (constant last, :262-263) like the reference.
    """
    lb = panel.lb
    if rff_x is None:
        T, N, F = panel.feats.shape
        rff_x = rff(panel.feats.reshape(T * N, F), panel.W).reshape(T, N, -1)
    if vol_scale is None:
        vol_scale = vol_scale_panel(panel)
    p = rff_x.shape[2]
    sel = np.flatnonzero(panel.valid[t])                     # :321-324 (ids ascending)
    n = len(sel)
    r = panel.ret_ld1[t, sel]                                # :330
    sigma = create_cov(panel.fct_load[t][sel], panel.fct_cov[t], panel.ivol[t, sel])    # :333
    K_Lambda = create_lambda(panel.lam[t, sel])              # :336
    w = panel.wealth[t]                                      # :339
    rf = panel.rf[t]                                         # :342
    gam = panel.gamma_rel
    m = m_func(w=w, mu=panel.mu, rf=rf, sigma_gam=sigma * gam, gam=gam, K_Lambda=K_Lambda,
               iterations=panel.iterations)                  # :345

    # standardise per (d, month) over the date-d universe, :357-369; key tau = months back from d
    signals = {}
    gtm = {}
    for tau in range(lb + 2):
        tt = t - tau
        x = rff_x[tt, sel, :]                                # rows of month tt for ids of date d
        x = x - x.mean(axis=0)                               # :368 (constant is not de-meaned)
        s = np.concatenate([x, np.ones((n, 1))], axis=1)     # feat_cons order: RFFs then constant, :263,:364
        s = s * np.sqrt(1 / (s ** 2).sum(axis=0))            # :369
        scales_mat = np.diag(1 / vol_scale[tt, sel])         # :384
        signals[tau] = scales_mat @ s                        # :386
        gt = (1 + panel.tr_ld0[tt, sel]) / (1 + panel.mu_ld0[tt, sel])   # :404
        gt = np.nan_to_num(gt, nan=1.0)                      # :405
        gtm[tau] = m @ np.diag(gt)                           # :406

    gtm_agg = {0: np.eye(n)}                                 # :420
    gtm_agg_l1 = {0: np.eye(n)}                              # :422
    for i in range(lb):                                      # :424-429
        gtm_agg[i + 1] = gtm_agg[i] @ gtm[i]
        gtm_agg_l1[i + 1] = gtm_agg_l1[i] @ gtm[i + 1]

    omega = 0
    const = 0
    omega_l1 = 0
    const_l1 = 0
    for theta in range(lb + 1):                              # :443-453
        omega = omega + gtm_agg[theta] @ signals[theta]
        const = const + gtm_agg[theta]
        omega_l1 = omega_l1 + gtm_agg_l1[theta] @ signals[theta + 1]
        const_l1 = const_l1 + gtm_agg_l1[theta]
    omega = np.linalg.solve(const, omega)                    # :455
    omega_l1 = np.linalg.solve(const_l1, omega_l1)           # :456
    gt0 = np.diag((1 + panel.tr_ld0[t, sel]) / (1 + panel.mu_ld0[t, sel]))     # :458 (no NaN->1, quirk Q8)
    omega_chg = omega - gt0 @ omega_l1                       # :459

    r_tilde = omega.T @ r                                    # :470
    risk = gam * omega.T @ sigma @ omega                     # :473
    tc = w * omega_chg.T @ K_Lambda @ omega_chg              # :477
    denom = risk + tc                                        # :481

    # re-index feat_cons -> feat_all (:484, :489): constant (last) moves to the front
    perm = np.concatenate([[p], np.arange(p)])
    out = dict(r_tilde=r_tilde[perm], denom=denom[np.ix_(perm, perm)], risk=risk[np.ix_(perm, perm)],
               tc=tc[np.ix_(perm, perm)], signal_t=signals[0][:, perm], ids=panel.ids[sel])
    if keep:
        out.update(m=m, omega=omega[:, perm], omega_chg=omega_chg[:, perm], sigma=sigma,
                   signals={k: v[:, perm] for k, v in signals.items()})
    return out


# ----------------------------------------------------------------------------------------------
# a11/a12: expanding-window sums and the ridge grid
# ----------------------------------------------------------------------------------------------
def denom_sum_fun(denoms):
    """PFML_Search_Coef.py:37-46: left-to-right reduce(+)."""
    it = iter(denoms)
    acc = next(it).copy()
    for d in it:
        acc = acc + d
    return acc


def search_coef(r_tilde, denom, month_ends, hp_years, p_vec, l_vec):
    """PFML_Search_Coef.py:68-137 on dense inputs.

    r_tilde [T, P], denom [T, P, P] for the hot dates ``month_ends`` [T] (datetime64[D]); returns
    {year: {p: beta[n_lambda, p+1]}} in pfml_feat_fun(p) order."""
    hp_years = sorted(int(y) for y in hp_years)
    p_max = r_tilde.shape[1] - 1
    end_bef = min(hp_years) - 1
    d_all = np.asarray(month_ends).astype("datetime64[D]")
    mask = d_all < np.datetime64(f"{end_bef - 1}-12-31")     # :72 (strict)
    idx = np.flatnonzero(mask)
    if len(idx) == 0:
        raise ValueError("empty burn-in window (pd.concat([]) raises in the reference, quirk Q13)")
    r_sum = r_tilde[idx].T.sum(axis=1)                       # :79-82 (concat columns, row sums)
    d_sum = denom_sum_fun(denom[i] for i in idx)             # :85
    n = len(idx)                                             # :88
    coef = {}
    for year in hp_years:                                    # :102
        lo = np.datetime64(f"{year - 2}-12-31")              # :105
        hi = np.datetime64(f"{year - 1}-11-30")              # :106
        idx = np.flatnonzero((d_all >= lo) & (d_all <= hi))  # :109
        if len(idx) == 0:
            raise ValueError(f"empty window for year {year} (quirk Q13)")
        n += len(idx)                                        # :112
        r_sum = r_sum + r_tilde[idx].T.sum(axis=1)           # :115-117
        d_sum = d_sum + denom_sum_fun(denom[i] for i in idx)  # :120-121
        by_p = {}
        for p in p_vec:                                      # :125
            fi = pfml_feat_idx(p, p_max)                     # :126
            r_sub = r_sum[fi] / n                            # :127
            d_sub = d_sum[np.ix_(fi, fi)] / n                # :128
            betas = np.empty((len(l_vec), p + 1))
            for j, l in enumerate(l_vec):                    # :131-133
                betas[j] = np.linalg.solve(d_sub + l * np.identity(p + 1), r_sub)
            by_p[p] = betas
        coef[year] = by_p
    return coef


# ----------------------------------------------------------------------------------------------
# a16/a17 (+a15): covariance estimator pieces
# ----------------------------------------------------------------------------------------------
def weighted_cov_wt(X, weights):
    """General_functions.py:745-784 (R cov.wt, center=TRUE, unbiased)."""
    w = np.asarray(weights, dtype=float)
    w_norm = w / w.sum()                                     # :766
    mu = np.average(X, axis=0, weights=w_norm)               # :769
    Xc = X - mu                                              # :772
    Xw = Xc * np.sqrt(w_norm)[:, np.newaxis]                 # :775
    denom = 1.0 - np.sum(w_norm ** 2)                        # :778
    return (Xw.T @ Xw) / denom                               # :779


def weighted_cor_wt(X, weights):
    """General_functions.py:786-835."""
    cov = weighted_cov_wt(X, weights)                        # :803-818
    sd = np.sqrt(np.diag(cov))                               # :821
    cor = cov / np.outer(sd, sd)                             # :824-827
    np.fill_diagonal(cor, 1.0)                               # :830
    return cor


def factor_cov(fct_ret_window, w_cor, w_var):
    """"Estimate Covariance Matrix.py":306-320: sd_diag @ cor @ sd_diag for one calc date."""
    t = len(fct_ret_window)
    cor = weighted_cor_wt(fct_ret_window, w_cor[-t:])        # :311
    var = weighted_cov_wt(fct_ret_window, w_var[-t:])        # :313
    sd = np.diag(np.sqrt(np.diag(var)))                      # :316
    return sd @ cor @ sd                                     # :319


def ewma_vol(x, lam, start):
    """"Estimate Covariance Matrix.py":345-386 (the numba kernel), in plain Python/numpy."""
    x = np.asarray(x, dtype=np.float64)
    n = len(x)
    vol = np.full(n, np.nan)
    if n <= start:                                           # :362
        return vol
    valid_x = x[:start]
    mask = ~np.isnan(valid_x)
    count = int(mask.sum())
    if count <= 1:                                           # :369
        return vol
    var = np.empty(n)
    var[:start] = np.nan
    var[start] = np.sum(valid_x[mask] ** 2) / (count - 1)    # :372-375
    for i in range(start + 1, n):                            # :378-383
        if np.isnan(x[i - 1]):
            var[i] = var[i - 1]
        else:
            var[i] = lam * var[i - 1] + (1 - lam) * x[i - 1] ** 2
    return np.sqrt(var)                                      # :385


def daily_ols(X, y):
    """"Estimate Covariance Matrix.py":224-233 for one day: coef, residuals (pinv fallback :229)."""
    XtX = X.T @ X
    Xty = X.T @ y
    try:
        coef = np.linalg.solve(XtX, Xty)
    except np.linalg.LinAlgError:
        coef = np.linalg.pinv(XtX) @ Xty
    return coef, y - X @ coef


# ----------------------------------------------------------------------------------------------
# f1: validation utility (PFML_hp_reals.py), f2: aim portfolio (PFML_aim_fun.py), f3: back-test (PFML_best_hps.py)
# ----------------------------------------------------------------------------------------------
def hp_reals(r_tilde, denom, month_ends, coef, hp_years, p_vec, l_vec):
    """PFML_hp_reals.py:73-122 for one g.  Returns a dict of equal-length arrays sorted by (p, l, eom_ret):
    eom, eom_ret (datetime64[D]), obj, l, p, hp_end, cum_obj, rank."""
    import pandas as pd
    d_all = np.asarray(month_ends).astype("datetime64[D]")
    p_max = r_tilde.shape[1] - 1
    rows = []
    for end in sorted(int(y) for y in hp_years):                                    # :73
        idx = np.flatnonzero((d_all >= np.datetime64(f"{end - 1}-12-31")) & (d_all <= np.datetime64(f"{end}-11-30")))   # :76
        for p in p_vec:                                                             # :81
            fi = pfml_feat_idx(p, p_max)
            for i in range(len(l_vec)):                                             # :88
                b = coef[end][p][i]
                for t in idx:                                                       # :91
                    rt, dn = r_tilde[t][fi], denom[t][np.ix_(fi, fi)]
                    r = rt.T @ b - 0.5 * b.T @ dn @ b                               # :95
                    rows.append((d_all[t], r, i, p, end))
    df = pd.DataFrame(rows, columns=["eom", "obj", "l", "p", "hp_end"])
    df["eom"] = pd.to_datetime(df["eom"])
    df["eom_ret"] = df["eom"] + pd.offsets.MonthEnd(1)                              # :97
    df = df.sort_values(["p", "l", "eom_ret"])                                      # :106
    df["cum_obj"] = df.groupby(["p", "l"])["obj"].expanding().mean().reset_index(level=[0, 1], drop=True)   # :109-115
    df["rank"] = df.groupby("eom_ret")["cum_obj"].rank(ascending=False, method="dense")                  # :118-122
    return df.reset_index(drop=True)


def aim_portfolios(validation, coef, signal_t, month_ends, dates_oos_idx, ids_by_date):
    """PFML_aim_fun.py:126-163 for one g: per OOS date the rank-1 (p, l) of the previous December and
    w_aim = signal_t[d][feat] @ beta.  ``signal_t[t]`` [n_t, P] in feat_all order; returns {t: (p, l, w_aim)}."""
    import pandas as pd
    v = validation
    opt = v[(v["eom_ret"].dt.month == 12) & (v["rank"] == 1)].assign(hp_end=lambda x: x["eom_ret"].dt.year)[["hp_end", "l", "p"]].reset_index(drop=True)
    out = {}
    for t in dates_oos_idx:
        d = pd.Timestamp(month_ends[t])
        oos_year = (d + pd.offsets.MonthEnd(1)).year                                # :136-138
        h = opt[opt["hp_end"] == oos_year - 1]
        p, l = int(h["p"].values[0]), int(h["l"].values[0])                         # first rank-1 row, :144-149
        P = signal_t[t].shape[1]
        fi = pfml_feat_idx(p, P - 1)
        b = coef[oos_year][p][l]
        out[t] = (p, l, signal_t[t][:, fi] @ b)                                     # :152
    return out


def pfml_w(panel, dates_idx, w_aim, vol_scale=None):
    """pfml_w of PFML_best_hps.py:168-216 on the dense panel: value-weighted start at the first date, then
    w = m w_start + (I - m) w_aim (:195), drift w_start' = w (1 + tr_ld1) / (1 + mu_ld1) (:199), new stocks start at 0.
    ``w_aim[t]`` is aligned with the valid ids of date t.  Returns {t: dict(ids, w_start, w)}."""
    out = {}
    prev = None
    for k, t in enumerate(dates_idx):
        sel = np.flatnonzero(panel.valid[t])
        n = len(sel)
        if k == 0:
            w_start = panel.me[t, sel] / panel.me[t, sel].sum()                     # :143-147
        else:
            w_start = np.zeros(n)                                                   # :210
            pos = {int(i): j for j, i in enumerate(prev[0])}
            for j, i in enumerate(sel):
                if int(i) in pos:
                    w_start[j] = prev[1][pos[int(i)]]
            w_start = np.nan_to_num(w_start, nan=0.0)                               # :210: NaN w_start (NaN tr_ld1 / mu_ld1) -> 0
        sigma = create_cov(panel.fct_load[t][sel], panel.fct_cov[t], panel.ivol[t, sel])
        m = m_func(panel.wealth[t], panel.mu, panel.rf[t], sigma * panel.gamma_rel, panel.gamma_rel,
                   create_lambda(panel.lam[t, sel]), panel.iterations)               # :186-190
        w = m @ w_start + (np.eye(n) - m) @ w_aim[t]                                # :195
        nxt = w * (1 + panel.tr_ld1[t, sel]) / (1 + panel.mu_ld1[t])                # :199 (mu_ld1 of wealth at eom d)
        out[t] = dict(ids=panel.ids[sel], w_start=w_start, w=w)
        prev = (sel, nxt)
    return out


def pf_stats(panel, weights):
    """compute_stats / pf_ts_fun of PFML_best_hps.py:220-259 per date: inv, shorting, turnover, r, tc."""
    out = {}
    for t, o in weights.items():
        sel = np.flatnonzero(panel.valid[t])
        w, ws = o["w"], o["w_start"]
        out[t] = dict(inv=np.sum(np.abs(w)), shorting=np.sum(np.abs(w[w < 0])), turnover=np.sum(np.abs(w - ws)),
                      r=np.sum(w * panel.ret_ld1[t, sel]), tc=(panel.wealth[t] / 2) * np.sum(panel.lam[t, sel] * (w - ws) ** 2))
    return out
