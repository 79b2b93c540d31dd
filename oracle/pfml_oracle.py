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
