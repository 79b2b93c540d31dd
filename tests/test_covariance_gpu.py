"""GPU: covariance-estimator kernels against the oracle and the reference's own functions (golden vectors).
Tolerance 1e-8 relative (observed 1e-13); ewma_vol follows the reference's operation order and is held to 1e-15."""
import os

import numpy as np
import pytest

import pfml_oracle as O
from conftest import GOLD, pkg, relerr

pytestmark = pytest.mark.gpu


def test_weighted_cov_cor_golden():
    ec = pkg("estimate_covariance")
    f = np.load(os.path.join(GOLD, "funcs_ref.npz"))
    assert relerr(ec.weighted_cov_wt(f["wc_X"], f["wc_w"]), f["cov_ref"]) < 1e-10
    cor = ec.weighted_cor_wt(f["wc_X"], f["wc_w"])
    assert relerr(cor, f["cor_ref"]) < 1e-10
    assert (np.diag(cor) == 1.0).all()


def test_factor_cov_windows_vs_oracle():
    ec = pkg("estimate_covariance")
    rng = np.random.default_rng(8)
    D, K, obs = 900, 25, 600
    A = rng.standard_normal((K, K))
    F = 0.005 * rng.standard_normal((D, K)) @ A.T / np.sqrt(K)
    ends = [40, 300, 599, 600, 899]                      # short window, exactly obs, longer than obs
    w_c, w_v = ec.ewma_weights(obs, 378), ec.ewma_weights(obs, 126)
    got = ec.factor_cov(F, ends, obs=obs, hl_cor=378, hl_var=126, scale=21.0)
    for k, e in enumerate(ends):
        win = F[max(0, e + 1 - obs):e + 1]
        ref = O.factor_cov(win, w_c, w_v) * 21.0
        assert relerr(got[k], ref) < 1e-10, e
        assert relerr(got[k], got[k].T) < 1e-15


def test_daily_ols_vs_oracle_with_empty_industry():
    ec = pkg("estimate_covariance")
    rng = np.random.default_rng(9)
    K, D = 25, 40
    rows, ptr = [], [0]
    Xs, ys = [], []
    for d in range(D):
        n = int(rng.integers(60, 140))
        ind = rng.integers(0, 12, n)
        if d == 7:
            ind[ind == 3] = 4                            # industry 3 has no stock that day -> singular X'X
        X = np.concatenate([np.eye(12)[ind], rng.standard_normal((n, 13))], axis=1)
        y = X @ (0.01 * rng.standard_normal(K)) + 0.02 * rng.standard_normal(n)
        Xs.append(X); ys.append(y); ptr.append(ptr[-1] + n)
    X, y = np.concatenate(Xs), np.concatenate(ys)
    coef, resid, status = ec.daily_ols(X, y, ptr)
    for d in range(D):
        c_ref, r_ref = O.daily_ols(Xs[d], ys[d])
        if d == 7:
            assert status[d] == 1
            c_ref = np.linalg.pinv(Xs[d].T @ Xs[d]) @ (Xs[d].T @ ys[d])     # what the reference's fallback returns
            r_ref = ys[d] - Xs[d] @ c_ref
            assert coef[d][3] == 0.0
        else:
            assert status[d] == 0
        assert relerr(coef[d], c_ref) < 1e-9, d
        assert relerr(resid[ptr[d]:ptr[d + 1]], r_ref) < 1e-9, d


def test_ewma_vol_golden_and_edge_cases():
    ec = pkg("estimate_covariance")
    f = np.load(os.path.join(GOLD, "funcs_ref.npz"))
    x, lam, ref = f["ev_x"], float(f["ev_lam"]), f["ev_ref"]
    rng = np.random.default_rng(10)
    short = rng.standard_normal(50)                      # n <= start -> all NaN
    allnan = np.full(100, np.nan)                        # count <= 1 -> all NaN
    other = rng.standard_normal(300) * 0.01
    out = ec.ewma_vol_batch([x, short, allnan, other], lam, 63)
    assert (np.isnan(out[0]) == np.isnan(ref)).all()
    assert np.nanmax(np.abs(out[0] - ref) / ref) < 1e-15
    assert np.isnan(out[1]).all() and np.isnan(out[2]).all()
    o3 = O.ewma_vol(other, lam, 63)
    assert np.nanmax(np.abs(out[3] - o3) / o3) < 1e-15


def test_ewma_wrapper_mirror():
    """Same-named drop-in of "Estimate Covariance Matrix.py":389-397: Series in, Series out, settings-driven lam / start."""
    import pandas as pd
    ec = pkg("estimate_covariance")
    GF = pkg("general_functions")
    settings, _ = GF.get_settings()
    rng = np.random.default_rng(12)
    s = pd.Series(0.01 * rng.standard_normal(400), index=pd.date_range("2020-01-01", periods=400, freq="B"))
    s.iloc[[70, 71, 200]] = np.nan
    got = ec.ewma_wrapper(s, settings)
    ref = O.ewma_vol(s.values.astype(float), 0.5 ** (1 / settings["cov_set"]["hl_stock_var"]), settings["cov_set"]["initial_var_obs"])
    assert got.index.equals(s.index)
    assert np.allclose(got.to_numpy(), ref, rtol=1e-13, atol=0, equal_nan=True)
