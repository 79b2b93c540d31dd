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
