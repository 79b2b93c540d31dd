"""The numerical part of PFML_best_hps.py on the GPU: pfml_w and pf_ts_fun.

``pfml_w`` (:168-216): one batched trading-speed computation for all out-of-sample months (the reference calls
m_func once per month, :190) and one kernel that walks the months in order,
w = m w_start + (I - m) w_aim, w_start' = w (1 + tr_ld1) / (1 + mu_ld1).  ``pf_ts_fun`` statistics (:220-259)
come out of the same kernel.  Plots and the summary tables are out of scope.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib
from ._lib import check
from .device import DevicePanel, universe


def _ptr(t):
    return C.c_void_p(t.data_ptr())


def pfml_w(panel, months, w_aim, device="cuda:0", dpanel=None):
    """Back-test over the consecutive month indices ``months`` of the dense panel.

    w_aim [B, NP] (or [B, >= n]) aligned with each date's valid ids.  Returns a dict with ids (list per date),
    w_start, w (lists of arrays per date), stats [B, 5] = inv, shorting, turnover, r, tc, and status [B]."""
    lib = _lib.load()
    dev = torch.device(device)
    months = [int(t) for t in months]
    dp = dpanel or DevicePanel(panel, dev)
    sel, n_sel, n_max, NP = universe(panel, months)
    B = len(months)
    f64 = dict(dtype=torch.float64, device=dev)
    wa = np.zeros((B, NP))
    for b in range(B):
        wa[b, :n_sel[b]] = np.asarray(w_aim[b])[:n_sel[b]]
    wa_d = torch.as_tensor(wa).to(dev)
    dt = torch.as_tensor(np.asarray(months, dtype=np.int32)).to(dev)
    sl = torch.as_tensor(sel).to(dev)
    ns = torch.as_tensor(n_sel).to(dev)
    me = torch.as_tensor(np.ascontiguousarray(panel.me.reshape(-1))).to(dev)
    tr1 = torch.as_tensor(np.ascontiguousarray(panel.tr_ld1.reshape(-1))).to(dev)
    mu1 = torch.as_tensor(np.ascontiguousarray(panel.mu_ld1)).to(dev)
    m = torch.empty((B, NP, NP), **f64)
    status = torch.zeros(B, dtype=torch.int32, device=dev)
    nbytes = lib.pfml_mfunc_dates_workspace_bytes(C.byref(dp.struct), B, n_max)
    ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    carry = torch.zeros(panel.N, **f64)
    w_start = torch.empty((B, NP), **f64)
    w = torch.empty((B, NP), **f64)
    stats = torch.empty((B, 5), **f64)
    with torch.cuda.device(dev):
        s = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        check(lib.pfml_mfunc_dates(C.byref(dp.struct), B, n_max, _ptr(dt), _ptr(sl), _ptr(ns), _ptr(m), _ptr(status), _ptr(ws),
                                   nbytes, s), "pfml_mfunc_dates")
        check(lib.pfml_backtest(C.byref(dp.struct), _ptr(me), _ptr(tr1), _ptr(mu1), B, n_max, _ptr(dt), _ptr(sl), _ptr(ns),
                                _ptr(m), _ptr(wa_d), _ptr(carry), _ptr(w_start), _ptr(w), _ptr(stats), s), "pfml_backtest")
        torch.cuda.synchronize(dev)
    ws_h, w_h = w_start.cpu().numpy(), w.cpu().numpy()
    return dict(ids=[panel.ids[sel[b, :n_sel[b]]] for b in range(B)], w_start=[ws_h[b, :n_sel[b]] for b in range(B)],
                w=[w_h[b, :n_sel[b]] for b in range(B)], stats=stats.cpu().numpy(), status=status.cpu().numpy(),
                m=m)


def pf_ts_fun(result):
    """The per-month statistics of pf_ts_fun / compute_stats (:220-259) as a dict of arrays."""
    names = ("inv", "shorting", "turnover", "r", "tc")
    return {n: result["stats"][:, k] for k, n in enumerate(names)}
