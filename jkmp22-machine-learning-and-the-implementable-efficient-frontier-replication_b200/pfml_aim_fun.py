"""PFML_aim_fun.py on the GPU: the aim portfolio of every out-of-sample month.

Host code picks, per month, the rank-1 (p, lambda) of the previous December from the validation table
(:130-149, first rank-1 row wins like ``.values[0]``) and scatters its coefficients into the feat_all layout;
the products ``signal_t[d][feat] @ beta`` (:150-160) are one kernel launch for all months.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import pandas as pd
import torch

from . import _lib
from ._lib import check
from .pfml_search_coef import pfml_feat_idx


def best_hps(validation):
    """Rows with rank 1 whose eom_ret is a December, keyed by that year (:130-134)."""
    v = validation
    opt = v[(v["eom_ret"].dt.month == 12) & (v["rank"] == 1)].assign(hp_end=lambda x: x["eom_ret"].dt.year)
    return opt[["hp_end", "l", "p"]].reset_index(drop=True)


def aim_portfolios(validation, coef, signal_t, n_sel, month_ends, device="cuda:0"):
    """w_aim for the dates of ``month_ends`` (the OOS months).

    signal_t: device tensor or array [B, NP, P] (rows >= n_sel[b] ignored), coef: {year: {p: beta[n_lambda, p+1]}}.
    Returns (w_aim [B, NP] numpy, chosen p [B], chosen lambda index [B])."""
    lib = _lib.load()
    dev = torch.device(device)
    S = signal_t if torch.is_tensor(signal_t) else torch.as_tensor(np.ascontiguousarray(signal_t)).to(dev)
    B, NP, P = S.shape
    opt = best_hps(validation)
    beta_full = np.zeros((B, P))
    ps, ls = np.zeros(B, dtype=np.int64), np.zeros(B, dtype=np.int64)
    for b, d in enumerate(pd.to_datetime(np.asarray(month_ends).astype("datetime64[D]"))):
        oos_year = (d + pd.offsets.MonthEnd(1)).year                     # :136-138
        h = opt[opt["hp_end"] == oos_year - 1]
        if len(h) == 0:
            raise IndexError(f"no rank-1 December row for hp year {oos_year - 1} (the reference fails at .values[0])")
        p, l = int(h["p"].values[0]), int(h["l"].values[0])
        beta_full[b, pfml_feat_idx(p, P - 1)] = coef[oos_year][p][l]
        ps[b], ls[b] = p, l
    bf = torch.as_tensor(beta_full).to(dev)
    ns = torch.as_tensor(np.asarray(n_sel, dtype=np.int32)).to(dev)
    w = torch.empty((B, NP), dtype=torch.float64, device=dev)
    with torch.cuda.device(dev):
        check(lib.pfml_aim_weights(C.c_void_p(S.data_ptr()), C.c_void_p(ns.data_ptr()), B, NP, P, C.c_void_p(bf.data_ptr()),
                                   C.c_void_p(w.data_ptr()), C.c_void_p(torch.cuda.current_stream().cuda_stream)),
              "pfml_aim_weights")
        torch.cuda.synchronize(dev)
    return w.cpu().numpy(), ps, ls


def run(path, settings=None, device="cuda:0"):
    """PFML_aim_fun.py: reads Data/validation.csv, Data/coef_dict_{g}.pkl, signal_t of Data/pfml_input_{g}.pkl and
    writes Data/aims.pkl = {g: {d: {'aim_pf': DataFrame(id, eom, w_aim), 'coef': Series}}} (:163-169).
    The out-of-sample months run from January of start_year + split_years to test_end - 1 month (:90-98; the
    reference overwrites test_end with 2023-12-31 at :94, here it comes from the settings)."""
    import os
    import pickle

    from . import general_functions as GF
    from .general_functions import pfml_feat_fun
    settings, _pf = settings or GF.get_settings()
    start_oos = settings["pf"]["dates"]["start_year"] + settings["pf"]["dates"]["split_years"]
    dates_oos = pd.date_range(start=pd.Timestamp(f"{start_oos}-01-01"),
                              end=settings["split"]["test_end"] + pd.Timedelta(days=1) - pd.DateOffset(months=1),
                              freq="MS") - pd.DateOffset(days=1)
    l_vec, p_vec = settings["pf_ml"]["l_vec"], settings["pf_ml"]["p_vec"]
    val_all = pd.read_csv(os.path.join(path, "Data", "validation.csv"), parse_dates=["eom", "eom_ret"])
    aims = {}
    for g_index, _g in enumerate(settings["pf_ml"]["g_vec"]):
        with open(os.path.join(path, "Data", f"pfml_input_{g_index}.pkl"), "rb") as f:
            signal_t = pickle.load(f)[g_index]["signal_t"]
        with open(os.path.join(path, "Data", f"coef_dict_{g_index}.pkl"), "rb") as f:
            hp_coef = pickle.load(f)[g_index]
        v = val_all[val_all["g"] == g_index]
        coef = {int(y): {int(p): np.stack([hp_coef[y][p][j].to_numpy() for j in range(len(l_vec))]) for p in p_vec}
                for y in hp_coef.keys()}
        sig = [signal_t[d] for d in dates_oos]
        n_sel = np.array([len(s) for s in sig], dtype=np.int32)
        P = sig[0].shape[1]
        S = np.zeros((len(sig), int(n_sel.max()), P))
        for b, s in enumerate(sig):
            S[b, :n_sel[b]] = s.to_numpy()
        me = np.array([np.datetime64(d.date()) for d in dates_oos])
        w, ps, ls = aim_portfolios(v, coef, S, n_sel, me, device=device)
        out = {}
        for b, d in enumerate(dates_oos):
            oos_year = (d + pd.offsets.MonthEnd(1)).year
            feat = pfml_feat_fun(int(ps[b]))
            aim_pf = pd.DataFrame({"id": sig[b].index.to_numpy(), "eom": d, "w_aim": w[b, :n_sel[b]]})
            out[d] = {"aim_pf": aim_pf, "coef": pd.Series(coef[oos_year][int(ps[b])][int(ls[b])], index=feat)}
        aims[g_index] = out
    with open(os.path.join(path, "Data", "aims.pkl"), "wb") as f:
        pickle.dump(aims, f)
    return aims
