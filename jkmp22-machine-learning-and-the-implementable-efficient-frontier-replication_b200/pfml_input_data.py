"""Drop-in for PFML_Input_Data.py: same artefacts in (``path + "Data/"``), same pickles out, arithmetic on the GPU.

    run(path)            the script: reads JKP_US_SP500.db:Factors_processed, FF_RF_monthly.csv,
                         wealth_processed.csv, Barra_Cov.pkl, rff_w.csv (:63-108, :245) and writes
                         Data/pfml_input_{g_index}.pkl with the schema of :484-495
    rff(X, p, g, W)      the function of :159-185

Host code only does what the reference does with pandas before and after the numerical loop (reading tables,
date ranges :133-154, id/date bookkeeping, building labelled DataFrames); every formula runs in libpfml_b200.so.
"""
from __future__ import annotations

import os
import pickle
import sqlite3

import numpy as np
import pandas as pd

from . import general_functions as GF
from . import synthetic


def rff(X, p=None, g=None, W=None, device="cuda:0"):
    """Random Fourier features (:159-185).  W is normally given (quirk Q1); if not, it is drawn N(0, g I) with
    numpy's default_rng -- the reference's own draw is not reproducible either (quirk Q2)."""
    import ctypes as C
    import torch
    from . import _lib
    lib = _lib.load()
    Xn = np.ascontiguousarray(np.asarray(X, dtype=np.float64))
    if W is None:
        W = np.sqrt(g) * np.random.default_rng().standard_normal((Xn.shape[1], p // 2))
    Wn = np.ascontiguousarray(np.asarray(W, dtype=np.float64))
    dev = torch.device(device)
    xd, wd = torch.as_tensor(Xn).to(dev), torch.as_tensor(Wn).to(dev)
    out = torch.empty((Xn.shape[0], 2 * Wn.shape[1]), dtype=torch.float64, device=dev)
    with torch.cuda.device(dev):
        _lib.check(lib.pfml_rff_sincos(C.c_void_p(xd.data_ptr()), C.c_void_p(wd.data_ptr()), Xn.shape[0], Xn.shape[1],
                                       Wn.shape[1], C.c_void_p(out.data_ptr()),
                                       C.c_void_p(torch.cuda.current_stream().cuda_stream)), "pfml_rff_sincos")
        torch.cuda.synchronize(dev)
    o = out.cpu().numpy()
    h = Wn.shape[1]
    return {"W": W, "X_cos": o[:, :h], "X_sin": o[:, h:]}


def read_panel(path, settings, pf_set):
    """The reading part of the script (:39-142) into a dense ``synthetic.Panel``."""
    features = GF.get_features(exclude_poor_coverage=True)
    data = os.path.join(path, "Data")
    con = sqlite3.connect(os.path.join(data, "JKP_US_SP500.db"))
    cols = "id, eom, valid, ret_ld1, tr_ld0, mu_ld0, lambda, " + ", ".join(features)
    chars = pd.read_sql_query(f"SELECT {cols} FROM Factors_processed", con, parse_dates=["eom"])
    con.close()
    chars["valid"] = chars["valid"].astype(bool)
    rf = pd.read_csv(os.path.join(data, "FF_RF_monthly.csv"), usecols=["yyyymm", "RF"])
    rf["rf"] = rf["RF"] / 100
    rf["eom"] = pd.to_datetime(rf["yyyymm"].astype(str) + "01", format="%Y%m%d") + pd.offsets.MonthEnd(0)
    wealth = pd.read_csv(os.path.join(data, "wealth_processed.csv"), parse_dates=["eom"])
    with open(os.path.join(data, "Barra_Cov.pkl"), "rb") as f:
        barra = pickle.load(f)
    W = pd.read_csv(os.path.join(data, "rff_w.csv")).drop(columns="Unnamed: 0").to_numpy()

    first_cov = pd.to_datetime(min(barra.keys()))
    lb = pf_set["lb_hor"]
    dates_m2 = pd.date_range(start=first_cov + pd.DateOffset(months=lb + 1),
                             end=settings["split"]["test_end"] + pd.Timedelta(days=1) - pd.DateOffset(months=1),
                             freq="MS") - pd.DateOffset(days=1)
    months = pd.date_range(start=first_cov, end=dates_m2.max(), freq="ME")
    chars = chars[(chars["eom"] >= months[0]) & (chars["eom"] <= months[-1])]
    ids = np.sort(chars["id"].unique()).astype(np.int64)
    T, N, F = len(months), len(ids), len(features)
    ti = pd.Index(months).get_indexer(chars["eom"])
    ii = pd.Index(ids).get_indexer(chars["id"])
    if (ti < 0).any():
        raise ValueError("Factors_processed has eom values that are not calendar month-ends (quirk Q16)")

    def dense(col, fill=np.nan, dtype=np.float64):
        a = np.full((T, N), fill, dtype=dtype)
        a[ti, ii] = chars[col].to_numpy()
        return a

    feats = np.full((T, N, F), np.nan)
    feats[ti, ii] = chars[features].to_numpy(dtype=np.float64)
    present = np.zeros((T, N), dtype=bool)
    present[ti, ii] = True
    valid = dense("valid", False, bool)
    K = barra[months[0]]["fct_cov"].shape[0]
    fct_load = np.zeros((T, N, K)); fct_cov = np.zeros((T, K, K)); ivol = np.zeros((T, N))
    has = np.zeros((T, N), dtype=bool)
    for t, d in enumerate(months):
        b = barra[d]
        bi = pd.Index(ids).get_indexer(np.asarray(b["fct_load"].index, dtype=np.int64))
        ok = bi >= 0
        fct_load[t, bi[ok]] = b["fct_load"].to_numpy(dtype=np.float64)[ok]
        ivol[t, bi[ok]] = b["ivol_vec"].loc[b["fct_load"].index].to_numpy(dtype=np.float64)[ok]
        fct_cov[t] = b["fct_cov"].to_numpy(dtype=np.float64)
        has[t, bi[ok]] = True
    w_by = wealth.set_index("eom")["wealth"].reindex(months).to_numpy()
    rf_by = rf.set_index("eom")["rf"].reindex(months).to_numpy()
    lam = dense("lambda")
    if not settings.get("Transaction_Costs", True):
        lam = np.where(present, 1e-16, np.nan)                       # :121-126
    panel = synthetic.Panel(ids=ids, month_ends=months.values.astype("datetime64[D]"), feats=feats, valid=valid,
                            ret_ld1=dense("ret_ld1"), tr_ld0=dense("tr_ld0"), mu_ld0=dense("mu_ld0"), lam=lam, wealth=w_by,
                            rf=rf_by, fct_load=fct_load, fct_cov=fct_cov, ivol=ivol, barra_has=has, W=W,
                            gamma_rel=float(pf_set["gamma_rel"]), mu=float(pf_set["mu"]), lb=int(lb), iterations=10)
    hot = pd.Index(months).get_indexer(dates_m2)
    # every id of a date's universe needs a row in all lb+2 cross-sections (the reference fails otherwise, :388)
    for t in hot:
        sel = valid[t]
        if not present[t - lb - 1:t + 1][:, sel].all():
            raise ValueError(f"ids valid at {months[t].date()} lack rows in the {lb + 2} look-back months")
        if not has[t][sel].all():
            raise KeyError(f"Barra_Cov[{months[t].date()}] lacks ids of that date's universe")
    return panel, hot, months


def run(path, settings=None, device="cuda:0", batch=64):
    """PFML_Input_Data.py for every g in settings['pf_ml']['g_vec'] (quirk Q1: W comes from rff_w.csv for every g,
    so the date-steps are computed once and written once per g_index, quirk Q10)."""
    from .device import DevicePanel, date_steps
    settings, pf_set = settings or GF.get_settings()
    panel, hot, months = read_panel(path, settings, pf_set)
    p_max = max(settings["pf_ml"]["p_vec"])
    if panel.p != p_max:
        raise ValueError(f"rff_w.csv has {panel.p // 2} columns but p_max = {p_max}")
    feat_all = GF.pfml_feat_fun(p_max)
    # NaN characteristics only matter on rows that enter a universe's look-back; zero the rest for the RFF pass
    panel.feats = np.nan_to_num(panel.feats, nan=0.0)
    dp = DevicePanel(panel, device)
    reals, signal_t = {}, {}
    for i in range(0, len(hot), batch):
        chunk = [int(t) for t in hot[i:i + batch]]
        for t, o in zip(chunk, date_steps(panel, chunk, device, dpanel=dp)):
            d = pd.Timestamp(months[t])
            if o["status"] & 3:
                raise np.linalg.LinAlgError(f"date-step {d.date()}: numerical failure on the device (status {o['status']})")
            reals[d] = {"r_tilde": pd.Series(o["r_tilde"], index=feat_all),
                        "denom": pd.DataFrame(o["denom"], index=feat_all, columns=feat_all),
                        "risk": pd.DataFrame(o["risk"], index=feat_all, columns=feat_all),
                        "tc": pd.DataFrame(o["tc"], index=feat_all, columns=feat_all)}
            signal_t[d] = pd.DataFrame(o["signal_t"], index=o["ids"].astype(int), columns=feat_all)
    out = {}
    for g_index, _g in enumerate(settings["pf_ml"]["g_vec"]):
        pfml_input = {g_index: {"reals": reals, "signal_t": signal_t, "rff_w": panel.W}}
        with open(os.path.join(path, "Data", f"pfml_input_{g_index}.pkl"), "wb") as f:
            pickle.dump(pfml_input, f)
        out[g_index] = pfml_input[g_index]
    return out


def run_store(path, settings=None, device=0, block=64):
    """PFML_Input_Data.py into the dense tensor store (tensor_store.py) instead of the pickle: blocks of dates go through
    ``pfml_date_steps_host_packed`` (pinned host buffers, packed upper triangles of risk / tc) straight into the
    memory-mapped arrays of ``Data/pfml_store_{g}/``.  Same information as ``run`` writes, a third of the bytes, no
    pandas objects; ``tensor_store.Store(path, g).to_legacy()`` rebuilds the reference's dict on demand."""
    from .host import HostBlock
    from .tensor_store import StoreWriter
    settings, pf_set = settings or GF.get_settings()
    panel, hot, months = read_panel(path, settings, pf_set)
    p_max = max(settings["pf_ml"]["p_vec"])
    if panel.p != p_max:
        raise ValueError(f"rff_w.csv has {panel.p // 2} columns but p_max = {p_max}")
    panel.feats = np.nan_to_num(panel.feats, nan=0.0)
    P = panel.p + 1
    n_max_all = int(max(panel.valid[int(t)].sum() for t in hot))
    from . import _lib
    NP = _lib.load().pfml_padded_n(n_max_all)
    writers = [StoreWriter(path, g_index, months[hot].values.astype("datetime64[D]"), P, NP, panel.W)
               for g_index, _g in enumerate(settings["pf_ml"]["g_vec"])]
    for i in range(0, len(hot), block):
        chunk = [int(t) for t in hot[i:i + block]]
        blk = HostBlock(panel, chunk, pin=True).run_packed(device=device)
        bad = blk.status.numpy() & 3
        if bad.any():
            raise np.linalg.LinAlgError(f"date-step {months[chunk[int(np.flatnonzero(bad)[0])]].date()}: numerical failure on the device")
        sig = np.zeros((blk.B, NP, P))
        sig[:, :blk.NP] = blk.out["signal_t"].numpy()
        ids_list = [panel.ids[blk.sel[b, :int(blk.n_sel[b])]] for b in range(blk.B)]
        for w in writers:                                  # quirk Q1 / Q10: W does not depend on g, every g_index gets the same tensors
            w.put_block(i, blk.out["r_tilde"].numpy(), blk.packed["risk_ut"].numpy(), blk.packed["tc_ut"].numpy(), sig, ids_list)
    for w in writers:
        w.close()
    return [w.dir for w in writers]
