"""PFML_hp_reals.py on the GPU: the validation utility table.

For every hp year, the 12 months after its training window are scored with that year's coefficients:
obj = r~' beta - 1/2 beta' denom beta per (p, lambda, month) (:90-102, one kernel launch per p), then the expanding
mean per (p, lambda) (:109-115) and the dense rank per eom_ret (:118-122) with pandas, as in the reference.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import pandas as pd
import torch

from . import _lib
from ._lib import check


def _ptr(t):
    return C.c_void_p(t.data_ptr())


def score_windows(month_ends, hp_years):
    """date -> index of the hp year that scores it: [(end-1)-12-31, end-11-30] (:76), -1 outside every window."""
    d = np.asarray(month_ends).astype("datetime64[D]")
    yrs = sorted(int(y) for y in hp_years)
    idx = np.full(len(d), -1, dtype=np.int32)
    for k, end in enumerate(yrs):
        idx[(d >= np.datetime64(f"{end - 1}-12-31")) & (d <= np.datetime64(f"{end}-11-30"))] = k
    return idx, yrs


def validation(r_tilde, denom, month_ends, coef, hp_years, p_vec, l_vec, device="cuda:0", g_index=0):
    """The validation DataFrame of one g: columns eom, eom_ret, obj, l, p, hp_end, cum_obj, rank, g.

    r_tilde [T,P], denom [T,P,P]: numpy or device tensors; coef: {year: {p: beta[n_lambda, p+1]}}."""
    lib = _lib.load()
    dev = torch.device(device)
    rt = r_tilde if torch.is_tensor(r_tilde) else torch.as_tensor(np.ascontiguousarray(r_tilde)).to(dev)
    dn = denom if torch.is_tensor(denom) else torch.as_tensor(np.ascontiguousarray(denom)).to(dev)
    T, P = rt.shape
    year_idx, yrs = score_windows(month_ends, hp_years)
    yi = torch.as_tensor(year_idx).to(dev)
    n_l = len(l_vec)
    me = pd.to_datetime(np.asarray(month_ends).astype("datetime64[D]"))
    frames = []
    with torch.cuda.device(dev):
        for p in p_vec:
            beta = torch.as_tensor(np.ascontiguousarray(np.stack([coef[y][int(p)] for y in yrs]))).to(dev)   # [Y][n_l][q]
            obj = torch.empty((T, n_l), dtype=torch.float64, device=dev)
            check(lib.pfml_validation_utility(_ptr(rt), _ptr(dn), T, P - 1, _ptr(yi), _ptr(beta), int(p), n_l, _ptr(obj),
                                              C.c_void_p(torch.cuda.current_stream().cuda_stream)), "pfml_validation_utility")
            o = obj.cpu().numpy()
            scored = np.flatnonzero(year_idx >= 0)
            for l in range(n_l):
                frames.append(pd.DataFrame({"eom": me[scored], "obj": o[scored, l], "l": l, "p": int(p),
                                            "hp_end": np.asarray(yrs)[year_idx[scored]]}))
    v = pd.concat(frames, ignore_index=True)
    v["eom_ret"] = v["eom"] + pd.offsets.MonthEnd(1)
    v = v.sort_values(["p", "l", "eom_ret"])
    v["cum_obj"] = v.groupby(["p", "l"])["obj"].expanding().mean().reset_index(level=[0, 1], drop=True)
    v["rank"] = v.groupby("eom_ret")["cum_obj"].rank(ascending=False, method="dense")
    return v.assign(g=g_index).reset_index(drop=True)


def run(path, settings=None, device="cuda:0", strict_reference_bug=False):
    """PFML_hp_reals.py: reads Data/pfml_input_{g}.pkl and Data/coef_dict_{g}.pkl, writes Data/validation.csv (:130).

    Quirk Q12: the reference initialises ``rows`` outside the g loop (:60 vs :62), so the table of g = 1 also contains
    the rows of g = 0 relabelled g = 1.  ``strict_reference_bug=True`` reproduces that; the default gives every g its
    own rows."""
    import os
    import pickle

    from . import general_functions as GF
    settings, _pf = settings or GF.get_settings()
    hp_years = np.arange(settings["pf"]["dates"]["start_year"], settings["pf"]["dates"]["end_yr"] + 1)
    p_vec, l_vec = settings["pf_ml"]["p_vec"], settings["pf_ml"]["l_vec"]
    tables, carried = [], []
    for g_index, _g in enumerate(settings["pf_ml"]["g_vec"]):
        with open(os.path.join(path, "Data", f"pfml_input_{g_index}.pkl"), "rb") as f:
            reals = pickle.load(f)[g_index]["reals"]
        with open(os.path.join(path, "Data", f"coef_dict_{g_index}.pkl"), "rb") as f:
            hp_coef = pickle.load(f)[g_index]
        d_all = sorted(reals.keys())
        r_t = np.stack([reals[d]["r_tilde"].to_numpy() for d in d_all])
        den = np.stack([reals[d]["denom"].to_numpy() for d in d_all])
        me = np.array([np.datetime64(pd.Timestamp(d).date()) for d in d_all])
        coef = {int(y): {int(p): np.stack([hp_coef[y][p][j].to_numpy() for j in range(len(l_vec))]) for p in p_vec}
                for y in hp_years}
        v = validation(r_t, den, me, coef, hp_years, p_vec, l_vec, device=device, g_index=g_index)
        if strict_reference_bug and carried:
            raw = pd.concat(carried + [v[["eom", "eom_ret", "obj", "l", "p", "hp_end"]]], ignore_index=True)
            raw = raw.sort_values(["p", "l", "eom_ret"], kind="stable")
            raw["cum_obj"] = raw.groupby(["p", "l"])["obj"].expanding().mean().reset_index(level=[0, 1], drop=True)
            raw["rank"] = raw.groupby("eom_ret")["cum_obj"].rank(ascending=False, method="dense")
            carried.append(v[["eom", "eom_ret", "obj", "l", "p", "hp_end"]])
            v = raw.assign(g=g_index)
        else:
            carried.append(v[["eom", "eom_ret", "obj", "l", "p", "hp_end"]])
        tables.append(v)
    out = pd.concat(tables)
    out[["eom", "eom_ret", "obj", "l", "p", "hp_end", "cum_obj", "rank", "g"]].to_csv(os.path.join(path, "Data", "validation.csv"),
                                                                                    index=False)
    return out
