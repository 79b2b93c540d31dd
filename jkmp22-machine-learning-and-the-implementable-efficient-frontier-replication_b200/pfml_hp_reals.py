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
