"""PFML_Search_Coef.py on the GPU: expanding-window sums and the ridge grid.

Host code keeps the reference's date logic (burn-in window :69-72, yearly windows :105-109, quirk Q13) and its
output schema ``{year: {p: {lambda_index: beta}}}``; the sums and the solves run in libpfml_b200.so.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib
from ._lib import check


def pfml_feat_idx(p, p_max):
    """Integer form of pfml_feat_fun(p) (General_functions.py:837-844) inside the pfml_feat_fun(p_max) layout."""
    h, hm = p // 2, p_max // 2
    return np.concatenate([[0], 1 + np.arange(h), 1 + hm + np.arange(h)]).astype(np.int64)


def pfml_feat_fun(p):
    """General_functions.py:837-844 (labels)."""
    return (["constant"] + [f"rff{x}_cos" for x in range(1, p // 2 + 1)] + [f"rff{x}_sin" for x in range(1, p // 2 + 1)])


def window_buckets(month_ends, hp_years):
    """Bucket of every date: 0 = burn-in (d < (min(hp_years)-2)-12-31, strict, :69-72), k = k-th hp year
    ([(year-2)-12-31, (year-1)-11-30], :105-109), -1 = in no window.  Raises like the reference on an empty window."""
    hp_years = sorted(int(y) for y in hp_years)
    d = np.asarray(month_ends).astype("datetime64[D]")
    bucket = np.full(len(d), -1, dtype=np.int32)
    end_bef = hp_years[0] - 1
    bucket[d < np.datetime64(f"{end_bef - 1}-12-31")] = 0
    if not (bucket == 0).any():
        raise ValueError("empty burn-in window (pd.concat([]) raises in the reference, quirk Q13)")
    for k, year in enumerate(hp_years, start=1):
        m = (d >= np.datetime64(f"{year - 2}-12-31")) & (d <= np.datetime64(f"{year - 1}-11-30"))
        if not m.any():
            raise ValueError(f"empty window for year {year} (quirk Q13)")
        bucket[m] = k
    return bucket, hp_years


def _ptr(t):
    return C.c_void_p(t.data_ptr())


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def bucket_sums(r_tilde, denom, bucket, n_buckets, r_sum=None, d_sum=None):
    """Accumulate per-window sums of device tensors r_tilde [B,P], denom [B,P,P] (bucket: int32 host/devices)."""
    lib = _lib.load()
    dev = denom.device
    B, P = r_tilde.shape
    if r_sum is None:
        r_sum = torch.zeros((n_buckets, P), dtype=torch.float64, device=dev)
        d_sum = torch.zeros((n_buckets, P, P), dtype=torch.float64, device=dev)
    bk = torch.as_tensor(np.asarray(bucket, dtype=np.int32)).to(dev)
    with torch.cuda.device(dev):
        check(lib.pfml_bucket_accumulate(_ptr(r_tilde), _ptr(denom), _ptr(bk), B, P, n_buckets, _ptr(r_sum), _ptr(d_sum),
                                         _stream()), "pfml_bucket_accumulate")
    return r_sum, d_sum


def ridge_grid(r_cum, d_cum, n_cum, p_vec, l_vec, comm=None, prob_range=None):
    """beta for every (year, p, lambda) from running sums r_cum [Y,P], d_cum [Y,P,P], n_cum [Y] (device tensors).
    With a ``comm`` (comm.Communicator) the problems are dealt over its ranks and beta is all-gathered
    (pfml_ridge_grid_sharded); every rank returns the whole grid.  ``prob_range`` = (lo, hi) solves only that range of
    the flat (year, lambda) index (pfml_ridge_grid_range; the rest of beta is zero)."""
    lib = _lib.load()
    dev = d_cum.device
    Y, P = r_cum.shape
    p_max = P - 1
    p_arr = np.asarray(p_vec, dtype=np.int32)
    l_dev = torch.as_tensor(np.asarray(l_vec, dtype=np.float64)).to(dev)
    n_l = len(l_vec)
    total = int(sum(Y * n_l * (int(p) + 1) for p in p_arr))
    beta = torch.zeros(total, dtype=torch.float64, device=dev)
    status = torch.zeros(len(p_arr) * Y * n_l, dtype=torch.int32, device=dev)
    nbytes = lib.pfml_ridge_grid_workspace_bytes(Y, p_max, n_l)
    ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        if prob_range is not None:
            check(lib.pfml_ridge_grid_range(_ptr(r_cum), _ptr(d_cum), _ptr(n_cum), Y, p_max, C.c_void_p(p_arr.ctypes.data), len(p_arr),
                                            _ptr(l_dev), n_l, int(prob_range[0]), int(prob_range[1]), _ptr(beta), _ptr(status),
                                            _ptr(ws), nbytes, _stream()), "pfml_ridge_grid_range")
        else:
            check(lib.pfml_ridge_grid_sharded(comm.handle if comm is not None else None, _ptr(r_cum), _ptr(d_cum), _ptr(n_cum), Y,
                                              p_max, C.c_void_p(p_arr.ctypes.data), len(p_arr), _ptr(l_dev), n_l, _ptr(beta),
                                              _ptr(status), _ptr(ws), nbytes, _stream()), "pfml_ridge_grid_sharded")
    out, off = {}, 0
    for p in p_arr:
        q = int(p) + 1
        out[int(p)] = beta[off:off + Y * n_l * q].view(Y, n_l, q)
        off += Y * n_l * q
    return out, status.view(len(p_arr), Y, n_l)


def coef_from_sums(r_sum, d_sum, cnt, p_vec, l_vec, comm=None, group=None, reduce=True):
    """Second half of the year loop (PFML_Search_Coef.py:112-135) from per-window partial sums of THIS RANK's dates:
    all-reduce over the ranks (``comm``: the library's NCCL communicator; ``group``: a torch.distributed group, used by
    the CPU/gloo tests), running sums over the windows, ridge grid (sharded over ``comm``).  Tensors are modified in place."""
    lib = _lib.load()
    dev = d_sum.device
    nb, P = r_sum.shape
    if reduce and comm is not None:
        comm.allreduce_sums(r_sum, d_sum, cnt)
    elif reduce and group is not None:
        import torch.distributed as dist
        for t in (r_sum, d_sum, cnt):
            dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group if group is not True else None)
    if (comm is not None or group is not None) and bool((cnt == 0).any()):
        raise ValueError("empty window (quirk Q13)")
    with torch.cuda.device(dev):
        check(lib.pfml_bucket_cumsum(_ptr(r_sum), _ptr(d_sum), P, nb, _stream()), "pfml_bucket_cumsum")
    n_cum = torch.cumsum(cnt, 0)
    return ridge_grid(r_sum[1:], d_sum[1:], n_cum[1:].contiguous(), p_vec, l_vec, comm=comm)


def search_coef(r_tilde, denom, month_ends, hp_years, p_vec, l_vec, device="cuda:0", group=None, comm=None):
    """The loop of PFML_Search_Coef.py:60-137 for one g.

    r_tilde [T,P], denom [T,P,P]: numpy arrays or device tensors holding THIS RANK's dates (``month_ends`` [T]);
    with ``comm`` (comm.Communicator: NCCL through the C ABI) or a torch.distributed ``group`` the per-window partial
    sums of all ranks are combined by an all-reduce (the one exchange step of the path); with ``comm`` the ridge
    problems are also dealt over the ranks and beta all-gathered.  Returns ``{year: {p: beta[n_lambda, p+1] (numpy)}}``."""
    lib = _lib.load()
    dev = torch.device(device)
    rt = r_tilde if torch.is_tensor(r_tilde) else torch.as_tensor(np.ascontiguousarray(r_tilde)).to(dev)
    dn = denom if torch.is_tensor(denom) else torch.as_tensor(np.ascontiguousarray(denom)).to(dev)
    P = rt.shape[1]
    hp_sorted = sorted(int(y) for y in hp_years)
    nb = len(hp_sorted) + 1
    multi = group is not None or comm is not None
    if len(month_ends):
        bucket, _ = window_buckets(month_ends, hp_years) if not multi else (_rank_buckets(month_ends, hp_sorted), None)
    else:
        bucket = np.zeros(0, dtype=np.int32)
    counts = np.bincount(bucket[bucket >= 0], minlength=nb).astype(np.float64)
    r_sum = torch.zeros((nb, P), dtype=torch.float64, device=dev)
    d_sum = torch.zeros((nb, P, P), dtype=torch.float64, device=dev)
    if len(bucket):
        bucket_sums(rt, dn, bucket, nb, r_sum, d_sum)
    cnt = torch.as_tensor(counts).to(dev)
    betas, status = coef_from_sums(r_sum, d_sum, cnt, p_vec, l_vec, comm=comm, group=group)
    torch.cuda.synchronize(dev)
    coef = {}
    for k, year in enumerate(hp_sorted):
        coef[year] = {int(p): betas[int(p)][k].cpu().numpy() for p in p_vec}
    return coef


def rank_buckets(month_ends, hp_sorted):
    """Windows of a rank's own dates (no emptiness check: that is global, after the all-reduce)."""
    return _rank_buckets(month_ends, hp_sorted)


def _rank_buckets(month_ends, hp_sorted):
    """Buckets of a rank's own dates; emptiness is checked globally after the all-reduce."""
    d = np.asarray(month_ends).astype("datetime64[D]")
    bucket = np.full(len(d), -1, dtype=np.int32)
    bucket[d < np.datetime64(f"{hp_sorted[0] - 2}-12-31")] = 0
    for k, year in enumerate(hp_sorted, start=1):
        bucket[(d >= np.datetime64(f"{year - 2}-12-31")) & (d <= np.datetime64(f"{year - 1}-11-30"))] = k
    return bucket


def denom_sum_fun(train):
    """PFML_Search_Coef.py:37-46 for a dict {date: {'denom': array-like}} (device sum in date order)."""
    dn = torch.as_tensor(np.stack([np.asarray(v["denom"]) for v in train.values()])).cuda()
    B, P, _ = dn.shape
    rt = torch.zeros((B, P), dtype=torch.float64, device=dn.device)
    _, d_sum = bucket_sums(rt, dn, np.zeros(B, dtype=np.int32), 1)
    torch.cuda.synchronize()
    return d_sum[0].cpu().numpy()


def run(path, settings=None, device="cuda:0", source="pickle"):
    """PFML_Search_Coef.py: reads Data/pfml_input_{g}.pkl (or, with source="store", the dense tensor store written by
    pfml_input_data.run_store), writes Data/coef_dict_{g}.pkl (:60-143), same schema
    ``{g: {year: {p: {lambda_index: Series}}}}``."""
    import os
    import pickle

    import pandas as pd

    from . import general_functions as GF
    settings, _pf = settings or GF.get_settings()
    hp_years = np.arange(settings["pf"]["dates"]["start_year"], settings["pf"]["dates"]["end_yr"] + 1)
    p_vec, l_vec = settings["pf_ml"]["p_vec"], settings["pf_ml"]["l_vec"]
    out = {}
    for g_index, _g in enumerate(settings["pf_ml"]["g_vec"]):
        if source == "store":
            from .tensor_store import Store
            st = Store(path, g_index)
            r_t, den, me = np.array(st.r_tilde), st.denom_stack(), st.dates
        else:
            with open(os.path.join(path, "Data", f"pfml_input_{g_index}.pkl"), "rb") as f:
                reals = pickle.load(f)[g_index]["reals"]
            d_all = sorted(reals.keys())
            r_t = np.stack([reals[d]["r_tilde"].to_numpy() for d in d_all])
            den = np.stack([reals[d]["denom"].to_numpy() for d in d_all])
            me = np.array([np.datetime64(pd.Timestamp(d).date()) for d in d_all])
        coef = search_coef(r_t, den, me, hp_years, p_vec, l_vec, device=device)
        coef_dict = {int(y): {int(p): {j: pd.Series(coef[int(y)][int(p)][j], index=pfml_feat_fun(int(p))) for j in range(len(l_vec))}
                              for p in p_vec} for y in hp_years}
        with open(os.path.join(path, "Data", f"coef_dict_{g_index}.pkl"), "wb") as f:
            pickle.dump({g_index: coef_dict}, f)
        out[g_index] = coef_dict
    return out
