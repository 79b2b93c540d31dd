"""The whole hp-grid pipeline on 1..N GPUs: PFML_Input_Data.py's date loop feeding PFML_Search_Coef.py's year loop.

For every bandwidth g (PFML_Input_Data.py:225; one W per g, quirk Q1):

  1. this rank's contiguous block of dates (sharding.shard_dates) runs through ``pfml_date_steps_multi_g`` in batches:
     r_tilde / denom of every date and every g (PFML_Input_Data.py:318-491; Sigma, m_func, the gtm chains and the two
     ``const`` inverses of a date are shared over g, quirk Q10) and, fused into the a9 finish kernel, the per-window
     partial sums of PFML_Search_Coef.py:77-88, :112-121 over this rank's dates;
  2. one FP64 NCCL all-reduce of the ``[windows] x (P + P^2 + 1)`` partial sums (``pfml_comm_allreduce_sums``) -- the one
     real exchange of the path;
  3. running sums over the windows (``pfml_bucket_cumsum``), then the ridge grid (``pfml_ridge_grid_sharded``): the
     (year, lambda) problems of every p dealt over the ranks, beta all-gathered (PFML_Search_Coef.py:125-135).

Nothing but the window sums and beta crosses NVLink.  Determinism: fixed date -> rank map, in-order sums inside a rank;
N-rank and 1-rank results differ by FP64 summation order only.  ``virtual_world`` runs the ranks of a world one after the
other on ONE GPU (partial sums added in rank order instead of the all-reduce, ridge ranges written into one beta): the
same host logic and kernels as the N-GPU run without the NCCL transport, for single-GPU test boxes.
"""
from __future__ import annotations

import dataclasses
import time

import numpy as np
import torch

from . import device as dm
from . import pfml_search_coef as sc
from . import sharding
from .comm import shard_range


def slice_panel(panel, t0, t1, W=None):
    """Months [t0, t1] of a synthetic.Panel (a rank only needs its dates plus lb + 1 look-back months resident)."""
    kw = {}
    for f in dataclasses.fields(panel):
        v = getattr(panel, f.name)
        if isinstance(v, np.ndarray) and f.name not in ("ids", "W") and v.shape[0] == panel.T:
            v = v[t0:t1 + 1]
        kw[f.name] = v
    if W is not None:
        kw["W"] = W
    return type(panel)(**kw)


def partial_sums(panel, W_by_g, months, mine, hp_sorted, dev, batch=148):
    """Per-window partial sums of one rank for every bandwidth g: date-steps of ``mine`` (a sub-list of ``months``) with
    the sums fused in; m, the chains and the two inverses of a date are computed once and shared over g
    (pfml_date_steps_multi_g, quirk Q10).  Returns r_sum [G, nb, P], d_sum [G, nb, P, P], cnt [nb] on ``dev``."""
    nb, G = len(hp_sorted) + 1, len(W_by_g)
    P = 2 * W_by_g[0].shape[1] + 1
    r_sum = torch.zeros((G, nb, P), dtype=torch.float64, device=dev)
    d_sum = torch.zeros((G, nb, P, P), dtype=torch.float64, device=dev)
    cnt = torch.zeros(nb, dtype=torch.float64, device=dev)
    if len(mine) == 0:
        return r_sum, d_sum, cnt
    t0, t1 = sharding.panel_slice_for(mine, panel.lb)
    sub = slice_panel(panel, t0, t1, W=W_by_g[0])
    dp = dm.DevicePanel(sub, dev, extra_W=W_by_g[1:])
    local = np.asarray(mine) - t0
    bucket_all = sc.rank_buckets(panel.month_ends[np.asarray(mine)], hp_sorted)
    cnt += torch.as_tensor(np.bincount(bucket_all[bucket_all >= 0], minlength=nb).astype(np.float64)).to(dev)
    sel, n_sel, n_max, NP = dm.universe(sub, local)
    bsz = min(batch, len(local))
    runner = dm.MultiGRunner(dp, bsz, n_max)
    for b0 in range(0, len(local), bsz):
        sl = slice(b0, min(b0 + bsz, len(local)))
        runner.set_batch(local[sl], sel[sl], n_sel[sl])
        bk = torch.as_tensor(bucket_all[sl].astype(np.int32)).to(dev)
        runner.run(bk, nb, r_sum, d_sum)
        if int(runner.status.abs().sum()) != 0:
            raise RuntimeError("numerical failure flag set in the date-steps")
    return r_sum, d_sum, cnt


def hp_grid_pipeline(panel, W_by_g, months, hp_years, p_vec, l_vec, comm=None, device="cuda:0", batch=148, timing=None,
                     virtual_world=None):
    """Returns ``{g_index: {year: {p: beta[n_lambda, p+1]}}}`` (numpy, identical on every rank).

    ``panel``: synthetic.Panel-like host panel (its W is ignored); ``W_by_g``: list of W matrices [F, p/2], one per g;
    ``months``: month indices of ALL hot dates (every rank passes the same list); ``comm``: comm.Communicator or None."""
    dev = torch.device(device)
    hp_sorted = sorted(int(y) for y in hp_years)
    months = np.asarray(list(months))
    sc.window_buckets(panel.month_ends[months], hp_years)             # raises on an empty window like the reference (Q13)
    rank, world = (comm.rank, comm.world) if comm is not None else (0, 1)
    out = {}
    tm = dict(date_steps_s=0.0, exchange_s=0.0, ridge_s=0.0, allreduce_bytes=0)
    torch.cuda.synchronize(dev)
    ta = time.perf_counter()
    if virtual_world:
        parts = [partial_sums(panel, W_by_g, months, sharding.shard_dates(months, r, virtual_world), hp_sorted, dev, batch)
                 for r in range(virtual_world)]
        r_all, d_all, cnt_all = parts[0]
        for pr, pd_, pc in parts[1:]:                              # what the all-reduce computes, in rank order
            r_all += pr; d_all += pd_; cnt_all += pc
    else:
        r_all, d_all, cnt_all = partial_sums(panel, W_by_g, months, sharding.shard_dates(months, rank, world), hp_sorted, dev, batch)
    torch.cuda.synchronize(dev)
    tm["date_steps_s"] = time.perf_counter() - ta
    for g_index in range(len(W_by_g)):
        r_sum, d_sum, cnt = r_all[g_index], d_all[g_index], cnt_all.clone()
        torch.cuda.synchronize(dev)
        tb = time.perf_counter()
        if comm is not None:
            comm.allreduce_sums(r_sum, d_sum, cnt)
            tm["allreduce_bytes"] += int(r_sum.numel() + d_sum.numel() + cnt.numel()) * 8
        torch.cuda.synchronize(dev)
        tc = time.perf_counter()
        if virtual_world:
            lib_nb, P = r_sum.shape
            from ._lib import check, load
            import ctypes as C
            with torch.cuda.device(dev):
                check(load().pfml_bucket_cumsum(C.c_void_p(r_sum.data_ptr()), C.c_void_p(d_sum.data_ptr()), P, lib_nb,
                                                C.c_void_p(torch.cuda.current_stream().cuda_stream)), "pfml_bucket_cumsum")
            n_cum = torch.cumsum(cnt, 0)
            nprob = len(hp_sorted) * len(l_vec)
            betas = None
            for r in range(virtual_world):                            # disjoint ranges: the sum of the pieces is the all-gather
                b_r, _ = sc.ridge_grid(r_sum[1:], d_sum[1:], n_cum[1:].contiguous(), p_vec, l_vec,
                                       prob_range=shard_range(nprob, r, virtual_world))
                betas = b_r if betas is None else {p: betas[p] + b_r[p] for p in betas}
        else:
            betas, _ = sc.coef_from_sums(r_sum, d_sum, cnt, p_vec, l_vec, comm=comm, reduce=False)
        torch.cuda.synchronize(dev)
        td = time.perf_counter()
        out[g_index] = {year: {int(p): betas[int(p)][k].cpu().numpy() for p in p_vec} for k, year in enumerate(hp_sorted)}
        tm["exchange_s"] += tc - tb
        tm["ridge_s"] += td - tc
    if timing is not None:
        timing.update(tm)
    return out
