"""GPU: expanding-window sums + ridge grid (PFML_Search_Coef.py) against the reference run and the oracle.

Tolerance 1e-8 relative for lambda > 0.  lambda = 0 systems are ill-conditioned by construction (quirk Q14):
there the bound is condition-aware, 1e-8 * max(1, cond * eps / 1e-8), and the normal-equation residual is checked."""
import os

import numpy as np
import pytest

import pfml_oracle as O
from conftest import GOLD, pkg, relerr

pytestmark = pytest.mark.gpu


def test_search_coef_golden_reference_run():
    sc = pkg("pfml_search_coef")
    z = np.load(os.path.join(GOLD, "pfml_ref_small.npz"))
    month_ends = z["panel_month_ends"][z["hot"]]
    p_vec, l_vec, years = [int(p) for p in z["p_vec"]], z["l_vec"], [int(y) for y in z["years"]]
    coef = sc.search_coef(z["r_tilde"], z["denom"], month_ends, years, p_vec, l_vec)
    bucket, _ = sc.window_buckets(month_ends, years)
    for k, y in enumerate(years, start=1):
        m = (bucket >= 0) & (bucket <= k)
        D = z["denom"][m].sum(0) / m.sum()
        for p in p_vec:
            ref = z[f"coef_{y}_{p}"]
            fi = sc.pfml_feat_idx(p, 16)
            assert (fi == O.pfml_feat_idx(p, 16)).all()
            for j, l in enumerate(l_vec):
                cond = np.linalg.cond(D[np.ix_(fi, fi)] + l * np.eye(p + 1))
                tol = 1e-8 * max(1.0, cond * 2.2e-16 / 1e-8)
                assert relerr(coef[y][p][j], ref[j]) < tol, (y, p, l, cond)


def test_search_coef_vs_oracle_p256(syn):
    sc = pkg("pfml_search_coef")
    rng = np.random.default_rng(3)
    T, p_max = 40, 256
    P = p_max + 1
    month_ends = syn.month_end_range("2011-01", T)
    r_t = rng.standard_normal((T, P)) * 1e-3
    G = rng.standard_normal((T, 2 * P, P)) * 0.1
    denom = np.einsum("tki,tkj->tij", G, G)
    years = [2013, 2014]
    p_vec = [64, 128, 256]
    l_vec = [0.0] + list(np.exp(np.linspace(-10, 10, 5)))
    ref = O.search_coef(r_t, denom, month_ends, years, p_vec, l_vec)
    got = sc.search_coef(r_t, denom, month_ends, years, p_vec, l_vec)
    for y in years:
        for p in p_vec:
            for j in range(len(l_vec)):
                assert relerr(got[y][p][j], ref[y][p][j]) < 1e-8, (y, p, j)
    with pytest.raises(ValueError):
        sc.search_coef(r_t, denom, month_ends, [2030], p_vec, l_vec)


def test_bucket_sums_order_and_skip():
    import torch
    sc = pkg("pfml_search_coef")
    rng = np.random.default_rng(4)
    B, P = 9, 33
    rt = rng.standard_normal((B, P)); dn = rng.standard_normal((B, P, P))
    bucket = np.array([0, 0, -1, 1, 1, 1, 2, -1, 2], dtype=np.int32)
    r_sum, d_sum = sc.bucket_sums(torch.as_tensor(rt).cuda(), torch.as_tensor(dn).cuda(), bucket, 3)
    torch.cuda.synchronize()
    for k in range(3):
        ref = O.denom_sum_fun(dn[i] for i in np.flatnonzero(bucket == k))
        assert (d_sum[k].cpu().numpy() == ref).all()           # same left-to-right order: bit-exact
        assert relerr(r_sum[k].cpu().numpy(), rt[bucket == k].sum(0)) < 1e-15
