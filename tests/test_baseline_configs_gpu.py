"""GPU: the CUDA path against the oracle AT THE CONFIGURATIONS BASELINE.json NAMES (through the C ABI).

  metric size     N = 500, p = 1024 (P = 1025 edge tiles, upper-triangle a9 launches, 1 025-column slicing), in all
                  three arithmetic modes of the n x n / n x P products (0 = FP64 DMMA, 1 = INT8 7 digits, 2 = INT8 6 digits)
  C3              N = 500, p = 512, three distinct bandwidths g (three W matrices)
  C4              ridge grid 15 lambda x p in {64 .. 1024} x 28 yearly refits on a P = 1025 layout
  C5 universe     N = 3000 (padded 3008: DMMA a9, three-pass slicing kernels), p = 64, one date, against the frozen
                  oracle vectors of oracle/make_golden_n3000.py
  second fixture  N = 150, p = 64 outputs of the exec'd reference script (oracle/make_golden_mid.py)

Tolerance 1e-8 relative (north_star), written below; lambda = 0 of the ridge grid condition-aware (quirk Q14).
"""
import os

import numpy as np
import pytest

import pfml_oracle as O
from conftest import GOLD, panel_from_golden, pkg, relerr

pytestmark = pytest.mark.gpu
TOL = 1e-8
NAMES = ("r_tilde", "denom", "risk", "tc", "signal_t")


@pytest.fixture(scope="module")
def dev():
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    return pkg("device")


def _pre(panel):
    T, N, F = panel.feats.shape
    return O.rff(panel.feats.reshape(T * N, F), panel.W).reshape(T, N, -1), O.vol_scale_panel(panel)


def test_metric_size_n500_p1024_all_modes(dev, syn):
    lib = pkg("_lib").load()
    panel = syn.make_panel(N=500, T=15, p=1024, seed=51, ragged=0.01)
    rff_x, vs = _pre(panel)
    months = [13, 14]
    ref = [O.date_step(panel, t, rff_x, vs) for t in months]
    dp = dev.DevicePanel(panel)
    worst = {}
    try:
        for mode in (0, 1, 2):
            lib.pfml_set_emulation(mode)
            got = dev.date_steps(panel, months, dpanel=dp)
            for g, o in zip(got, ref):
                assert g["status"] == 0, (mode, g["status"])
                assert (g["ids"] == o["ids"]).all()
                for name in NAMES:
                    e = relerr(g[name], o[name])
                    worst[(mode, name)] = max(worst.get((mode, name), 0.0), e)
                    assert e < TOL, (mode, name, e)
                assert (g["denom"] == g["denom"].T).all() and (g["risk"] == g["risk"].T).all() and (g["tc"] == g["tc"].T).all()
                assert (g["risk"] + g["tc"] == g["denom"]).all()
    finally:
        lib.pfml_set_emulation(1)
    print("N=500 p=1024 max rel err vs oracle by (mode, output):", {k: f"{v:.1e}" for k, v in worst.items()})


def test_c3_three_bandwidths_p512(dev, syn):
    for g in (np.exp(-3.0), np.exp(-2.0), np.exp(-1.0)):
        panel = syn.make_panel(N=500, T=14, p=512, seed=52, g=float(g))
        rff_x, vs = _pre(panel)
        o = O.date_step(panel, 13, rff_x, vs)
        got = dev.date_steps(panel, [13])[0]
        assert got["status"] == 0
        for name in NAMES:
            assert relerr(got[name], o[name]) < TOL, (g, name, relerr(got[name], o[name]))


def test_c4_ridge_grid_p1024_15_lambda_28_years(syn):
    sc = pkg("pfml_search_coef")
    rng = np.random.default_rng(53)
    years = list(range(1993, 1993 + 28))
    T = 24 + 12 * len(years)                                 # 1990-01 ..: burn-in 1990-01 .. 1991-11, then one window per year
    month_ends = syn.month_end_range("1990-01", T)
    p_max = 1024
    P = p_max + 1
    r_t = rng.standard_normal((T, P)) * 1e-3
    A = rng.standard_normal((T, P, 6)) * 0.2
    diag = rng.uniform(0.5, 1.5, (T, P)) * 1e-2
    denom = np.einsum("tik,tjk->tij", A, A)
    denom[:, np.arange(P), np.arange(P)] += diag
    p_vec = [64, 128, 256, 512, 1024]
    l_vec = [0.0] + list(np.exp(np.linspace(-10, 10, 14)))
    ref = O.search_coef(r_t, denom, month_ends, years, p_vec, l_vec)
    got = sc.search_coef(r_t, denom, month_ends, years, p_vec, l_vec)
    bucket, _ = sc.window_buckets(month_ends, years)
    worst = 0.0
    for k, y in enumerate(years, start=1):
        for p in p_vec:
            for j in range(1, len(l_vec)):
                e = relerr(got[y][p][j], ref[y][p][j])
                worst = max(worst, e)
                assert e < TOL, (y, p, j, e)
        if y in (years[0], years[-1]):                       # lambda = 0: condition-aware bound (quirk Q14)
            m = (bucket >= 0) & (bucket <= k)
            D = denom[m].sum(0) / m.sum()
            for p in p_vec:
                fi = sc.pfml_feat_idx(p, p_max)
                cond = np.linalg.cond(D[np.ix_(fi, fi)])
                assert relerr(got[y][p][0], ref[y][p][0]) < 1e-8 * max(1.0, cond * 2.2e-16 / 1e-8), (y, p, cond)
    print(f"C4 grid ({len(years)} years x {len(p_vec)} p x {len(l_vec)} lambda): max rel err vs oracle (lambda > 0) {worst:.1e}")


def test_c5_universe_n3000_p64(dev, syn):
    z = np.load(os.path.join(GOLD, "oracle_n3000_p64.npz"))
    N, T, p, seed, month = (int(x) for x in z["params"])
    panel = syn.make_panel(N=N, T=T, p=p, seed=seed, ragged=0.01)
    assert np.allclose([panel.feats.sum(), panel.W.sum(), panel.lam.sum()], z["feats_checksum"], rtol=1e-14), \
        "the seeded panel differs from the one the fixture was made from"
    got = dev.date_steps(panel, [month])[0]
    assert got["status"] == 0
    assert (got["ids"] == z["ids"]).all()
    for name in ("r_tilde", "denom", "risk", "tc"):
        assert relerr(got[name], z[name]) < TOL, (name, relerr(got[name], z[name]))
    assert relerr(got["signal_t"][::8], z["signal_t_every8"]) < TOL


def test_second_reference_run_fixture_n150_p64(dev, syn):
    """CUDA vs the exec'd reference script at N = 150 (padded 192: uneven two-level block recursion), p = 64."""
    z = np.load(os.path.join(GOLD, "pfml_ref_mid.npz"))
    panel = panel_from_golden(z, syn)
    hot = [int(t) for t in z["hot"]]
    got = dev.date_steps(panel, hot)
    for k, g in enumerate(got):
        n = int(z["n_d"][k])
        assert g["status"] == 0
        for name in ("r_tilde", "denom", "risk", "tc"):
            assert relerr(g[name], z[name][k]) < TOL, (name, k)
        assert relerr(g["signal_t"], z["signal_t"][k, :n]) < TOL


def test_one_bad_date_in_a_healthy_batch(dev, syn):
    """A NaN characteristic in month 5 makes the standardised slabs of the dates whose 13-month look-back covers month 5
    non-finite.  Those dates come back NaN with status bit 4; every other date of the same batch has status 0 and is
    bit-identical to the same date computed in a batch without the bad data (INT8 path, n = 500)."""
    panel = syn.make_panel(N=500, T=22, p=32, seed=54)
    months = list(range(13, 22))
    clean = dev.date_steps(panel, months)
    panel.feats[5, 3, 7] = np.nan
    got = dev.date_steps(panel, months)
    for t, g, c in zip(months, got, clean):
        if t - panel.lb - 1 <= 5:                            # look-back months t-12 .. t cover month 5
            assert g["status"] & 16, (t, g["status"])
            for name in ("r_tilde", "denom", "risk", "tc"):
                assert np.isnan(g[name]).all(), (t, name)
        else:
            assert g["status"] == 0, (t, g["status"])
            for name in NAMES:
                assert np.array_equal(g[name], c[name]), (t, name)
