"""GPU: the CUDA path (through the C ABI) against the numpy oracle and the reference-generated golden vectors.

Tolerance: 1e-8 relative (BASELINE.json north_star) on r_tilde, denom, risk, tc, signal_t, m; observed ~1e-13.
"""
import os

import numpy as np
import pytest

import pfml_oracle as O
from conftest import GOLD, panel_from_golden, pkg, relerr

pytestmark = pytest.mark.gpu
TOL = 1e-8


@pytest.fixture(scope="module")
def dev():
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    return pkg("device")


def test_dgemm_all_layouts(dev):
    rng = np.random.default_rng(0)
    for (M, N, K) in ((64, 64, 64), (130, 70, 33), (257, 129, 100), (500, 1025, 512), (25, 300, 515)):
        for ta in (False, True):
            for tb in (False, True):
                A = rng.standard_normal((3, K, M) if ta else (3, M, K))
                B = rng.standard_normal((3, N, K) if tb else (3, K, N))
                ref = np.matmul(A.transpose(0, 2, 1) if ta else A, B.transpose(0, 2, 1) if tb else B)
                got = dev.dgemm(A, B, ta, tb, alpha=0.5)
                assert relerr(got, 0.5 * ref) < 1e-13, (M, N, K, ta, tb)


def test_inverse(dev):
    rng = np.random.default_rng(1)
    for n in (1, 17, 64, 100, 192, 500):
        A = rng.standard_normal((4, n, n)) * (0.5 / np.sqrt(n)) + 2.0 * np.eye(n)      # cond ~ 2, like x + y - m o sigma_gr
        X, st = dev.inverse(A)
        assert (st == 0).all()
        assert relerr(X, np.linalg.inv(A)) < 1e-11, n
    X, st = dev.inverse(np.zeros((1, 64, 64)))
    assert st[0] & 1                       # singular pivot reported, not hidden


def test_m_func_vs_oracle_and_golden(dev, syn):
    f = np.load(os.path.join(GOLD, "funcs_ref.npz"))
    m, st = dev.m_func_batched(float(f["m_w"]), float(f["m_mu"]), float(f["m_rf"]), f["m_sigma"][None] * 10.0, 10.0,
                               f["m_lam"][None], 10)
    assert st[0] == 0
    assert relerr(m[0], f["m_ref"]) < TOL          # golden vector from the reference's own m_func
    pn = syn.make_panel(N=200, T=14, p=8, seed=4)
    sig = np.stack([O.create_cov(pn.fct_load[t], pn.fct_cov[t], pn.ivol[t]) for t in (12, 13)])
    m, st = dev.m_func_batched(pn.wealth[[12, 13]], pn.mu, pn.rf[[12, 13]], sig * 10.0, 10.0, pn.lam[[12, 13]], 10)
    for k, t in enumerate((12, 13)):
        ref = O.m_func(pn.wealth[t], pn.mu, pn.rf[t], sig[k] * 10.0, 10.0, np.diag(pn.lam[t]), 10)
        assert relerr(m[k], ref) < TOL
        ev = np.linalg.eigvals(m[k]).real
        assert 0 < ev.min() and ev.max() < 1


def _check_dates(got, panel, months, rff_x, vs, tol=TOL):
    for g, t in zip(got, months):
        o = O.date_step(panel, int(t), rff_x, vs)
        assert g["status"] == 0
        assert (g["ids"] == o["ids"]).all()
        for name in ("r_tilde", "denom", "risk", "tc", "signal_t"):
            assert relerr(g[name], o[name]) < tol, (name, t, relerr(g[name], o[name]))
        assert (g["denom"] == g["denom"].T).all()           # mirrored epilogue: exactly symmetric
        assert relerr(g["risk"] + g["tc"], g["denom"]) < 1e-15


def _panel_pre(panel):
    T, N, F = panel.feats.shape
    return O.rff(panel.feats.reshape(T * N, F), panel.W).reshape(T, N, -1), O.vol_scale_panel(panel)


def test_date_steps_golden_reference_run(dev, syn):
    """CUDA vs the outputs of the exec'd reference script (ragged universe, Barra median fill)."""
    z = np.load(os.path.join(GOLD, "pfml_ref_small.npz"))
    panel = panel_from_golden(z, syn)
    hot = [int(t) for t in z["hot"]]
    got = dev.date_steps(panel, hot)
    for k, g in enumerate(got):
        n = int(z["n_d"][k])
        for name in ("r_tilde", "denom", "risk", "tc"):
            assert relerr(g[name], z[name][k]) < TOL, (name, k)
        assert relerr(g["signal_t"], z["signal_t"][k, :n]) < TOL


def test_date_steps_vs_oracle_balanced(dev, syn):
    panel = syn.make_panel(N=100, T=16, p=32, seed=21)
    rff_x, vs = _panel_pre(panel)
    months = [12, 13, 14]
    _check_dates(dev.date_steps(panel, months), panel, months, rff_x, vs)


def test_date_steps_vs_oracle_ragged_medianfill(dev, syn):
    panel = syn.make_panel(N=150, T=15, p=64, seed=22, ragged=0.05, barra_missing=0.2)
    rff_x, vs = _panel_pre(panel)
    months = [12, 13]
    _check_dates(dev.date_steps(panel, months), panel, months, rff_x, vs)


def test_date_steps_n500_p128(dev, syn):
    panel = syn.make_panel(N=500, T=14, p=128, seed=23, ragged=0.02)
    rff_x, vs = _panel_pre(panel)
    _check_dates(dev.date_steps(panel, [12]), panel, [12], rff_x, vs)


def test_date_steps_n700_uneven_recursion(dev, syn):
    """n = 700 -> NP = 704 = 11 blocks of 64: uneven splits of the recursive block inverse (352 + 352 -> 192 + 160 ...)."""
    panel = syn.make_panel(N=700, T=14, p=16, seed=28, ragged=0.01)
    rff_x, vs = _panel_pre(panel)
    _check_dates(dev.date_steps(panel, [12, 13]), panel, [12, 13], rff_x, vs)


def test_nan_growth_semantics_q8(dev, syn):
    """NaN tr_ld0 in look-back months -> g = 1 (PFML_Input_Data.py:405); at date d it propagates to tc (:458)."""
    panel = syn.make_panel(N=80, T=15, p=16, seed=24, nan_tr=0.03)
    panel.tr_ld0[13] = np.nan_to_num(panel.tr_ld0[13], nan=0.01)       # date 13 itself clean
    rff_x, vs = _panel_pre(panel)
    _check_dates(dev.date_steps(panel, [13]), panel, [13], rff_x, vs)
    panel.tr_ld0[13, 5] = np.nan
    g = dev.date_steps(panel, [13])[0]
    o = O.date_step(panel, 13, rff_x, vs)
    assert np.isnan(o["tc"]).all() and np.isnan(g["tc"]).all()
    assert relerr(g["risk"], o["risk"]) < TOL


def test_nan_growth_semantics_q8_n500(dev, syn):
    """The same at n = 500 (padded 512), where risk / tc run on the INT8 tensor cores: a NaN growth rate at date d is
    not an error (status 0) and makes every element of tc and denom NaN, as in the reference's dense product."""
    panel = syn.make_panel(N=500, T=14, p=16, seed=26)
    panel.tr_ld0[13, 7] = np.nan
    rff_x, vs = _panel_pre(panel)
    g = dev.date_steps(panel, [13])[0]
    o = O.date_step(panel, 13, rff_x, vs)
    assert g["status"] == 0
    assert np.isnan(o["tc"]).all() and np.isnan(g["tc"]).all() and np.isnan(g["denom"]).all()
    assert relerr(g["risk"], o["risk"]) < TOL


def test_rff_and_vol_scale(dev, syn):
    panel = syn.make_panel(N=90, T=14, p=64, seed=25, barra_missing=0.3, ragged=0.3)
    dp = dev.DevicePanel(panel)
    rff_x, vs = _panel_pre(panel)
    assert relerr(dp.rff.cpu().numpy().reshape(rff_x.shape), rff_x) < 1e-14
    assert relerr(dp.vol_scale.cpu().numpy().reshape(vs.shape), vs) < 1e-14


def test_host_entry_point_matches_device_path(dev, syn):
    host = pkg("host")
    panel = syn.make_panel(N=100, T=18, p=32, seed=26, ragged=0.04, barra_missing=0.1)
    months = [13, 14, 15, 16]
    a = dev.date_steps(panel, months)
    for chunk in (0, 3):
        b = host.date_steps_host(panel, months, chunk=chunk)
        for x, y in zip(a, b):
            for name in ("r_tilde", "denom", "risk", "tc", "signal_t"):
                assert relerr(y[name], x[name]) < 1e-13, (name, chunk)


def test_host_entry_points_bit_identical_to_resident_path(dev, syn):
    """Full-square entry (packed transfer + host rebuild, several chunks / unpack threads), packed entry + pfml_unpack_reals
    and the old full-square download all return exactly the bits the device-resident path holds."""
    import torch
    host = pkg("host")
    panel = syn.make_panel(N=100, T=24, p=64, seed=29, ragged=0.04)
    months = list(range(13, 24))
    a = dev.date_steps(panel, months)
    blk = host.HostBlock(panel, months, pin=True)
    for chunk, rebuild in ((0, "1"), (4, "1"), (4, "0")):
        os.environ["PFML_HOST_REBUILD"] = rebuild          # packed transfer + host rebuild / plain download of the squares
        for v in blk.out.values():
            v.fill_(float("nan"))
        blk.run(chunk=chunk)
        assert int(blk.status.abs().sum()) == 0
        for b, x in enumerate(a):
            n = x["signal_t"].shape[0]
            for name in ("r_tilde", "denom", "risk", "tc"):
                assert np.array_equal(blk.out[name][b].numpy(), x[name]), (name, b, chunk)
            assert np.array_equal(blk.out["signal_t"][b, :n].numpy(), x["signal_t"])
    # r_tilde + denom only: one triangle travels and is mirrored
    os.environ["PFML_HOST_REBUILD"] = "1"
    blk.out["denom"].fill_(float("nan"))
    blk.run(chunk=3, full_outputs=False)
    os.environ.pop("PFML_HOST_REBUILD")
    for b, x in enumerate(a):
        assert np.array_equal(blk.out["denom"][b].numpy(), x["denom"])
    blk.run_packed(chunk=5)
    for b, x in enumerate(a):
        risk, tc, denom = blk.unpack(b)
        assert np.array_equal(risk, x["risk"]) and np.array_equal(tc, x["tc"]) and np.array_equal(denom, x["denom"])
    assert blk.d2h_bytes < blk.d2h_bytes_full_squares


def test_host_entry_point_validates_before_upload(dev, syn):
    host = pkg("host")
    L = pkg("_lib")
    panel = syn.make_panel(N=70, T=16, p=8, seed=30)
    for field, bad in (("date_t", 3), ("n_sel", 0), ("n_sel", 10 ** 6), ("sel", -1), ("sel", 70)):
        blk = host.HostBlock(panel, [13, 14], pin=False)
        getattr(blk, field).reshape(-1)[0] = bad
        with pytest.raises(L.PfmlError, match="code -2"):
            blk.run()
    blk = host.HostBlock(panel, [13, 14], pin=False)
    blk.struct.K = 40
    with pytest.raises(L.PfmlError, match="code -2"):
        blk.run()


def test_full_size_properties(dev, syn):
    """BASELINE metric size (N=500, p=1024): properties that need no oracle run."""
    panel = syn.make_panel(N=500, T=14, p=1024, seed=27)
    g = dev.date_steps(panel, [12, 13])
    for o in g:
        assert o["status"] == 0
        n = o["signal_t"].shape[0]
        assert (o["denom"] == o["denom"].T).all()
        assert relerr(o["risk"] + o["tc"], o["denom"]) < 1e-15
        ev = np.linalg.eigvalsh(o["denom"])
        assert ev.min() > -1e-9 * ev.max()
        # constant column = 1/sqrt(n) / vol (quirk Q4); standardised columns: sum of (vol * col) = 0, l2 = 1
        vs = O.vol_scale_panel(panel)
    t = 13
    s = g[1]["signal_t"] * vs[t][:, None]
    assert np.allclose(s[:, 0], 1 / np.sqrt(500), rtol=1e-12)
    assert np.abs(s[:, 1:].sum(0)).max() < 1e-12
    assert np.allclose((s ** 2).sum(0), 1.0, rtol=1e-12)
