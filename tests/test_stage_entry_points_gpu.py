"""GPU: the stage-level entry points against the oracle's intermediate results and the reference's own.

  pfml_standardize_scale     vs oracle signals[tau] for every tau (PFML_Input_Data.py:357-388)
  pfml_omega_batched         vs oracle omega / omega_chg / m (:400-462, General_functions.py:919-962) and vs the omega,
                             omega_chg and m the exec'd reference script left in its namespace (pfml_ref_mid.npz)
  pfml_rtilde_denom_batched  a9 fed with the ORACLE's omega vs oracle r_tilde / risk / tc / denom (:470-481)
  pfml_date_steps_bucketed   fused expanding-window sums: bit-identical to pfml_bucket_accumulate on the outputs

Without these, a compensating error between the chain and the solves could hide behind r_tilde / denom.
Tolerance 1e-8 relative.
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


def _pre(panel):
    T, N, F = panel.feats.shape
    return O.rff(panel.feats.reshape(T * N, F), panel.W).reshape(T, N, -1), O.vol_scale_panel(panel)


CASES = [dict(N=150, T=15, p=64, seed=61, ragged=0.05, barra_missing=0.2),      # DMMA a9, ragged, 3 blocks
         dict(N=500, T=14, p=128, seed=62, ragged=0.02)]                         # INT8 path end to end (padded 512)


@pytest.mark.parametrize("case", CASES, ids=["n150_p64", "n500_p128"])
def test_stages_vs_oracle(dev, syn, case):
    panel = syn.make_panel(**case)
    rff_x, vs = _pre(panel)
    months = [12, 13]
    ref = [O.date_step(panel, t, rff_x, vs, keep=True) for t in months]
    dp = dev.DevicePanel(panel)

    sig = dev.stage_signals(panel, months, dpanel=dp)
    for s, o in zip(sig, ref):
        for tau in range(panel.lb + 2):
            assert relerr(s[tau], o["signals"][tau]) < TOL, ("signals", tau)

    om = dev.stage_omega(panel, months, dpanel=dp)
    for g, o in zip(om, ref):
        assert g["status"] == 0
        for name in ("omega", "omega_chg", "m"):
            assert relerr(g[name], o[name]) < TOL, (name, relerr(g[name], o[name]))

    a9 = dev.stage_rtilde_denom(panel, months, [o["omega"] for o in ref], [o["omega_chg"] for o in ref], dpanel=dp)
    for g, o in zip(a9, ref):
        assert g["status"] == 0
        for name in ("r_tilde", "denom", "risk", "tc"):
            assert relerr(g[name], o[name]) < TOL, (name, relerr(g[name], o[name]))
        assert (g["denom"] == g["denom"].T).all()


def test_omega_and_m_vs_reference_script_namespace(dev, syn):
    z = np.load(os.path.join(GOLD, "pfml_ref_mid.npz"))
    panel = panel_from_golden(z, syn)
    last = int(z["hot"][-1])
    g = dev.stage_omega(panel, [last])[0]
    assert g["status"] == 0
    assert relerr(g["omega"], z["omega_last"]) < TOL
    assert relerr(g["omega_chg"], z["omega_chg_last"]) < TOL
    assert relerr(g["m"], z["m_last"]) < TOL
    a9 = dev.stage_rtilde_denom(panel, [last], [z["omega_last"]], [z["omega_chg_last"]])[0]
    for name in ("r_tilde", "denom", "risk", "tc"):
        assert relerr(a9[name], z[name][-1]) < TOL, name


def test_a9_stage_q8_nan_omega_chg(dev, syn):
    """omega_chg with NaN rows (quirk Q8) handed to the a9 stage: tc and denom all NaN, risk and r_tilde unaffected."""
    panel = syn.make_panel(N=500, T=14, p=16, seed=63)
    rff_x, vs = _pre(panel)
    o = O.date_step(panel, 13, rff_x, vs, keep=True)
    omc = o["omega_chg"].copy()
    omc[7] = np.nan
    g = dev.stage_rtilde_denom(panel, [13], [o["omega"]], [omc])[0]
    assert g["status"] == 0
    assert np.isnan(g["tc"]).all() and np.isnan(g["denom"]).all()
    assert relerr(g["risk"], o["risk"]) < TOL and relerr(g["r_tilde"], o["r_tilde"]) < TOL


@pytest.mark.parametrize("n,p,mode", [(500, 64, 1), (100, 32, 0)], ids=["int8_fused_finish", "dmma_separate_pass"])
def test_bucketed_date_steps_bit_exact(dev, syn, n, p, mode):
    import torch
    sc = pkg("pfml_search_coef")
    lib = pkg("_lib").load()
    lib.pfml_set_emulation(mode)
    try:
        _bucketed_case(dev, syn, sc, torch, n, p)
    finally:
        lib.pfml_set_emulation(1)


def _bucketed_case(dev, syn, sc, torch, n, p):
    panel = syn.make_panel(N=n, T=22, p=p, seed=64, ragged=0.02)
    months = list(range(13, 22))
    dp = dev.DevicePanel(panel)
    sel, n_sel, n_max, NP = dev.universe(panel, months)
    run = dev.DateStepRunner(dp, len(months), n_max)
    run.set_batch(months, sel, n_sel)
    run.run()
    torch.cuda.synchronize()
    plain = {k: getattr(run, k).clone() for k in ("r_tilde", "denom", "risk", "tc", "signal_t")}
    bucket = np.array([0, 0, 2, -1, 1, 1, 2, 5, 0], dtype=np.int32)          # unordered windows, one date outside, one index too large
    nb, P = 3, p + 1
    r_ref, d_ref = sc.bucket_sums(plain["r_tilde"], plain["denom"], np.where(bucket < nb, bucket, -1), nb)
    r_sum = torch.full((nb, P), 1.5, dtype=torch.float64, device=dp.device)      # accumulated INTO: start from a non-zero value
    d_sum = torch.full((nb, P, P), -0.25, dtype=torch.float64, device=dp.device)
    run.run_bucketed(torch.as_tensor(bucket).to(dp.device), nb, r_sum, d_sum)
    torch.cuda.synchronize()
    for k, v in plain.items():
        assert torch.equal(getattr(run, k), v), k                              # the outputs themselves are unchanged
    assert torch.equal(r_sum, r_ref + 1.5)
    assert torch.equal(d_sum, d_ref - 0.25)
    assert int(run.status.abs().sum()) == 0


def test_multi_g_shares_m_and_chain_bit_identical(dev, syn):
    """pfml_date_steps_multi_g: three bandwidths of the same dates with m, the chains and the inverses computed once
    (quirk Q10).  Every g's outputs and window sums equal the single-g call on that g's panel BIT FOR BIT, in the INT8
    direct form (n = 500) and on the per-g fallback (FP64 DMMA mode)."""
    import dataclasses
    import torch
    lib = pkg("_lib").load()
    panel = syn.make_panel(N=500, T=18, p=32, seed=65, ragged=0.02)
    rng = np.random.default_rng(66)
    Ws = [panel.W, 0.3 * rng.standard_normal(panel.W.shape), 0.6 * rng.standard_normal(panel.W.shape)]
    months = [13, 14, 15, 16]
    sel, n_sel, n_max, NP = dev.universe(panel, months)
    bucket = torch.as_tensor(np.array([0, 1, 1, 0], dtype=np.int32)).cuda()
    P = panel.p + 1
    for mode in (1, 0):
        lib.pfml_set_emulation(mode)
        try:
            dp = dev.DevicePanel(panel, extra_W=Ws[1:])
            mg = dev.MultiGRunner(dp, len(months), n_max, want_risk_tc=True, want_signal=True)
            mg.set_batch(months, sel, n_sel)
            r_sum = torch.zeros((3, 2, P), dtype=torch.float64, device="cuda")
            d_sum = torch.zeros((3, 2, P, P), dtype=torch.float64, device="cuda")
            mg.run(bucket, 2, r_sum, d_sum)
            torch.cuda.synchronize()
            assert int(mg.status.abs().sum()) == 0
            for g, W in enumerate(Ws):
                one = dev.DateStepRunner(dev.DevicePanel(dataclasses.replace(panel, W=W)), len(months), n_max)
                one.set_batch(months, sel, n_sel)
                rs, ds = torch.zeros((2, P), dtype=torch.float64, device="cuda"), torch.zeros((2, P, P), dtype=torch.float64, device="cuda")
                one.run_bucketed(bucket, 2, rs, ds)
                torch.cuda.synchronize()
                for name in ("r_tilde", "denom", "risk", "tc", "signal_t"):
                    assert torch.equal(getattr(mg, name)[g], getattr(one, name)), (mode, g, name)
                assert torch.equal(r_sum[g], rs) and torch.equal(d_sum[g], ds), (mode, g)
        finally:
            lib.pfml_set_emulation(1)
    # and one g against the oracle
    pg = dataclasses.replace(panel, W=Ws[2])
    rff_x, vs = _pre(pg)
    o = O.date_step(pg, 14, rff_x, vs)
    for name in ("r_tilde", "denom", "risk", "tc"):
        assert relerr(getattr(mg, name)[2, 1].cpu().numpy(), o[name]) < TOL, name
