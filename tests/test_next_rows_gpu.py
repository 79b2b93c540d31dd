"""GPU: next rows -- validation utility, aim portfolio, back-test weights and statistics -- against the
vectors frozen from the reference scripts / functions (tests/golden/pfml_ref_small.npz).  Tolerance 1e-8 relative."""
import os

import numpy as np
import pytest

import pfml_oracle as O
from conftest import GOLD, panel_from_golden, pkg, relerr

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def gold():
    return np.load(os.path.join(GOLD, "pfml_ref_small.npz"))


def _coef(gold):
    return {int(y): {int(p): gold[f"coef_{y}_{p}"] for p in gold["p_vec"]} for y in gold["years"]}


def test_validation_table(gold, syn):
    panel = panel_from_golden(gold, syn)
    hot = [int(t) for t in gold["hot"]]
    v = pkg("pfml_hp_reals").validation(gold["r_tilde"], gold["denom"], panel.month_ends[hot], _coef(gold),
                                        [int(y) for y in gold["years"]], [int(p) for p in gold["p_vec"]], gold["l_vec"])
    assert (v["l"].to_numpy() == gold["val_l"]).all() and (v["p"].to_numpy() == gold["val_p"]).all()
    assert (v["hp_end"].to_numpy() == gold["val_hp_end"]).all()
    assert (v["eom"].to_numpy().astype("datetime64[D]") == gold["val_eom"]).all()
    assert relerr(v["obj"].to_numpy(), gold["val_obj"]) < 1e-8
    assert relerr(v["cum_obj"].to_numpy(), gold["val_cum_obj"]) < 1e-8
    assert (v["rank"].to_numpy() == gold["val_rank"]).all()


def test_aim_and_backtest(gold, syn):
    panel = panel_from_golden(gold, syn)
    hot = [int(t) for t in gold["hot"]]
    oos = [int(t) for t in gold["oos_idx"]]
    years, p_vec = [int(y) for y in gold["years"]], [int(p) for p in gold["p_vec"]]
    coef = _coef(gold)
    v = O.hp_reals(gold["r_tilde"], gold["denom"], panel.month_ends[hot], coef, years, p_vec, gold["l_vec"])
    dev = pkg("device")
    dp = dev.DevicePanel(panel)
    outs = dev.date_steps(panel, oos, dpanel=dp)                       # signal_t of the OOS months from the device path
    NP = max(o["signal_t"].shape[0] for o in outs)
    P = panel.p + 1
    S = np.zeros((len(oos), NP, P))
    n_sel = np.array([o["signal_t"].shape[0] for o in outs], dtype=np.int32)
    for b, o in enumerate(outs):
        S[b, :n_sel[b]] = o["signal_t"]
    w_aim, ps, ls = pkg("pfml_aim_fun").aim_portfolios(v, coef, S, n_sel, panel.month_ends[oos])
    assert (ps == gold["aim_p"]).all() and (ls == gold["aim_l"]).all()
    for b, t in enumerate(oos):
        sel = np.flatnonzero(panel.valid[t])
        assert relerr(w_aim[b, :n_sel[b]], gold["w_aim"][b, sel]) < 1e-8
    res = pkg("pfml_best_hps").pfml_w(panel, oos, [gold["w_aim"][b, np.flatnonzero(panel.valid[t])] for b, t in enumerate(oos)],
                                      dpanel=dp)
    assert (res["status"] == 0).all()
    for b, t in enumerate(oos):
        sel = np.flatnonzero(panel.valid[t])
        assert (res["ids"][b] == panel.ids[sel]).all()
        assert relerr(res["w_start"][b], gold["bt_w_start"][b, sel]) < 1e-8, b
        assert relerr(res["w"][b], gold["bt_w"][b, sel]) < 1e-8, b
    assert relerr(res["stats"], gold["bt_stats"]) < 1e-8
    # trading-speed matrices of the batch equal the dense m_func (oracle) for one month
    t = oos[3]
    sel = np.flatnonzero(panel.valid[t])
    sig = O.create_cov(panel.fct_load[t][sel], panel.fct_cov[t], panel.ivol[t, sel])
    m_ref = O.m_func(panel.wealth[t], panel.mu, panel.rf[t], sig * panel.gamma_rel, panel.gamma_rel, np.diag(panel.lam[t, sel]), 10)
    assert relerr(res["m"][3, :len(sel), :len(sel)].cpu().numpy(), m_ref) < 1e-8


def test_backtest_nan_drift_starts_at_zero(syn):
    """A missing tr_ld1 for a held stock: the reference writes the NaN drifted weight into next month's w_start and then
    sets NaN w_start to 0 (PFML_best_hps.py:199-210), so the stock restarts at 0 and nothing else is contaminated."""
    panel = syn.make_panel(N=60, T=20, p=8, seed=81, ragged=0.05)
    oos = [13, 14, 15, 16]
    rng = np.random.default_rng(82)
    w_aim = [rng.standard_normal(int(panel.valid[t].sum())) * 0.01 for t in oos]
    held = np.flatnonzero(panel.valid[13] & panel.valid[14])[4]
    panel.tr_ld1[13, held] = np.nan
    ref = O.pfml_w(panel, oos, {t: w for t, w in zip(oos, w_aim)})
    res = pkg("pfml_best_hps").pfml_w(panel, oos, w_aim)
    for b, t in enumerate(oos):
        assert np.isfinite(res["w"][b]).all() and np.isfinite(res["w_start"][b]).all()
        assert relerr(res["w_start"][b], ref[t]["w_start"]) < 1e-8 and relerr(res["w"][b], ref[t]["w"]) < 1e-8
    j = list(np.flatnonzero(panel.valid[14])).index(held)
    assert res["w_start"][1][j] == 0.0
