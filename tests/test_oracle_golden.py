"""CPU: the numpy oracle (oracle/pfml_oracle.py) against vectors frozen from the reference itself.

tests/golden/*.npz were produced by oracle/make_golden.py, which exec's the unmodified reference scripts
from /root/reference on a synthetic panel.  Tolerances: 1e-10 relative (the oracle restates the same
arithmetic in the same order; observed 1e-14)."""
import os

import numpy as np
import pytest

import pfml_oracle as O
from conftest import GOLD, panel_from_golden, relerr


@pytest.fixture(scope="module")
def gold():
    return np.load(os.path.join(GOLD, "pfml_ref_small.npz"), allow_pickle=False)


@pytest.fixture(scope="module")
def funcs():
    return np.load(os.path.join(GOLD, "funcs_ref.npz"))


def test_date_step_matches_reference_script(gold, syn):
    panel = panel_from_golden(gold, syn)
    T, N, F = panel.feats.shape
    rff_x = O.rff(panel.feats.reshape(T * N, F), panel.W).reshape(T, N, -1)
    vs = O.vol_scale_panel(panel)
    hot = gold["hot"]
    assert (hot == panel.hot_months()).all()
    for k in (0, len(hot) // 2, len(hot) - 1):
        o = O.date_step(panel, int(hot[k]), rff_x, vs)
        n = int(gold["n_d"][k])
        assert o["signal_t"].shape[0] == n
        for name in ("r_tilde", "denom", "risk", "tc"):
            assert relerr(o[name], gold[name][k]) < 1e-10, name
        assert relerr(o["signal_t"], gold["signal_t"][k, :n]) < 1e-10
        # analytic properties the reference implies
        assert relerr(o["risk"] + o["tc"], o["denom"]) < 1e-14
        assert relerr(o["denom"], o["denom"].T) < 1e-9
        assert np.linalg.eigvalsh(0.5 * (o["denom"] + o["denom"].T)).min() > -1e-10 * np.abs(o["denom"]).max()


def test_search_coef_matches_reference_script(gold):
    hot = gold["hot"]
    month_ends = gold["panel_month_ends"][hot]
    p_vec, l_vec, years = gold["p_vec"], gold["l_vec"], gold["years"]
    coef = O.search_coef(gold["r_tilde"], gold["denom"], month_ends, years, p_vec, l_vec)
    for y in years:
        for p in p_vec:
            ref = gold[f"coef_{y}_{p}"]
            for j, l in enumerate(l_vec):
                tol = 1e-9 if l > 0 else 1e-5        # lambda = 0 is condition-limited (quirk Q14)
                assert relerr(coef[int(y)][int(p)][j], ref[j]) < tol


def test_empty_window_raises(gold):
    hot = gold["hot"]
    month_ends = gold["panel_month_ends"][hot]
    with pytest.raises(ValueError):
        O.search_coef(gold["r_tilde"], gold["denom"], month_ends, [2030], gold["p_vec"], gold["l_vec"])


def test_m_func_known_answer(funcs):
    m = O.m_func(float(funcs["m_w"]), float(funcs["m_mu"]), float(funcs["m_rf"]), funcs["m_sigma"] * 10.0, 10.0,
                 np.diag(funcs["m_lam"]), 10)
    assert relerr(m, funcs["m_ref"]) < 1e-12
    ev = np.linalg.eigvals(m).real
    assert ev.min() > 0 and ev.max() < 1          # trading-speed matrix


def test_weighted_cov_cor_known_answer(funcs):
    assert relerr(O.weighted_cov_wt(funcs["wc_X"], funcs["wc_w"]), funcs["cov_ref"]) < 1e-13
    cor = O.weighted_cor_wt(funcs["wc_X"], funcs["wc_w"])
    assert relerr(cor, funcs["cor_ref"]) < 1e-13
    assert (np.diag(cor) == 1.0).all()


def test_ewma_vol_known_answer(funcs):
    ev = O.ewma_vol(funcs["ev_x"], float(funcs["ev_lam"]), 63)
    ref = funcs["ev_ref"]
    assert (np.isnan(ev) == np.isnan(ref)).all()
    assert np.nanmax(np.abs(ev - ref)) < 1e-15
    assert np.isnan(O.ewma_vol(funcs["ev_x"][:60], 0.99, 63)).all()      # n <= start


def test_feature_index_map_bit_exact(funcs):
    for p in (4, 8, 16):
        assert (O.pfml_feat_idx(p, 16) == funcs[f"feat_idx_{p}"]).all()


def _golden_coef(gold):
    return {int(y): {int(p): gold[f"coef_{y}_{p}"] for p in gold["p_vec"]} for y in gold["years"]}


def test_hp_reals_aim_backtest_match_reference(gold, syn):
    """Next rows: validation utility table, aim portfolio, back-test weights and statistics."""
    panel = panel_from_golden(gold, syn)
    hot = [int(t) for t in gold["hot"]]
    p_vec, l_vec, years = [int(p) for p in gold["p_vec"]], gold["l_vec"], [int(y) for y in gold["years"]]
    coef = _golden_coef(gold)
    v = O.hp_reals(gold["r_tilde"], gold["denom"], panel.month_ends[hot], coef, years, p_vec, l_vec)
    assert (v["l"].to_numpy() == gold["val_l"]).all() and (v["p"].to_numpy() == gold["val_p"]).all()
    assert (v["eom"].to_numpy().astype("datetime64[D]") == gold["val_eom"]).all()
    assert relerr(v["obj"].to_numpy(), gold["val_obj"]) < 1e-10
    assert relerr(v["cum_obj"].to_numpy(), gold["val_cum_obj"]) < 1e-10
    assert (v["rank"].to_numpy() == gold["val_rank"]).all()
    oos = [int(t) for t in gold["oos_idx"]]
    sig = {t: gold["signal_t"][hot.index(t), :int(gold["n_d"][hot.index(t)])] for t in oos}
    aims = O.aim_portfolios(v, coef, sig, panel.month_ends, oos, None)
    w_aim = {}
    for k, t in enumerate(oos):
        sel = np.flatnonzero(panel.valid[t])
        assert aims[t][0] == gold["aim_p"][k] and aims[t][1] == gold["aim_l"][k]
        assert relerr(aims[t][2], gold["w_aim"][k, sel]) < 1e-12
        w_aim[t] = gold["w_aim"][k, sel]
    w = O.pfml_w(panel, oos, w_aim)
    st = O.pf_stats(panel, w)
    for k, t in enumerate(oos):
        sel = np.flatnonzero(panel.valid[t])
        assert relerr(w[t]["w"], gold["bt_w"][k, sel]) < 1e-10
        assert relerr(w[t]["w_start"], gold["bt_w_start"][k, sel]) < 1e-10
        assert relerr(np.array([st[t][n] for n in ("inv", "shorting", "turnover", "r", "tc")]), gold["bt_stats"][k]) < 1e-10
    assert abs(w[oos[0]]["w_start"].sum() - 1.0) < 1e-14            # value-weighted start


def test_date_step_matches_second_reference_run_n150_p64(syn):
    """Second reference-run fixture (oracle/make_golden_mid.py): N = 150, p = 64, ragged, Barra ids missing; outputs of every
    hot date plus omega / omega_chg / m of the last date read from the exec'd script's namespace."""
    z = np.load(os.path.join(GOLD, "pfml_ref_mid.npz"))
    panel = panel_from_golden(z, syn)
    T, N, F = panel.feats.shape
    rff_x = O.rff(panel.feats.reshape(T * N, F), panel.W).reshape(T, N, -1)
    vs = O.vol_scale_panel(panel)
    hot = [int(t) for t in z["hot"]]
    for k, t in enumerate(hot):
        last = k == len(hot) - 1
        o = O.date_step(panel, t, rff_x, vs, keep=last)
        n = int(z["n_d"][k])
        for name in ("r_tilde", "denom", "risk", "tc"):
            assert relerr(o[name], z[name][k]) < 1e-10, (name, k)
        assert relerr(o["signal_t"], z["signal_t"][k, :n]) < 1e-10
        if last:
            assert relerr(o["omega"], z["omega_last"]) < 1e-10
            assert relerr(o["omega_chg"], z["omega_chg_last"]) < 1e-10
            assert relerr(o["m"], z["m_last"]) < 1e-10
