"""Generate tests/golden/*.npz by running the reference itself (build container only; needs /root/reference).

TEST INFRASTRUCTURE.  Usage:  python oracle/make_golden.py
Writes
  tests/golden/pfml_ref_small.npz   panel inputs + outputs of the exec'd PFML_Input_Data.py / PFML_Search_Coef.py
  tests/golden/funcs_ref.npz        known-answer vectors of the reference's importable functions
and prints the oracle-vs-reference deviations (the pinning check that tests/test_oracle_golden.py repeats).
"""
from __future__ import annotations

import importlib
import os
import sys
import tempfile

import numpy as np
import pandas as pd

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
PKG = "jkmp22-machine-learning-and-the-implementable-efficient-frontier-replication_b200"

import pfml_oracle as O          # noqa: E402
import ref_harness as H          # noqa: E402


def panel_to_dict(panel):
    keys = ["ids", "month_ends", "feats", "valid", "ret_ld1", "tr_ld0", "mu_ld0", "lam", "wealth", "rf",
            "fct_load", "fct_cov", "ivol", "barra_has", "W", "me", "tr_ld1", "mu_ld1"]
    d = {"panel_" + k: getattr(panel, k) for k in keys}
    d["panel_scalars"] = np.array([panel.gamma_rel, panel.mu, panel.lb, panel.iterations, panel.meta["g"]])
    return d


def relerr(a, b):
    return float(np.max(np.abs(a - b)) / max(np.max(np.abs(b)), 1e-300))


def main():
    syn = importlib.import_module(PKG + ".synthetic")
    gold = os.path.join(ROOT, "tests", "golden")
    os.makedirs(gold, exist_ok=True)

    # ---------------- script-level golden: PFML_Input_Data.py .. PFML_best_hps.py::pfml_w -----------------
    p_vec = [4, 8, 16]
    l_vec = [0.0, float(np.exp(-10)), 1e-2, 1.0]
    panel = syn.make_panel(N=20, T=60, p=16, seed=11, ragged=0.08, barra_missing=0.1, start="2010-01", quantize=True)
    with tempfile.TemporaryDirectory() as tmp:
        out, glob = H.run_pfml_input(panel, tmp + "/", p_vec, l_vec, start_year=2013, end_yr=2014)
        coef = H.run_search_coef(panel, tmp + "/", p_vec, l_vec, start_year=2013, end_yr=2014)
        validation = H.run_hp_reals(panel, tmp + "/", p_vec, l_vec, start_year=2013, end_yr=2014)
        aims = H.run_aim_fun(panel, tmp + "/", p_vec, l_vec, start_year=2013, end_yr=2014, split_years=1)
        # back-test functions of PFML_best_hps.py on the aims of g = 0
        bt = H.backtest_functions()
        import sqlite3, pickle
        con = sqlite3.connect(os.path.join(tmp, "Data", "JKP_US_SP500.db"))
        chars = pd.read_sql_query("SELECT id, eom, me, lambda, tr_ld1, ret_ld1, valid FROM Factors_processed", con, parse_dates=["eom"])
        con.close()
        chars["valid"] = chars["valid"].astype(bool)
        dates_oos = sorted(aims.keys())
        with open(os.path.join(tmp, "Data", "Barra_Cov.pkl"), "rb") as f:
            cov_list = pickle.load(f)
        lambda_list = {d: chars[chars["eom"] == d].sort_values("id").set_index("id")["lambda"].to_dict() for d in dates_oos}
        me = pd.to_datetime(panel.month_ends)
        risk_free = pd.DataFrame({"eom": me, "rf": panel.rf})
        wealth = pd.DataFrame({"eom": me, "wealth": panel.wealth, "mu_ld1": panel.mu_ld1})
        aims_df = pd.concat([aims[d]["aim_pf"] for d in dates_oos], ignore_index=True)
        w_ref = bt["pfml_w"](data=chars[(chars["eom"].isin(dates_oos)) & (chars["valid"])], dates=dates_oos, cov_list=cov_list,
                             lambda_list=lambda_list, gamma_rel=panel.gamma_rel, iterations=10, risk_free=risk_free,
                             wealth=wealth, mu=panel.mu, aims=aims_df)
        pf_ref = bt["pf_ts_fun"](weights=w_ref, data=chars, wealth=wealth, gam=panel.gamma_rel)
    dates = sorted(out["reals"].keys())
    hot = panel.hot_months()
    assert len(dates) == len(hot)
    assert all(pd.Timestamp(panel.month_ends[t]) == d for t, d in zip(hot, dates))
    P = panel.p + 1
    r_t = np.stack([out["reals"][d]["r_tilde"].to_numpy() for d in dates])
    den = np.stack([out["reals"][d]["denom"].to_numpy() for d in dates])
    risk = np.stack([out["reals"][d]["risk"].to_numpy() for d in dates])
    tc = np.stack([out["reals"][d]["tc"].to_numpy() for d in dates])
    n_max = panel.N
    sig = np.full((len(dates), n_max, P), np.nan)
    n_d = np.zeros(len(dates), dtype=np.int64)
    for k, d in enumerate(dates):
        s = out["signal_t"][d].to_numpy()
        n_d[k] = s.shape[0]
        sig[k, :s.shape[0]] = s
    # feature labels must be feat_all
    GF = H._import_general_functions()
    assert list(out["reals"][dates[0]]["r_tilde"].index) == GF.pfml_feat_fun(panel.p)

    # oracle vs reference (pinning)
    T, N, F = panel.feats.shape
    rff_x = O.rff(panel.feats.reshape(T * N, F), panel.W).reshape(T, N, -1)
    vs = O.vol_scale_panel(panel)
    worst = {}
    o_rt, o_den, o_sig = [], [], {}
    for k, t in enumerate(hot):
        o = O.date_step(panel, t, rff_x, vs)
        o_rt.append(o["r_tilde"]); o_den.append(o["denom"]); o_sig[int(t)] = o["signal_t"]
        for name, ref in (("r_tilde", r_t[k]), ("denom", den[k]), ("risk", risk[k]), ("tc", tc[k]),
                          ("signal_t", sig[k, :n_d[k]])):
            worst[name] = max(worst.get(name, 0.0), relerr(o[name], ref))
    print("oracle vs exec'd PFML_Input_Data.py, max rel err:", worst)
    years = sorted(coef.keys())
    o_coef = O.search_coef(np.stack(o_rt), np.stack(o_den), panel.month_ends[hot], years, p_vec, l_vec)
    cw = 0.0
    coef_arrs = {}
    ref_coef = {}
    for y in years:
        ref_coef[y] = {}
        for p in p_vec:
            ref_b = np.stack([coef[y][p][j].to_numpy() for j in range(len(l_vec))])
            coef_arrs[f"coef_{y}_{p}"] = ref_b
            ref_coef[y][p] = ref_b
            for j in range(1, len(l_vec)):          # lambda=0 is condition-limited (quirk Q14)
                cw = max(cw, relerr(o_coef[y][p][j], ref_b[j]))
            assert list(coef[y][p][0].index) == GF.pfml_feat_fun(p)
    print("oracle vs exec'd PFML_Search_Coef.py (lambda>0), max rel err:", cw)

    # f1: validation table (the reference's own coefficients in, so lambda = 0 noise does not enter)
    v_or = O.hp_reals(r_t, den, panel.month_ends[hot], ref_coef, years, p_vec, l_vec)
    v_ref = validation.sort_values(["p", "l", "eom_ret"]).reset_index(drop=True)
    assert (v_or[["l", "p", "hp_end"]].to_numpy() == v_ref[["l", "p", "hp_end"]].to_numpy()).all()
    assert (v_or["eom_ret"].to_numpy() == v_ref["eom_ret"].to_numpy()).all()
    print("oracle vs exec'd PFML_hp_reals.py: obj", relerr(v_or["obj"].to_numpy(), v_ref["obj"].to_numpy()),
          "cum_obj", relerr(v_or["cum_obj"].to_numpy(), v_ref["cum_obj"].to_numpy()),
          "rank equal", bool((v_or["rank"].to_numpy() == v_ref["rank"].to_numpy()).all()))
    # f2: aim portfolios
    oos_idx = [int(np.flatnonzero(panel.month_ends == np.datetime64(d.date()))[0]) for d in dates_oos]
    a_or = O.aim_portfolios(v_ref, ref_coef, {t: sig[list(hot).index(t), :n_d[list(hot).index(t)]] for t in oos_idx},
                            panel.month_ends, oos_idx, None)
    aw = 0.0
    aim_p, aim_l = [], []
    w_aim = np.full((len(oos_idx), N), np.nan)
    for k, (t, d) in enumerate(zip(oos_idx, dates_oos)):
        ref_w = aims[d]["aim_pf"]["w_aim"].to_numpy()
        sel = np.flatnonzero(panel.valid[t])
        assert list(aims[d]["aim_pf"]["id"]) == list(panel.ids[sel])
        w_aim[k, sel] = ref_w
        aim_p.append(a_or[t][0]); aim_l.append(a_or[t][1])
        assert len(aims[d]["coef"]) == a_or[t][0] + 1
        aw = max(aw, relerr(a_or[t][2], ref_w))
    print("oracle vs exec'd PFML_aim_fun.py: w_aim", aw)
    # f3: back-test weights and statistics
    w_or = O.pfml_w(panel, oos_idx, {t: w_aim[k, np.flatnonzero(panel.valid[t])] for k, t in enumerate(oos_idx)})
    st_or = O.pf_stats(panel, w_or)
    ww = 0.0
    w_out = np.full((len(oos_idx), N), np.nan); ws_out = np.full((len(oos_idx), N), np.nan)
    for k, (t, d) in enumerate(zip(oos_idx, dates_oos)):
        r = w_ref[w_ref["eom"] == d].sort_values("id")
        sel = np.flatnonzero(panel.valid[t])
        assert list(r["id"]) == list(panel.ids[sel])
        w_out[k, sel] = r["w"].to_numpy(); ws_out[k, sel] = r["w_start"].to_numpy()
        ww = max(ww, relerr(w_or[t]["w"], r["w"].to_numpy()), relerr(w_or[t]["w_start"], r["w_start"].to_numpy()))
    pf_arr = pf_ref.sort_values("eom_ret")[["inv", "shorting", "turnover", "r", "tc"]].to_numpy()
    st_arr = np.array([[st_or[t][k] for k in ("inv", "shorting", "turnover", "r", "tc")] for t in oos_idx])
    print("oracle vs PFML_best_hps.py pfml_w / pf_ts_fun: weights", ww, "stats", relerr(st_arr, pf_arr))
    pdict = panel_to_dict(panel)
    pdict["panel_feats_q"] = np.round(pdict.pop("panel_feats") * 1000.0).astype(np.uint16)
    np.savez_compressed(os.path.join(gold, "pfml_ref_small.npz"), r_tilde=r_t, denom=den, risk=risk, tc=tc,
                        signal_t=sig, n_d=n_d, hot=hot, p_vec=np.array(p_vec), l_vec=np.array(l_vec),
                        years=np.array(years), **coef_arrs, **pdict,
                        val_eom=v_ref["eom"].to_numpy().astype("datetime64[D]"), val_obj=v_ref["obj"].to_numpy(),
                        val_l=v_ref["l"].to_numpy(), val_p=v_ref["p"].to_numpy(), val_hp_end=v_ref["hp_end"].to_numpy(),
                        val_cum_obj=v_ref["cum_obj"].to_numpy(), val_rank=v_ref["rank"].to_numpy(),
                        oos_idx=np.array(oos_idx), aim_p=np.array(aim_p), aim_l=np.array(aim_l), w_aim=w_aim,
                        bt_w=w_out, bt_w_start=ws_out, bt_stats=pf_arr,
                        versions=np.array([np.__version__, pd.__version__, __import__("scipy").__version__]))

    # ---------------- function-level known answers -----------------
    rng = np.random.default_rng(5)
    n = 30
    pn = syn.make_panel(N=n, T=14, p=8, seed=7)
    sigma = GF.create_cov({"fct_load": pd.DataFrame(pn.fct_load[13], index=pn.ids),
                           "fct_cov": pd.DataFrame(pn.fct_cov[13]),
                           "ivol_vec": pd.Series(pn.ivol[13], index=pn.ids)}, ids=pn.ids[::2])
    lam_m = GF.create_lambda(dict(zip(pn.ids, pn.lam[13])), ids=pn.ids[::2])
    m_ref = GF.m_func(w=pn.wealth[13], mu=pn.mu, rf=pn.rf[13], sigma_gam=sigma.to_numpy() * 10.0, gam=10.0,
                      K_Lambda=lam_m, iterations=10)
    m_or = O.m_func(pn.wealth[13], pn.mu, pn.rf[13], sigma.to_numpy() * 10.0, 10.0, lam_m, 10)
    print("m_func oracle vs reference:", relerr(m_or, m_ref))
    X = rng.standard_normal((300, 6)) * 0.01
    wts = (0.5 ** (1 / 126)) ** np.arange(300, 0, -1)
    cov_ref = GF.weighted_cov_wt(pd.DataFrame(X), wts).to_numpy()
    cor_ref = GF.weighted_cor_wt(pd.DataFrame(X), wts).to_numpy()
    print("wcov/wcor oracle vs reference:", relerr(O.weighted_cov_wt(X, wts), cov_ref),
          relerr(O.weighted_cor_wt(X, wts), cor_ref))
    # ewma_vol: exec only the function definition out of the covariance script
    src = open(os.path.join(H.REF, "Estimate Covariance Matrix.py")).read()
    a = src.index("@njit\ndef ewma_vol")
    b = src.index("#wrapper that pandas.apply() can call")
    ns = {}
    exec("import numpy as np\nfrom numba import njit\n" + src[a:b], ns)
    xs = rng.standard_normal(400) * 0.02
    xs[rng.random(400) < 0.1] = np.nan
    lam_e = 0.5 ** (1 / 126)
    ev_ref = ns["ewma_vol"](xs, lam_e, 63)
    ev_or = O.ewma_vol(xs, lam_e, 63)
    print("ewma_vol oracle vs reference:", float(np.nanmax(np.abs(ev_or - ev_ref))),
          bool((np.isnan(ev_or) == np.isnan(ev_ref)).all()))
    feat_idx = {p: np.array([GF.pfml_feat_fun(16).index(f) for f in GF.pfml_feat_fun(p)]) for p in (4, 8, 16)}
    for p, fi in feat_idx.items():
        assert (fi == O.pfml_feat_idx(p, 16)).all()
    np.savez_compressed(os.path.join(gold, "funcs_ref.npz"),
                        m_sigma=sigma.to_numpy(), m_lam=np.diag(lam_m), m_w=pn.wealth[13], m_rf=pn.rf[13],
                        m_mu=pn.mu, m_ref=m_ref, wc_X=X, wc_w=wts, cov_ref=cov_ref, cor_ref=cor_ref,
                        ev_x=xs, ev_lam=lam_e, ev_ref=ev_ref,
                        feat_idx_4=feat_idx[4], feat_idx_8=feat_idx[8], feat_idx_16=feat_idx[16])
    # ---------------- script-level golden: "Estimate Covariance Matrix.py" -----------------
    daily = syn.make_daily()
    cov_set = dict(obs=400, hl_cor=150, hl_var=60, hl_stock_var=60, initial_var_obs=40)
    with tempfile.TemporaryDirectory() as tmp:
        barra = H.run_cov_script(daily, tmp + "/", cov_set)
    keys = sorted(barra.keys())
    N = len(daily["ids"])
    K = barra[keys[0]]["fct_cov"].shape[0]
    load = np.full((len(keys), N, K), np.nan); ivol = np.full((len(keys), N), np.nan)
    fcov = np.stack([barra[k]["fct_cov"].to_numpy() for k in keys])
    for c, k in enumerate(keys):
        ii = np.searchsorted(daily["ids"], np.asarray(barra[k]["fct_load"].index))
        load[c, ii] = barra[k]["fct_load"].to_numpy()
        ivol[c, ii] = barra[k]["ivol_vec"].to_numpy()
    np.savez_compressed(os.path.join(gold, "cov_ref_small.npz"), calc_dates=np.array([np.datetime64(k.date()) for k in keys]),
                        fct_load=load, fct_cov=fcov, ivol=ivol, factor_cols=np.array(list(barra[keys[0]]["fct_cov"].columns)),
                        cov_set=np.array([cov_set[k] for k in ("obs", "hl_cor", "hl_var", "hl_stock_var", "initial_var_obs")]),
                        daily_feats_q=np.round(daily["feats"] * 1000.0).astype(np.uint16),
                        **{"daily_" + k: v for k, v in daily.items() if k not in ("seed", "feats")})
    print("covariance script golden:", len(keys), "calc dates", keys[0].date(), "..", keys[-1].date())
    print("golden fixtures written to", gold)


if __name__ == "__main__":
    main()
