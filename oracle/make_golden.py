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
            "fct_load", "fct_cov", "ivol", "barra_has", "W"]
    d = {"panel_" + k: getattr(panel, k) for k in keys}
    d["panel_scalars"] = np.array([panel.gamma_rel, panel.mu, panel.lb, panel.iterations, panel.meta["g"]])
    return d


def relerr(a, b):
    return float(np.max(np.abs(a - b)) / max(np.max(np.abs(b)), 1e-300))


def main():
    syn = importlib.import_module(PKG + ".synthetic")
    gold = os.path.join(ROOT, "tests", "golden")
    os.makedirs(gold, exist_ok=True)

    # ---------------- script-level golden: PFML_Input_Data.py + PFML_Search_Coef.py -----------------
    p_vec = [4, 8, 16]
    l_vec = [0.0, float(np.exp(-10)), 1e-2, 1.0]
    panel = syn.make_panel(N=20, T=48, p=16, seed=11, ragged=0.08, barra_missing=0.1, start="2010-01")
    with tempfile.TemporaryDirectory() as tmp:
        out, glob = H.run_pfml_input(panel, tmp + "/", p_vec, l_vec, start_year=2013, end_yr=2014)
        coef = H.run_search_coef(panel, tmp + "/", p_vec, l_vec, start_year=2013, end_yr=2014)
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
    o_rt, o_den = [], []
    for k, t in enumerate(hot):
        o = O.date_step(panel, t, rff_x, vs)
        o_rt.append(o["r_tilde"]); o_den.append(o["denom"])
        for name, ref in (("r_tilde", r_t[k]), ("denom", den[k]), ("risk", risk[k]), ("tc", tc[k]),
                          ("signal_t", sig[k, :n_d[k]])):
            worst[name] = max(worst.get(name, 0.0), relerr(o[name], ref))
    print("oracle vs exec'd PFML_Input_Data.py, max rel err:", worst)
    years = sorted(coef.keys())
    o_coef = O.search_coef(np.stack(o_rt), np.stack(o_den), panel.month_ends[hot], years, p_vec, l_vec)
    cw = 0.0
    coef_arrs = {}
    for y in years:
        for p in p_vec:
            ref_b = np.stack([coef[y][p][j].to_numpy() for j in range(len(l_vec))])
            coef_arrs[f"coef_{y}_{p}"] = ref_b
            for j in range(1, len(l_vec)):          # lambda=0 is condition-limited (quirk Q14)
                cw = max(cw, relerr(o_coef[y][p][j], ref_b[j]))
            assert list(coef[y][p][0].index) == GF.pfml_feat_fun(p)
    print("oracle vs exec'd PFML_Search_Coef.py (lambda>0), max rel err:", cw)
    np.savez_compressed(os.path.join(gold, "pfml_ref_small.npz"), r_tilde=r_t, denom=den, risk=risk, tc=tc,
                        signal_t=sig, n_d=n_d, hot=hot, p_vec=np.array(p_vec), l_vec=np.array(l_vec),
                        years=np.array(years), **coef_arrs, **panel_to_dict(panel),
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
