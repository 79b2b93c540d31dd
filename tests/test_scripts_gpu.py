"""GPU: the script-level drop-ins (path contract of Main.py:21-22) against the outputs of the reference scripts.

The artefacts under Data/ are written from the golden panel with the same writer the reference run used; the
drop-ins read them, compute on the GPU and write pfml_input_0.pkl / coef_dict_0.pkl, which must match what the
exec'd PFML_Input_Data.py / PFML_Search_Coef.py wrote (tests/golden/pfml_ref_small.npz) to 1e-8."""
import os
import pickle

import numpy as np
import pandas as pd
import pytest

from conftest import GOLD, panel_from_golden, pkg, relerr

pytestmark = pytest.mark.gpu


def test_pfml_input_data_and_search_coef_scripts(tmp_path, syn):
    z = np.load(os.path.join(GOLD, "pfml_ref_small.npz"))
    panel = panel_from_golden(z, syn)
    path = str(tmp_path) + "/"
    syn.write_reference_artefacts(panel, path)
    GF = pkg("general_functions")
    settings, pf_set = GF.get_settings()
    settings["split"]["test_end"] = pd.Timestamp(panel.month_ends[-1])
    settings["pf"]["dates"].update(start_year=2013, end_yr=2014)
    settings["pf_ml"].update(g_vec=[panel.meta["g"]], p_vec=[int(p) for p in z["p_vec"]], l_vec=[float(l) for l in z["l_vec"]])
    out = pkg("pfml_input_data").run(path, (settings, pf_set))[0]
    dates = sorted(out["reals"].keys())
    assert len(dates) == len(z["hot"])
    assert [pd.Timestamp(panel.month_ends[t]) for t in z["hot"]] == dates
    feat_all = GF.pfml_feat_fun(panel.p)
    for k, d in enumerate(dates):
        r = out["reals"][d]
        assert list(r["r_tilde"].index) == feat_all and list(r["denom"].columns) == feat_all
        for name in ("r_tilde", "denom", "risk", "tc"):
            assert relerr(r[name].to_numpy(), z[name][k]) < 1e-8, (name, k)
        n = int(z["n_d"][k])
        s = out["signal_t"][d]
        assert s.shape == (n, panel.p + 1) and list(s.index) == list(panel.ids[panel.valid[z["hot"][k]]])
        assert relerr(s.to_numpy(), z["signal_t"][k, :n]) < 1e-8
    with open(os.path.join(path, "Data", "pfml_input_0.pkl"), "rb") as f:
        assert set(pickle.load(f)[0].keys()) == {"reals", "signal_t", "rff_w"}
    coef = pkg("pfml_search_coef").run(path, (settings, pf_set))[0]
    for y in z["years"]:
        for p in z["p_vec"]:
            ref = z[f"coef_{y}_{p}"]
            for j, l in enumerate(z["l_vec"]):
                s = coef[int(y)][int(p)][j]
                assert list(s.index) == GF.pfml_feat_fun(int(p))
                if l > 0:
                    assert relerr(s.to_numpy(), ref[j]) < 1e-8, (y, p, l)


def test_hp_reals_and_aim_fun_scripts(tmp_path, syn):
    """validation.csv and aims.pkl written by the drop-ins equal what the exec'd reference scripts wrote."""
    z = np.load(os.path.join(GOLD, "pfml_ref_small.npz"))
    panel = panel_from_golden(z, syn)
    path = str(tmp_path) + "/"
    syn.write_reference_artefacts(panel, path)
    GF = pkg("general_functions")
    settings, pf_set = GF.get_settings()
    settings["split"]["test_end"] = pd.Timestamp(panel.month_ends[-1])
    settings["pf"]["dates"].update(start_year=2013, end_yr=2014, split_years=1)
    settings["pf_ml"].update(g_vec=[panel.meta["g"]], p_vec=[int(p) for p in z["p_vec"]], l_vec=[float(l) for l in z["l_vec"]])
    st = (settings, pf_set)
    pkg("pfml_input_data").run(path, st)
    # score with the reference's own coefficients so that lambda = 0 conditioning does not enter the comparison
    coef = {0: {int(y): {int(p): {j: pd.Series(z[f"coef_{y}_{p}"][j], index=GF.pfml_feat_fun(int(p))) for j in range(len(z["l_vec"]))}
                         for p in z["p_vec"]} for y in z["years"]}}
    with open(os.path.join(path, "Data", "coef_dict_0.pkl"), "wb") as f:
        pickle.dump(coef, f)
    v = pkg("pfml_hp_reals").run(path, st)
    v2 = pd.read_csv(os.path.join(path, "Data", "validation.csv"), parse_dates=["eom", "eom_ret"])
    assert list(v2.columns) == ["eom", "eom_ret", "obj", "l", "p", "hp_end", "cum_obj", "rank", "g"]
    assert relerr(v["obj"].to_numpy(), z["val_obj"]) < 1e-8 and (v["rank"].to_numpy() == z["val_rank"]).all()
    aims = pkg("pfml_aim_fun").run(path, st)[0]
    dates = sorted(aims.keys())
    assert [np.datetime64(d.date()) for d in dates] == list(panel.month_ends[z["oos_idx"]])
    for b, d in enumerate(dates):
        sel = np.flatnonzero(panel.valid[z["oos_idx"][b]])
        a = aims[d]["aim_pf"]
        assert list(a["id"]) == list(panel.ids[sel]) and list(a.columns) == ["id", "eom", "w_aim"]
        assert relerr(a["w_aim"].to_numpy(), z["w_aim"][b, sel]) < 1e-8
        assert len(aims[d]["coef"]) == z["aim_p"][b] + 1


def test_function_mirrors(syn):
    GF = pkg("general_functions")
    f = np.load(os.path.join(GOLD, "funcs_ref.npz"))
    m = GF.m_func(float(f["m_w"]), float(f["m_mu"]), float(f["m_rf"]), f["m_sigma"] * 10.0, 10.0, np.diag(f["m_lam"]), 10)
    assert relerr(m, f["m_ref"]) < 1e-8
    pn = syn.make_panel(N=40, T=14, p=8, seed=5)
    x = {"fct_load": pd.DataFrame(pn.fct_load[13], index=pn.ids), "fct_cov": pd.DataFrame(pn.fct_cov[13]),
         "ivol_vec": pd.Series(pn.ivol[13], index=pn.ids)}
    ids = pn.ids[::3]
    sig = GF.create_cov(x, ids)
    L = x["fct_load"].loc[ids].to_numpy()
    assert relerr(sig.to_numpy(), L @ pn.fct_cov[13] @ L.T + np.diag(pn.ivol[13][::3])) < 1e-13
    assert list(sig.index) == list(ids)
    r = pkg("pfml_input_data").rff(pn.feats[0], W=pn.W)
    assert relerr(r["X_cos"], np.cos(pn.feats[0] @ pn.W)) < 1e-14 and relerr(r["X_sin"], np.sin(pn.feats[0] @ pn.W)) < 1e-14
    assert GF.create_lambda({1: 0.1, 2: 0.2, 3: 0.3}, [3, 1]).tolist() == [[0.3, 0.0], [0.0, 0.1]]


def test_tensor_store_equals_pickle_outputs(tmp_path, syn):
    """pfml_input_data.run_store (packed host entry point -> memory-mapped store) holds exactly what run() pickles, and
    PFML_Search_Coef from the store equals PFML_Search_Coef from the pickle."""
    z = np.load(os.path.join(GOLD, "pfml_ref_small.npz"))
    panel = panel_from_golden(z, syn)
    path = str(tmp_path) + "/"
    syn.write_reference_artefacts(panel, path)
    GF = pkg("general_functions")
    settings, pf_set = GF.get_settings()
    settings["split"]["test_end"] = pd.Timestamp(panel.month_ends[-1])
    settings["pf"]["dates"].update(start_year=2013, end_yr=2014)
    settings["pf_ml"].update(g_vec=[panel.meta["g"]], p_vec=[int(p) for p in z["p_vec"]], l_vec=[float(l) for l in z["l_vec"]])
    st = (settings, pf_set)
    ref = pkg("pfml_input_data").run(path, st)[0]
    dirs = pkg("pfml_input_data").run_store(path, st, block=7)
    assert os.path.exists(os.path.join(dirs[0], "meta.json"))
    store = pkg("tensor_store").Store(path, 0)
    leg = store.to_legacy()
    assert sorted(leg["reals"].keys()) == sorted(ref["reals"].keys())
    for d in ref["reals"]:
        for name in ("r_tilde", "denom", "risk", "tc"):
            a, b = leg["reals"][d][name], ref["reals"][d][name]
            assert np.array_equal(a.to_numpy(), b.to_numpy()), (d, name)
            assert list(a.index) == list(b.index)
        assert leg["signal_t"][d].equals(ref["signal_t"][d])
    assert np.array_equal(leg["rff_w"], ref["rff_w"])
    size_store = sum(os.path.getsize(os.path.join(dirs[0], f)) for f in os.listdir(dirs[0]))
    assert size_store < os.path.getsize(os.path.join(path, "Data", "pfml_input_0.pkl"))
    c_pkl = pkg("pfml_search_coef").run(path, st)[0]
    c_sto = pkg("pfml_search_coef").run(path, st, source="store")[0]
    for y in c_pkl:
        for p in c_pkl[y]:
            for j in c_pkl[y][p]:
                assert np.array_equal(c_pkl[y][p][j].to_numpy(), c_sto[y][p][j].to_numpy())


def test_percentile_rank_bit_exact_vs_pandas():
    """pfml_percentile_rank vs the reference's own three lines (Prepare_Data.py:338-345, :368): ties, zeros, NaNs, a month
    with a single non-missing value, an all-missing column."""
    rng = np.random.default_rng(9)
    T, N, F = 5, 300, 7
    x = np.round(rng.standard_normal((T, N, F)), 1)              # one decimal: many ties
    x[rng.random((T, N, F)) < 0.1] = np.nan
    x[rng.random((T, N, F)) < 0.05] = 0.0
    x[1, :, 2] = np.nan; x[1, 17, 2] = 3.0                       # single value
    x[2, :, 4] = np.nan                                          # nothing to rank
    x[3, 5, 0] = np.inf; x[3, 6, 0] = -np.inf
    got = pkg("prepare_data").percentile_rank(x, impute=True)
    got_nan = pkg("prepare_data").percentile_rank(x, impute=False)
    feats = [f"f{k}" for k in range(F)]
    chars = pd.DataFrame(x.reshape(T * N, F), columns=feats)
    chars["eom"] = np.repeat(np.arange(T), N)
    ranked = chars.groupby("eom")[feats].rank(pct=True)
    mask = chars[feats] == 0.0
    ranked[mask] = 0
    want_nan = ranked.to_numpy().reshape(T, N, F)
    want = ranked.apply(lambda c: c.fillna(0.5)).to_numpy().reshape(T, N, F)
    assert np.array_equal(got, want)
    assert np.array_equal(got_nan, want_nan, equal_nan=True)
