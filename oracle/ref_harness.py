"""Run the UNMODIFIED reference scripts from /root/reference on a synthetic panel (build container only).

TEST INFRASTRUCTURE.  Used by oracle/make_golden.py to pin the numpy oracle against the reference
itself; never imported by the product package, the GPU tests, smoke() or bench.py (the reference tree
does not exist on the GPU box).

Procedure: write the artefacts the scripts read under ``path + "Data/"``, stub
the one off-path missing import (statsmodels ECDF, General_functions.py:14), wrap
``General_functions.get_settings`` to set dates/grids, patch the pandas alias ``freq='M'`` ->
``'ME'`` (PFML_Input_Data.py:231, rejected by pandas 3), then ``exec`` each script with the global
``path`` exactly as Main.py:21-22 does.
"""
from __future__ import annotations

import os
import pickle
import sqlite3
import sys
import types

import numpy as np
import pandas as pd

REF = "/root/reference"


def _import_general_functions():
    if "statsmodels" not in sys.modules:
        sm = types.ModuleType("statsmodels")
        sd = types.ModuleType("statsmodels.distributions")
        se = types.ModuleType("statsmodels.distributions.empirical_distribution")
        se.ECDF = object
        sys.modules.update({"statsmodels": sm, "statsmodels.distributions": sd,
                            "statsmodels.distributions.empirical_distribution": se})
    if REF not in sys.path:
        sys.path.insert(0, REF)
    import General_functions as GF
    return GF


def write_reference_inputs(panel, path):
    """Write Factors_processed, FF_RF_monthly.csv, wealth_processed.csv, Barra_Cov.pkl, rff_w.csv (the writer lives
    in the package so that the GPU tests can use it without the reference tree)."""
    import importlib
    GF = _import_general_functions()
    syn = importlib.import_module("jkmp22-machine-learning-and-the-implementable-efficient-frontier-replication_b200.synthetic")
    syn.write_reference_artefacts(panel, path, GF.get_features(True))


def _settings_wrapper(GF, panel, p_vec, l_vec, start_year, end_yr, n_g=1):
    orig = GF._orig_get_settings if hasattr(GF, "_orig_get_settings") else GF.get_settings
    GF._orig_get_settings = orig

    def get_settings():
        settings, pf_set = orig()
        settings["split"]["test_end"] = pd.Timestamp(panel.month_ends[-1])
        settings["pf"]["dates"]["start_year"] = start_year
        settings["pf"]["dates"]["end_yr"] = end_yr
        settings["pf_ml"]["g_vec"] = [panel.meta.get("g", 0.05)] * n_g
        settings["pf_ml"]["p_vec"] = list(p_vec)
        settings["pf_ml"]["l_vec"] = list(l_vec)
        pf_set["gamma_rel"] = panel.gamma_rel
        pf_set["mu"] = panel.mu
        pf_set["lb_hor"] = panel.lb
        return settings, pf_set

    GF.get_settings = get_settings


def run_script(name, path, patches=()):
    with open(os.path.join(REF, name)) as f:
        src = f.read()
    for a, b in patches:
        assert a in src
        src = src.replace(a, b)
    glob = {"path": path if path.endswith("/") else path + "/", "__name__": "__main__"}
    exec(compile(src, name, "exec"), glob)
    return glob


def run_pfml_input(panel, path, p_vec, l_vec=(0.0,), start_year=2013, end_yr=2013):
    GF = _import_general_functions()
    write_reference_inputs(panel, path)
    _settings_wrapper(GF, panel, p_vec, l_vec, start_year, end_yr)
    glob = run_script("PFML_Input_Data.py", path, patches=[("freq='M')", "freq='ME')")])
    with open(os.path.join(path, "Data", "pfml_input_0.pkl"), "rb") as f:
        out = pickle.load(f)[0]
    return out, glob


def run_search_coef(panel, path, p_vec, l_vec, start_year, end_yr):
    GF = _import_general_functions()
    _settings_wrapper(GF, panel, p_vec, l_vec, start_year, end_yr)
    run_script("PFML_Search_Coef.py", path)
    with open(os.path.join(path, "Data", "coef_dict_0.pkl"), "rb") as f:
        return pickle.load(f)[0]


def run_cov_script(daily, path, cov_set):
    """exec "Estimate Covariance Matrix.py" on a synthetic daily panel; returns the Barra_Cov dict it pickles."""
    import importlib
    GF = _import_general_functions()
    syn = importlib.import_module("jkmp22-machine-learning-and-the-implementable-efficient-frontier-replication_b200.synthetic")
    syn.write_covariance_artefacts(daily, path, GF.get_features(True))
    orig = GF._orig_get_settings if hasattr(GF, "_orig_get_settings") else GF.get_settings
    GF._orig_get_settings = orig

    def get_settings():
        settings, pf_set = orig()
        settings["cov_set"].update(cov_set)
        return settings, pf_set

    GF.get_settings = get_settings
    run_script("Estimate Covariance Matrix.py", path)
    with open(os.path.join(path, "Data", "Barra_Cov.pkl"), "rb") as f:
        return pickle.load(f)


def run_hp_reals(panel, path, p_vec, l_vec, start_year, end_yr):
    """exec PFML_hp_reals.py (needs pfml_input_0.pkl and coef_dict_0.pkl under path); returns validation.csv."""
    GF = _import_general_functions()
    _settings_wrapper(GF, panel, p_vec, l_vec, start_year, end_yr)
    run_script("PFML_hp_reals.py", path)
    return pd.read_csv(os.path.join(path, "Data", "validation.csv"), parse_dates=["eom", "eom_ret"])


def run_aim_fun(panel, path, p_vec, l_vec, start_year, end_yr, split_years):
    """exec PFML_aim_fun.py; the script hard-codes test_end = 2023-12-31 at :94, which is patched to the panel's."""
    GF = _import_general_functions()
    _settings_wrapper(GF, panel, p_vec, l_vec, start_year, end_yr)
    inner = GF.get_settings

    def get_settings():
        settings, pf_set = inner()
        settings["pf"]["dates"]["split_years"] = split_years
        return settings, pf_set

    GF.get_settings = get_settings
    te = pd.Timestamp(panel.month_ends[-1]).strftime("%Y-%m-%d")
    run_script("PFML_aim_fun.py", path, patches=[("settings['split']['test_end'] = pd.to_datetime('2023-12-31')",
                                                  f"settings['split']['test_end'] = pd.to_datetime('{te}')")])
    with open(os.path.join(path, "Data", "aims.pkl"), "rb") as f:
        return pickle.load(f)[0]


def backtest_functions():
    """The function definitions pfml_w / initial_weights_new / compute_stats / pf_ts_fun of PFML_best_hps.py:137-259,
    exec'd on their own (the script itself imports plotnine, which is not installed)."""
    GF = _import_general_functions()
    src = open(os.path.join(REF, "PFML_best_hps.py")).read()
    a = src.index("def initial_weights_new(")
    b = src.index("#%% Get best Hyperparameter combination")
    ns = {"np": np, "pd": pd, "create_cov": GF.create_cov, "create_lambda": GF.create_lambda, "m_func": GF.m_func}
    seg = src[a:b]
    # pandas >= 3 drops the grouping column inside groupby().apply(); bring 'eom' back from the group index
    # (second pandas-compatibility patch besides freq='M' -> 'ME'; the arithmetic is untouched)
    old = ".reset_index(drop=True)[[\"id\", \"eom\", \"w_start\"]]"
    assert seg.count(old) == 2
    seg = seg.replace(old, ".reset_index(level=0).reset_index(drop=True)[[\"id\", \"eom\", \"w_start\"]]")
    exec(compile(seg, "PFML_best_hps.py[functions]", "exec"), ns)
    return ns
