import importlib
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)

PKG = "jkmp22-machine-learning-and-the-implementable-efficient-frontier-replication_b200"
GOLD = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def pkg(sub=None):
    return importlib.import_module(PKG + ("." + sub if sub else ""))


@pytest.fixture(scope="session")
def syn():
    return pkg("synthetic")


def relerr(a, b):
    import numpy as np
    return float(np.max(np.abs(np.asarray(a) - np.asarray(b))) / max(float(np.max(np.abs(b))), 1e-300))


def panel_from_golden(z, syn_mod):
    sc = z["panel_scalars"]
    feats = z["panel_feats_q"].astype("float64") / 1000.0
    return syn_mod.Panel(me=z["panel_me"], tr_ld1=z["panel_tr_ld1"], mu_ld1=z["panel_mu_ld1"],
                         ids=z["panel_ids"], month_ends=z["panel_month_ends"], feats=feats,
                         valid=z["panel_valid"], ret_ld1=z["panel_ret_ld1"], tr_ld0=z["panel_tr_ld0"],
                         mu_ld0=z["panel_mu_ld0"], lam=z["panel_lam"], wealth=z["panel_wealth"], rf=z["panel_rf"],
                         fct_load=z["panel_fct_load"], fct_cov=z["panel_fct_cov"], ivol=z["panel_ivol"],
                         barra_has=z["panel_barra_has"], W=z["panel_W"], gamma_rel=float(sc[0]), mu=float(sc[1]),
                         lb=int(sc[2]), iterations=int(sc[3]), meta=dict(g=float(sc[4])))
