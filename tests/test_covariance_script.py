"""The "Estimate Covariance Matrix.py" drop-in against the Barra_Cov.pkl the reference script wrote
(tests/golden/cov_ref_small.npz, made by oracle/make_golden.py).  Tolerance 1e-8 relative.

CPU variant: the host (pandas) logic with the three numerical loops served by the oracle -- it checks joins,
calendars, the coverage rule and the median imputation without a GPU.  GPU variant: the real kernels."""
import os

import numpy as np
import pandas as pd
import pytest

import pfml_oracle as O
from conftest import GOLD, pkg, relerr


def _setup(tmp_path, syn):
    z = np.load(os.path.join(GOLD, "cov_ref_small.npz"))
    daily = {k[len("daily_"):]: z[k] for k in z.files if k.startswith("daily_") and k != "daily_feats_q"}
    daily["feats"] = z["daily_feats_q"].astype(np.float64) / 1000.0
    path = str(tmp_path) + "/"
    syn.write_covariance_artefacts(daily, path)
    settings, pf_set = pkg("general_functions").get_settings()
    settings["cov_set"].update(dict(zip(("obs", "hl_cor", "hl_var", "hl_stock_var", "initial_var_obs"), (int(v) for v in z["cov_set"]))))
    return z, daily, path, (settings, pf_set)


def _check(barra, z, daily):
    keys = sorted(barra.keys())
    assert [np.datetime64(k.date()) for k in keys] == list(z["calc_dates"])
    for c, k in enumerate(keys):
        b = barra[k]
        assert list(b["fct_cov"].columns) == list(z["factor_cols"]) == list(b["fct_load"].columns)
        present = ~np.isnan(z["ivol"][c])
        assert list(b["fct_load"].index) == list(daily["ids"][present]) == list(b["ivol_vec"].index)
        assert relerr(b["fct_load"].to_numpy(), z["fct_load"][c][present]) < 1e-12
        assert relerr(b["fct_cov"].to_numpy(), z["fct_cov"][c]) < 1e-8, c
        assert relerr(b["ivol_vec"].to_numpy(), z["ivol"][c][present]) < 1e-8, c


def test_cov_script_host_logic_with_oracle_numerics(tmp_path, syn, monkeypatch):
    ec = pkg("estimate_covariance")
    z, daily, path, st = _setup(tmp_path, syn)

    def ols(X, y, ptr, device=None):
        coef, res = [], []
        for d in range(len(ptr) - 1):
            c, r = O.daily_ols(X[ptr[d]:ptr[d + 1]], y[ptr[d]:ptr[d + 1]])
            coef.append(c); res.append(r)
        return np.stack(coef), np.concatenate(res), np.zeros(len(ptr) - 1, dtype=np.int32)

    def fcov(F, ends, obs, hl_cor, hl_var, scale, device=None):
        wc, wv = ec.ewma_weights(obs, hl_cor), ec.ewma_weights(obs, hl_var)
        return np.stack([O.factor_cov(F[max(0, e + 1 - obs):e + 1], wc, wv) * scale for e in ends])

    monkeypatch.setattr(ec, "daily_ols", ols)
    monkeypatch.setattr(ec, "factor_cov", fcov)
    monkeypatch.setattr(ec, "ewma_vol_batch", lambda s, lam, start, device=None: [O.ewma_vol(x, lam, start) for x in s])
    _check(ec.run(path, st), z, daily)


@pytest.mark.gpu
def test_cov_script_gpu(tmp_path, syn):
    z, daily, path, st = _setup(tmp_path, syn)
    barra = pkg("estimate_covariance").run(path, st)
    _check(barra, z, daily)
    # the written pickle feeds PFML_Input_Data.py: same schema as the reference's
    import pickle
    with open(os.path.join(path, "Data", "Barra_Cov.pkl"), "rb") as f:
        b = pickle.load(f)
    assert set(next(iter(b.values())).keys()) == {"fct_load", "fct_cov", "ivol_vec"}
