"""Second reference-run fixture: tests/golden/pfml_ref_mid.npz (build container only; needs /root/reference).

TEST INFRASTRUCTURE.  Usage:  python oracle/make_golden_mid.py

The first fixture (oracle/make_golden.py, N = 20, p = 16) never leaves one 64-block of the block inverse.  This one
exec's the unmodified PFML_Input_Data.py on N = 150 stocks (padded 192 = 3 blocks: recursion depth 2, uneven split),
p = 64 (P = 65 > 64: two column blocks), a ragged universe that changes by ~10 % of the ids per month and 20 % of the
Barra ids missing (eom-median fill), and freezes r_tilde / denom / risk / tc / signal_t of its hot dates together with
the panel that produced them.  It also prints the oracle-vs-reference deviation on these outputs (the pinning check
tests/test_oracle_golden.py repeats) and, for the last date, omega / omega_chg / m read from the exec'd script's
globals (the stage-level vectors behind tests/test_stage_entry_points_gpu.py).
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
from make_golden import panel_to_dict, relerr   # noqa: E402


def main():
    syn = importlib.import_module(PKG + ".synthetic")
    gold = os.path.join(ROOT, "tests", "golden")
    # 13 look-back months + 3 hot dates; dates_m2 ends one month before test_end, hence T = 17
    panel = syn.make_panel(N=150, T=17, p=64, seed=31, ragged=0.10, barra_missing=0.2, start="2012-01", quantize=True)
    with tempfile.TemporaryDirectory() as tmp:
        out, glob = H.run_pfml_input(panel, tmp + "/", [64], [0.0], start_year=2013, end_yr=2013)
    dates = sorted(out["reals"].keys())
    hot = [int(t) for t in panel.hot_months()]
    assert len(dates) == len(hot), (len(dates), len(hot))
    assert all(pd.Timestamp(panel.month_ends[t]) == d for t, d in zip(hot, dates))
    P = panel.p + 1
    r_t = np.stack([out["reals"][d]["r_tilde"].to_numpy() for d in dates])
    den = np.stack([out["reals"][d]["denom"].to_numpy() for d in dates])
    risk = np.stack([out["reals"][d]["risk"].to_numpy() for d in dates])
    tc = np.stack([out["reals"][d]["tc"].to_numpy() for d in dates])
    sig = np.full((len(dates), panel.N, P), np.nan)
    n_d = np.zeros(len(dates), dtype=np.int64)
    for k, d in enumerate(dates):
        s = out["signal_t"][d].to_numpy()
        n_d[k] = s.shape[0]
        sig[k, :s.shape[0]] = s
    # stage-level vectors of the LAST date from the exec'd script's namespace (feat_cons order -> feat_all order)
    perm = np.concatenate([[panel.p], np.arange(panel.p)])
    omega = np.asarray(glob["omega"])[:, perm]
    omega_chg = np.asarray(glob["omega_chg"])[:, perm]
    m_last = np.asarray(glob["m"])

    T, N, F = panel.feats.shape
    rff_x = O.rff(panel.feats.reshape(T * N, F), panel.W).reshape(T, N, -1)
    vs = O.vol_scale_panel(panel)
    worst = {}
    for k, t in enumerate(hot):
        o = O.date_step(panel, t, rff_x, vs, keep=(k == len(hot) - 1))
        for name, ref in (("r_tilde", r_t[k]), ("denom", den[k]), ("risk", risk[k]), ("tc", tc[k]), ("signal_t", sig[k, :n_d[k]])):
            worst[name] = max(worst.get(name, 0.0), relerr(o[name], ref))
        if k == len(hot) - 1:
            worst["omega"] = relerr(o["omega"], omega)
            worst["omega_chg"] = relerr(o["omega_chg"], omega_chg)
            worst["m"] = relerr(o["m"], m_last)
    print("oracle vs exec'd PFML_Input_Data.py (N=150, p=64), max rel err:", worst)
    assert max(worst.values()) < 1e-9, worst
    pdict = panel_to_dict(panel)
    pdict["panel_feats_q"] = np.round(pdict.pop("panel_feats") * 1000.0).astype(np.uint16)      # 3-decimal ranks (quantize=True)
    np.savez_compressed(os.path.join(gold, "pfml_ref_mid.npz"), r_tilde=r_t, denom=den, risk=risk, tc=tc, signal_t=sig, n_d=n_d,
                        hot=np.asarray(hot), omega_last=omega, omega_chg_last=omega_chg, m_last=m_last,
                        versions=np.array([np.__version__, pd.__version__]), **pdict)
    print("wrote tests/golden/pfml_ref_mid.npz", os.path.getsize(os.path.join(gold, "pfml_ref_mid.npz")) // 1024, "KiB")


if __name__ == "__main__":
    main()
