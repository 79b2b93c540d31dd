"""CPU: the dense tensor store (tensor_store.py) -- memory-mapped arrays of packed triangles written block by
block and read back as the reference's dict of DataFrames.  No GPU: the rebuild of the squares is host code of the C-ABI
library (pfml_unpack_reals)."""
import numpy as np
import pandas as pd

from conftest import pkg


def test_store_round_trip(tmp_path, syn):
    ts = pkg("tensor_store")
    GF = pkg("general_functions")
    rng = np.random.default_rng(3)
    T, p, NP = 5, 8, 64
    P = p + 1
    dates = syn.month_end_range("2015-01", T)
    sym = lambda a: np.triu(a) + np.triu(a, 1).T
    risk = np.stack([sym(rng.standard_normal((P, P))) for _ in range(T)])
    tc = np.stack([sym(rng.standard_normal((P, P))) for _ in range(T)])
    r_t = rng.standard_normal((T, P))
    n_sel = np.array([40, 37, 64, 1, 50])
    sig = np.zeros((T, NP, P))
    ids = []
    for t in range(T):
        sig[t, :n_sel[t]] = rng.standard_normal((n_sel[t], P))
        ids.append(10001 + np.sort(rng.choice(500, n_sel[t], replace=False)))
    iu = np.triu_indices(P)
    W = rng.standard_normal((115, p // 2))
    w = ts.StoreWriter(str(tmp_path), 0, dates, P, NP, W)
    w.put_block(0, r_t[:2], risk[:2][:, iu[0], iu[1]], tc[:2][:, iu[0], iu[1]], sig[:2], ids[:2])
    w.put_block(2, r_t[2:], risk[2:][:, iu[0], iu[1]], tc[2:][:, iu[0], iu[1]], sig[2:], ids[2:])
    w.close()
    st = ts.Store(str(tmp_path), 0)
    assert st.T == T and st.P == P and (st.dates == dates).all()
    for t in range(T):
        r, den, rk, c = st.reals(t)
        assert np.array_equal(r, r_t[t]) and np.array_equal(rk, risk[t]) and np.array_equal(c, tc[t])
        assert np.array_equal(den, risk[t] + tc[t])
    assert np.array_equal(st.denom_stack(), risk + tc)
    leg = st.to_legacy()
    feat_all = GF.pfml_feat_fun(p)
    d2 = pd.Timestamp(dates[2])
    assert list(leg["reals"][d2]["denom"].columns) == feat_all and list(leg["reals"][d2]["r_tilde"].index) == feat_all
    assert leg["signal_t"][d2].shape == (64, P) and list(leg["signal_t"][d2].index) == list(ids[2])
    assert np.array_equal(leg["signal_t"][pd.Timestamp(dates[3])].to_numpy(), sig[3, :1])
    assert np.array_equal(leg["rff_w"], W)
