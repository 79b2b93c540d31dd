"""Dense on-disk tensor store for the outputs of PFML_Input_Data.py, beside the legacy pickle.

The reference pickles, per g, a dict of ~360 dates x 4 pandas objects of 1 025 x 1 025 doubles plus the signal frames
(``pfml_input_{g}.pkl``, PFML_Input_Data.py:484-495): about 9 GB per g, written and re-read through pandas.  The store
keeps the same information as a directory of memory-mappable ``.npy`` arrays, one directory per g:

    Data/pfml_store_{g}/
        meta.json          schema version, P, NP, dates (ISO), feature labels are pfml_feat_fun(p) (not stored)
        r_tilde.npy        [T, P]
        risk_ut.npy        [T, P (P + 1) / 2]   upper triangle of risk, row-major (row r holds columns r .. P-1)
        tc_ut.npy          [T, P (P + 1) / 2]   upper triangle of tc;  denom = risk + tc exactly, all three symmetric
        signal_t.npy       [T, NP, P]           rows >= n_sel[t] are zero
        ids.npy            [T, NP] int64        stock ids of each date's universe (ascending), -1 beyond n_sel[t]
        n_sel.npy          [T] int32
        rff_w.npy          [F, p / 2]

This is the layout ``pfml_date_steps_host_packed`` fills directly (a third of the bytes of the three squares).  A reader
memory-maps the arrays; full squares are rebuilt per date on demand by the C-ABI host helper ``pfml_unpack_reals`` (exact
copies and one IEEE addition), and ``to_legacy()`` produces the reference's dict-of-DataFrames for drop-in consumers.
"""
from __future__ import annotations

import ctypes as C
import json
import os

import numpy as np

from . import _lib
from ._lib import check

SCHEMA = 1


def store_dir(path, g_index):
    return os.path.join(path, "Data", f"pfml_store_{g_index}")


class StoreWriter:
    """Pre-sized, memory-mapped arrays that a producer fills date by date (or block by block)."""

    def __init__(self, path, g_index, dates, P, NP, rff_w):
        self.dir = store_dir(path, g_index)
        os.makedirs(self.dir, exist_ok=True)
        T, PT = len(dates), P * (P + 1) // 2
        mk = lambda name, shape, dt=np.float64: np.lib.format.open_memmap(os.path.join(self.dir, name), mode="w+", dtype=dt, shape=shape)
        self.r_tilde, self.risk_ut, self.tc_ut = mk("r_tilde.npy", (T, P)), mk("risk_ut.npy", (T, PT)), mk("tc_ut.npy", (T, PT))
        self.signal_t = mk("signal_t.npy", (T, NP, P))
        self.ids = mk("ids.npy", (T, NP), np.int64)
        self.ids[:] = -1
        self.n_sel = mk("n_sel.npy", (T,), np.int32)
        np.save(os.path.join(self.dir, "rff_w.npy"), np.asarray(rff_w, dtype=np.float64))
        self.meta = {"schema": SCHEMA, "P": int(P), "NP": int(NP), "T": int(T),
                     "dates": [str(np.datetime64(d, "D")) for d in dates],
                     "layout": "risk_ut / tc_ut: row-major upper triangles; denom = risk + tc; columns in pfml_feat_fun(p) order"}

    def put_block(self, t0, r_tilde, risk_ut, tc_ut, signal_t, ids_list):
        """Dates t0 .. t0 + b - 1 from the packed host outputs of one call (arrays [b, ...]; ids_list: b id arrays)."""
        b = len(r_tilde)
        self.r_tilde[t0:t0 + b] = r_tilde
        self.risk_ut[t0:t0 + b] = risk_ut
        self.tc_ut[t0:t0 + b] = tc_ut
        self.signal_t[t0:t0 + b] = signal_t
        for k, ids in enumerate(ids_list):
            self.n_sel[t0 + k] = len(ids)
            self.ids[t0 + k, :len(ids)] = ids

    def close(self):
        for a in (self.r_tilde, self.risk_ut, self.tc_ut, self.signal_t, self.ids, self.n_sel):
            a.flush()
        with open(os.path.join(self.dir, "meta.json"), "w") as f:
            json.dump(self.meta, f)


class Store:
    """Memory-mapped reader."""

    def __init__(self, path, g_index):
        self.dir = store_dir(path, g_index)
        with open(os.path.join(self.dir, "meta.json")) as f:
            self.meta = json.load(f)
        if self.meta.get("schema") != SCHEMA:
            raise ValueError(f"{self.dir}: unknown store schema {self.meta.get('schema')}")
        ld = lambda name: np.load(os.path.join(self.dir, name), mmap_mode="r")
        self.r_tilde, self.risk_ut, self.tc_ut, self.signal_t = ld("r_tilde.npy"), ld("risk_ut.npy"), ld("tc_ut.npy"), ld("signal_t.npy")
        self.ids, self.n_sel, self.rff_w = ld("ids.npy"), ld("n_sel.npy"), np.load(os.path.join(self.dir, "rff_w.npy"))
        self.P, self.NP, self.T = self.meta["P"], self.meta["NP"], self.meta["T"]
        self.dates = np.array(self.meta["dates"], dtype="datetime64[D]")

    def reals(self, t):
        """(r_tilde [P], denom, risk, tc [P, P]) of date index t as fresh numpy arrays."""
        lib = _lib.load()
        P = self.P
        risk, tc, denom = (np.empty((P, P)) for _ in range(3))
        a, b = np.ascontiguousarray(self.risk_ut[t]), np.ascontiguousarray(self.tc_ut[t])
        p = lambda x: C.c_void_p(x.ctypes.data)
        check(lib.pfml_unpack_reals(p(a), p(b), P, p(risk), p(tc), p(denom)), "pfml_unpack_reals")
        return np.array(self.r_tilde[t]), denom, risk, tc

    def denom_stack(self):
        """denom [T, P, P] (what PFML_Search_Coef.py consumes), rebuilt date by date."""
        out = np.empty((self.T, self.P, self.P))
        for t in range(self.T):
            out[t] = self.reals(t)[1]
        return out

    def to_legacy(self):
        """The reference's ``pfml_input[g]`` dict (PFML_Input_Data.py:484-491): {'reals': {date: {...DataFrames}}, 'signal_t': ..., 'rff_w'}."""
        import pandas as pd
        from .general_functions import pfml_feat_fun
        feat_all = pfml_feat_fun(self.P - 1)
        reals, signal_t = {}, {}
        for t, d in enumerate(self.dates):
            ts = pd.Timestamp(d)
            r, den, risk, tc = self.reals(t)
            reals[ts] = {"r_tilde": pd.Series(r, index=feat_all), "denom": pd.DataFrame(den, index=feat_all, columns=feat_all),
                         "risk": pd.DataFrame(risk, index=feat_all, columns=feat_all), "tc": pd.DataFrame(tc, index=feat_all, columns=feat_all)}
            n = int(self.n_sel[t])
            signal_t[ts] = pd.DataFrame(np.array(self.signal_t[t, :n]), index=np.array(self.ids[t, :n]).astype(int), columns=feat_all)
        return {"reals": reals, "signal_t": signal_t, "rff_w": self.rff_w}
