"""End-to-end entry over HOST buffers: the call a drop-in user of PFML_Input_Data.py makes.

``date_steps_host`` hands numpy inputs of one block of months to ``pfml_date_steps_host`` (C ABI), which
uploads them, computes RFF features, vol_scale and the date-steps on the GPU and writes the outputs back
into (pinned) host buffers.  Host<->device copies are part of the call.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib
from ._lib import HostBlockStruct, check
from .device import universe


def _c(a, dtype=np.float64):
    return np.ascontiguousarray(a, dtype=dtype)


class HostBlock:
    """Contiguous FP64 host copies of the months [t0, t1] of a panel + the universe of each requested date."""

    def __init__(self, panel, months, pin=True):
        lib = _lib.load()
        months = np.asarray(list(months))
        t0, t1 = int(months.min()) - panel.lb - 1, int(months.max())
        if t0 < 0:
            raise ValueError("date-steps need lb+1 look-back months inside the panel")
        sl = slice(t0, t1 + 1)
        self.t0 = t0
        T = t1 + 1 - t0
        N, F = panel.feats.shape[1:]
        K = panel.fct_cov.shape[1]
        self.T, self.N, self.F, self.K, self.p = T, N, F, K, panel.p
        self.P = panel.p + 1
        to = (lambda a: torch.from_numpy(a).pin_memory()) if pin else torch.from_numpy
        self.arrs = {
            "feats": to(_c(panel.feats[sl]).reshape(T * N, F)), "W": to(_c(panel.W)),
            "tr_ld0": to(_c(panel.tr_ld0[sl]).reshape(-1)), "mu_ld0": to(_c(panel.mu_ld0[sl]).reshape(-1)),
            "lam": to(_c(panel.lam[sl]).reshape(-1)), "ret_ld1": to(_c(panel.ret_ld1[sl]).reshape(-1)),
            "fct_load": to(_c(panel.fct_load[sl]).reshape(T * N, K)), "fct_cov": to(_c(panel.fct_cov[sl])),
            "ivol": to(_c(panel.ivol[sl]).reshape(-1)), "wealth": to(_c(panel.wealth[sl])), "rf": to(_c(panel.rf[sl])),
        }
        self.has = None
        if not bool(panel.barra_has[sl].all()):
            self.has = to(_c(panel.barra_has[sl].reshape(-1), np.uint8))
        self.sel, self.n_sel, self.n_max, self.NP = universe(panel, months)
        self.date_t = (months - t0).astype(np.int32)
        self.B = len(months)
        self.struct = HostBlockStruct(T=T, N=N, F=F, p=panel.p, K=K,
                                      barra_has=self.has.data_ptr() if self.has is not None else None,
                                      gamma_rel=float(panel.gamma_rel), mu=float(panel.mu), lb=int(panel.lb),
                                      iterations=int(panel.iterations),
                                      **{k: v.data_ptr() for k, v in self.arrs.items()})
        self.h2d_bytes = sum(v.numel() * v.element_size() for v in self.arrs.values()) + \
            (self.has.numel() if self.has is not None else 0) + self.sel.nbytes + self.n_sel.nbytes + self.date_t.nbytes
        B, P, NP = self.B, self.P, self.NP
        mk = (lambda *s: torch.empty(s, dtype=torch.float64).pin_memory()) if pin else (lambda *s: torch.empty(s, dtype=torch.float64))
        self.out = {"r_tilde": mk(B, P), "denom": mk(B, P, P), "risk": mk(B, P, P), "tc": mk(B, P, P),
                    "signal_t": mk(B, NP, P)}
        self.status = torch.zeros(B, dtype=torch.int32)
        self.mk = mk
        self.packed = None
        PT = P * (P + 1) // 2
        # bytes that cross PCIe per call: r_tilde, signal_t and the two packed triangles (both entry points)
        self.d2h_bytes = 8 * B * (P + NP * P + 2 * PT) + 4 * B
        self.d2h_bytes_full_squares = sum(v.numel() * 8 for v in self.out.values()) + 4 * B

    def run(self, chunk=0, device=0, full_outputs=True):
        lib = _lib.load()
        o = self.out
        opt = (lambda k: C.c_void_p(o[k].data_ptr())) if full_outputs else (lambda k: C.c_void_p(0))
        check(lib.pfml_date_steps_host(C.byref(self.struct), self.B, self.n_max, C.c_void_p(self.date_t.ctypes.data),
                                       C.c_void_p(self.sel.ctypes.data), C.c_void_p(self.n_sel.ctypes.data),
                                       C.c_void_p(o["r_tilde"].data_ptr()), C.c_void_p(o["denom"].data_ptr()),
                                       opt("risk"), opt("tc"), opt("signal_t"), C.c_void_p(self.status.data_ptr()),
                                       int(chunk), int(device)), "pfml_date_steps_host")
        return self


    def run_packed(self, chunk=0, device=0):
        """pfml_date_steps_host_packed: r_tilde, signal_t and the packed upper triangles of risk and tc (self.packed)."""
        lib = _lib.load()
        B, P = self.B, self.P
        if self.packed is None:
            PT = P * (P + 1) // 2
            self.packed = {"risk_ut": self.mk(B, PT), "tc_ut": self.mk(B, PT)}
        o, q = self.out, self.packed
        check(lib.pfml_date_steps_host_packed(C.byref(self.struct), B, self.n_max, C.c_void_p(self.date_t.ctypes.data),
                                              C.c_void_p(self.sel.ctypes.data), C.c_void_p(self.n_sel.ctypes.data),
                                              C.c_void_p(o["r_tilde"].data_ptr()), C.c_void_p(q["risk_ut"].data_ptr()),
                                              C.c_void_p(q["tc_ut"].data_ptr()), C.c_void_p(o["signal_t"].data_ptr()),
                                              C.c_void_p(self.status.data_ptr()), int(chunk), int(device)),
              "pfml_date_steps_host_packed")
        return self

    def unpack(self, b):
        """risk, tc, denom [P, P] of date b from the packed triangles (pfml_unpack_reals, host only)."""
        lib = _lib.load()
        P = self.P
        risk, tc, denom = (np.empty((P, P)) for _ in range(3))
        check(lib.pfml_unpack_reals(C.c_void_p(self.packed["risk_ut"][b].data_ptr()), C.c_void_p(self.packed["tc_ut"][b].data_ptr()), P,
                                    C.c_void_p(risk.ctypes.data), C.c_void_p(tc.ctypes.data), C.c_void_p(denom.ctypes.data)),
              "pfml_unpack_reals")
        return risk, tc, denom


def date_steps_host(panel, months, chunk=0, device=0, pin=True):
    """reals / signal_t of the given month indices (PFML_Input_Data.py:484-491) as numpy arrays."""
    blk = HostBlock(panel, months, pin=pin).run(chunk=chunk, device=device)
    out = []
    for b in range(blk.B):
        n = int(blk.n_sel[b])
        out.append({k: v[b].numpy() for k, v in blk.out.items() if k != "signal_t"})
        out[-1]["signal_t"] = blk.out["signal_t"][b, :n].numpy()
        out[-1]["ids"] = panel.ids[blk.sel[b, :n]]
        out[-1]["status"] = int(blk.status[b])
    return out
