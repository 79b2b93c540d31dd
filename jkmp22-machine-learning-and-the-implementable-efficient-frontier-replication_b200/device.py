"""Device-side driver of the PFML date-steps: torch owns memory and streams, libpfml_b200.so does the math.

``DevicePanel`` mirrors the panel-level part of PFML_Input_Data.py (rff :245-256, vol_scale :274-305) and
``date_steps`` the per-date loop (:318-491).  Every arithmetic step is a call through the C ABI.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib
from ._lib import PanelStruct, check


def _stream_ptr():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)


def _dev(a, device, dtype=torch.float64):
    return torch.as_tensor(np.ascontiguousarray(a), dtype=dtype).to(device)


class DevicePanel:
    """Panel resident in HBM: characteristics -> RFF features, Barra triples -> vol_scale, g_t, ...

    Layout: dense [T*N] rows (month-major), FP64, row-major; see pfml_panel_t in include/pfml_b200.h.
    """

    def __init__(self, panel, device="cuda:0", extra_W=()):
        """``extra_W``: further random-feature matrices (one per additional bandwidth g, quirk Q10); their feature panels
        are kept in ``self.rffs`` (``self.rffs[0]`` is the panel of ``panel.W``)."""
        lib = _lib.load()
        self.device = torch.device(device)
        self.host = panel
        T, N, F = panel.feats.shape
        K = panel.fct_cov.shape[1]
        self.T, self.N, self.F, self.K, self.p = T, N, F, K, panel.p
        with torch.cuda.device(self.device):
            s = _stream_ptr()
            feats = _dev(panel.feats.reshape(T * N, F), self.device)
            W = _dev(panel.W, self.device)
            self.W = W
            self.rff = torch.empty((T * N, self.p), dtype=torch.float64, device=self.device)
            check(lib.pfml_rff_sincos(_ptr(feats), _ptr(W), T * N, F, self.p // 2, _ptr(self.rff), s), "pfml_rff_sincos")
            self.rffs = [self.rff]
            for Wx in extra_W:
                Wd = _dev(Wx, self.device)
                r = torch.empty((T * N, self.p), dtype=torch.float64, device=self.device)
                check(lib.pfml_rff_sincos(_ptr(feats), _ptr(Wd), T * N, F, self.p // 2, _ptr(r), s), "pfml_rff_sincos")
                self.rffs.append(r)
            self.fct_load = _dev(panel.fct_load.reshape(T * N, K), self.device)
            self.fct_cov = _dev(panel.fct_cov, self.device)
            self.ivol = _dev(panel.ivol.reshape(-1), self.device)
            has = _dev(panel.barra_has.reshape(-1).astype(np.uint8), self.device, torch.uint8)
            self.vol_scale = torch.empty(T * N, dtype=torch.float64, device=self.device)
            has_arg = _ptr(has) if not bool(panel.barra_has.all()) else C.c_void_p(0)
            check(lib.pfml_vol_scale(_ptr(self.fct_load), _ptr(self.fct_cov), _ptr(self.ivol), has_arg, T, N, K,
                                     _ptr(self.vol_scale), s), "pfml_vol_scale")
            tr = _dev(panel.tr_ld0.reshape(-1), self.device)
            mu = _dev(panel.mu_ld0.reshape(-1), self.device)
            self.gt = torch.empty(T * N, dtype=torch.float64, device=self.device)
            check(lib.pfml_growth_ratio(_ptr(tr), _ptr(mu), T * N, _ptr(self.gt), s), "pfml_growth_ratio")
            self.lam = _dev(panel.lam.reshape(-1), self.device)
            self.ret_ld1 = _dev(panel.ret_ld1.reshape(-1), self.device)
            self.wealth = _dev(panel.wealth, self.device)
            self.rf = _dev(panel.rf, self.device)
            torch.cuda.current_stream().synchronize()
        self.struct = PanelStruct(T=T, N=N, F=F, p=self.p, K=K, rff=self.rff.data_ptr(),
                                  vol_scale=self.vol_scale.data_ptr(), gt=self.gt.data_ptr(), lam=self.lam.data_ptr(),
                                  ret_ld1=self.ret_ld1.data_ptr(), fct_load=self.fct_load.data_ptr(),
                                  fct_cov=self.fct_cov.data_ptr(), ivol=self.ivol.data_ptr(),
                                  wealth=self.wealth.data_ptr(), rf=self.rf.data_ptr(),
                                  gamma_rel=float(panel.gamma_rel), mu=float(panel.mu), lb=int(panel.lb),
                                  iterations=int(panel.iterations))


def universe(panel, months):
    """Host-side index logic of PFML_Input_Data.py:321-327: the valid ids of each date, ascending."""
    lib = _lib.load()
    sels = [np.flatnonzero(panel.valid[t]).astype(np.int32) for t in months]
    n_sel = np.array([len(s) for s in sels], dtype=np.int32)
    n_max = int(n_sel.max())
    NP = lib.pfml_padded_n(n_max)
    sel = np.zeros((len(months), NP), dtype=np.int32)
    for b, s in enumerate(sels):
        sel[b, :len(s)] = s
    return sel, n_sel, n_max, NP


class MultiGRunner:
    """B date-steps for the G bandwidths of a DevicePanel built with ``extra_W`` in one call (pfml_date_steps_multi_g):
    m, the chains and the two inverses are shared over g.  Outputs per g: r_tilde [G, B, P], denom [G, B, P, P] (risk / tc
    / signal_t optional); ``run(bucket, nb, r_sum, d_sum)`` also accumulates the per-window sums r_sum [G, nb, P],
    d_sum [G, nb, P, P]."""

    def __init__(self, dpanel, B, n_max, want_risk_tc=False, want_signal=False):
        lib = _lib.load()
        self.dp, self.B, self.n_max = dpanel, B, n_max
        self.G = len(dpanel.rffs)
        self.NP = lib.pfml_padded_n(n_max)
        P = self.P = dpanel.p + 1
        dev = dpanel.device
        f64 = dict(dtype=torch.float64, device=dev)
        G = self.G
        self.r_tilde = torch.empty((G, B, P), **f64)
        self.denom = torch.empty((G, B, P, P), **f64)
        self.risk = torch.empty((G, B, P, P), **f64) if want_risk_tc else None
        self.tc = torch.empty((G, B, P, P), **f64) if want_risk_tc else None
        self.signal_t = torch.empty((G, B, self.NP, P), **f64) if want_signal else None
        self.status = torch.zeros(B, dtype=torch.int32, device=dev)
        self.ws_bytes = lib.pfml_date_steps_workspace_bytes(C.byref(dpanel.struct), B, n_max)
        self.ws = torch.empty(self.ws_bytes, dtype=torch.uint8, device=dev)
        self.date_t = torch.empty(B, dtype=torch.int32, device=dev)
        self.sel = torch.empty((B, self.NP), dtype=torch.int32, device=dev)
        self.n_sel = torch.empty(B, dtype=torch.int32, device=dev)

    set_batch = None      # bound below (same as DateStepRunner.set_batch)

    @staticmethod
    def _ptrs(t, G):
        arr = (C.c_void_p * G)()
        for g in range(G):
            arr[g] = t[g].data_ptr() if t is not None else None
        return arr

    def run(self, bucket=None, n_buckets=0, r_sum=None, d_sum=None):
        lib = _lib.load()
        G = self.G
        rff = self._ptrs(self.dp.rffs, G)
        opt = lambda t: self._ptrs(t, G) if t is not None else None
        with torch.cuda.device(self.dp.device):
            check(lib.pfml_date_steps_multi_g(C.byref(self.dp.struct), G, rff, self.b, self.n_max, _ptr(self.date_t), _ptr(self.sel),
                                              _ptr(self.n_sel), _ptr(bucket) if bucket is not None else None, int(n_buckets),
                                              opt(r_sum), opt(d_sum), self._ptrs(self.r_tilde, G), self._ptrs(self.denom, G),
                                              opt(self.risk), opt(self.tc), opt(self.signal_t), _ptr(self.status), _ptr(self.ws),
                                              self.ws_bytes, _stream_ptr()), "pfml_date_steps_multi_g")


class DateStepRunner:
    """Pre-allocated outputs + workspace for batches of B date-steps on one GPU."""

    def __init__(self, dpanel: DevicePanel, B, n_max, want_risk_tc=True, want_signal=True):
        lib = _lib.load()
        self.dp = dpanel
        self.B, self.n_max = B, n_max
        self.NP = lib.pfml_padded_n(n_max)
        P = dpanel.p + 1
        self.P = P
        dev = dpanel.device
        f64 = dict(dtype=torch.float64, device=dev)
        self.r_tilde = torch.empty((B, P), **f64)
        self.denom = torch.empty((B, P, P), **f64)
        self.risk = torch.empty((B, P, P), **f64) if want_risk_tc else None
        self.tc = torch.empty((B, P, P), **f64) if want_risk_tc else None
        self.signal_t = torch.empty((B, self.NP, P), **f64) if want_signal else None
        self.status = torch.zeros(B, dtype=torch.int32, device=dev)
        self.ws_bytes = lib.pfml_date_steps_workspace_bytes(C.byref(dpanel.struct), B, n_max)
        self.ws = torch.empty(self.ws_bytes, dtype=torch.uint8, device=dev)
        self.date_t = torch.empty(B, dtype=torch.int32, device=dev)
        self.sel = torch.empty((B, self.NP), dtype=torch.int32, device=dev)
        self.n_sel = torch.empty(B, dtype=torch.int32, device=dev)

    def set_batch(self, months, sel, n_sel):
        b = len(months)
        assert b <= self.B and sel.shape[1] == self.NP
        self.date_t[:b].copy_(torch.as_tensor(np.asarray(months, dtype=np.int32)))
        self.sel[:b].copy_(torch.as_tensor(sel))
        self.n_sel[:b].copy_(torch.as_tensor(n_sel))
        self.b = b

    def run(self):
        lib = _lib.load()
        with torch.cuda.device(self.dp.device):
            check(lib.pfml_date_steps(C.byref(self.dp.struct), self.b, self.n_max, _ptr(self.date_t), _ptr(self.sel),
                                      _ptr(self.n_sel), _ptr(self.r_tilde), _ptr(self.denom), _ptr(self.risk),
                                      _ptr(self.tc), _ptr(self.signal_t), _ptr(self.status), _ptr(self.ws),
                                      self.ws_bytes, _stream_ptr()), "pfml_date_steps")


    def run_bucketed(self, bucket, n_buckets, r_sum, d_sum):
        """The same step with the expanding-window sums fused in (pfml_date_steps_bucketed): ``bucket`` int32 device
        tensor [b], ``r_sum`` [n_buckets, P] / ``d_sum`` [n_buckets, P, P] are accumulated into."""
        lib = _lib.load()
        with torch.cuda.device(self.dp.device):
            check(lib.pfml_date_steps_bucketed(C.byref(self.dp.struct), self.b, self.n_max, _ptr(self.date_t), _ptr(self.sel),
                                               _ptr(self.n_sel), _ptr(bucket), int(n_buckets), _ptr(r_sum), _ptr(d_sum),
                                               _ptr(self.r_tilde), _ptr(self.denom), _ptr(self.risk), _ptr(self.tc),
                                               _ptr(self.signal_t), _ptr(self.status), _ptr(self.ws), self.ws_bytes, _stream_ptr()),
                  "pfml_date_steps_bucketed")


MultiGRunner.set_batch = DateStepRunner.set_batch


def date_steps(panel, months, device="cuda:0", dpanel=None):
    """Run the date-steps of ``months`` (month indices into the panel) and return numpy outputs per date."""
    months = list(months)
    dpanel = dpanel or DevicePanel(panel, device)
    sel, n_sel, n_max, NP = universe(panel, months)
    runner = DateStepRunner(dpanel, len(months), n_max)
    runner.set_batch(months, sel, n_sel)
    runner.run()
    torch.cuda.synchronize(dpanel.device)
    status = runner.status.cpu().numpy()
    out = []
    for b, t in enumerate(months):
        n = int(n_sel[b])
        out.append(dict(r_tilde=runner.r_tilde[b].cpu().numpy(), denom=runner.denom[b].cpu().numpy(),
                        risk=runner.risk[b].cpu().numpy(), tc=runner.tc[b].cpu().numpy(),
                        signal_t=runner.signal_t[b, :n].cpu().numpy(), ids=panel.ids[sel[b, :n]], status=int(status[b])))
    return out


class _StageCall:
    """Shared set-up of the stage-level entry points: device copies of date_t / sel / n_sel and the date-step workspace."""

    def __init__(self, panel, months, device="cuda:0", dpanel=None):
        self.lib = _lib.load()
        self.months = [int(t) for t in months]
        self.dp = dpanel or DevicePanel(panel, device)
        self.sel, self.n_sel, self.n_max, self.NP = universe(panel, self.months)
        dev = self.dp.device
        self.B, self.P = len(self.months), self.dp.p + 1
        self.date_t = torch.as_tensor(np.asarray(self.months, dtype=np.int32)).to(dev)
        self.sel_d = torch.as_tensor(self.sel).to(dev)
        self.n_sel_d = torch.as_tensor(self.n_sel).to(dev)
        self.status = torch.zeros(self.B, dtype=torch.int32, device=dev)
        self.ws_bytes = self.lib.pfml_date_steps_workspace_bytes(C.byref(self.dp.struct), self.B, self.n_max)
        self.ws = torch.empty(self.ws_bytes, dtype=torch.uint8, device=dev)

    def head(self):
        return (C.byref(self.dp.struct), self.B, self.n_max, _ptr(self.date_t), _ptr(self.sel_d), _ptr(self.n_sel_d))

    def tail(self):
        return (_ptr(self.ws), self.ws_bytes, _stream_ptr())

    def empty(self, *shape):
        return torch.empty(shape, dtype=torch.float64, device=self.dp.device)


def stage_signals(panel, months, device="cuda:0", dpanel=None):
    """a6 alone (pfml_standardize_scale): signals[tau] = diag(1/vol_scale) s, PFML_Input_Data.py:357-388.
    Returns one array [lb+2, n, P] per date (columns in feat_all order)."""
    c = _StageCall(panel, months, device, dpanel)
    nt = panel.lb + 2
    out = c.empty(c.B, nt, c.NP, c.P)
    with torch.cuda.device(c.dp.device):
        check(c.lib.pfml_standardize_scale(*c.head(), _ptr(out), *c.tail()), "pfml_standardize_scale")
        torch.cuda.synchronize(c.dp.device)
    return [out[b, :, :int(c.n_sel[b])].cpu().numpy() for b in range(c.B)]


def stage_omega(panel, months, device="cuda:0", dpanel=None):
    """a3..a8 alone (pfml_omega_batched): omega, omega_chg (PFML_Input_Data.py:455-459) and m (General_functions.py:962)."""
    c = _StageCall(panel, months, device, dpanel)
    om, omc, m = c.empty(c.B, c.NP, c.P), c.empty(c.B, c.NP, c.P), c.empty(c.B, c.NP, c.NP)
    with torch.cuda.device(c.dp.device):
        check(c.lib.pfml_omega_batched(*c.head(), _ptr(om), _ptr(omc), _ptr(m), _ptr(c.status), *c.tail()), "pfml_omega_batched")
        torch.cuda.synchronize(c.dp.device)
    st = c.status.cpu().numpy()
    return [dict(omega=om[b, :n].cpu().numpy(), omega_chg=omc[b, :n].cpu().numpy(), m=m[b, :n, :n].cpu().numpy(), status=int(st[b]))
            for b, n in enumerate(int(x) for x in c.n_sel)]


def stage_rtilde_denom(panel, months, omegas, omega_chgs, device="cuda:0", dpanel=None):
    """a9 alone (pfml_rtilde_denom_batched) from given omega / omega_chg [n, P] per date (PFML_Input_Data.py:470-481)."""
    c = _StageCall(panel, months, device, dpanel)
    om, omc = np.zeros((c.B, c.NP, c.P)), np.zeros((c.B, c.NP, c.P))
    for b in range(c.B):
        om[b, :omegas[b].shape[0]] = omegas[b]
        omc[b, :omega_chgs[b].shape[0]] = omega_chgs[b]
    om_d, omc_d = _dev(om, c.dp.device), _dev(omc, c.dp.device)
    rt, dn, rk, tc = c.empty(c.B, c.P), c.empty(c.B, c.P, c.P), c.empty(c.B, c.P, c.P), c.empty(c.B, c.P, c.P)
    with torch.cuda.device(c.dp.device):
        check(c.lib.pfml_rtilde_denom_batched(*c.head(), _ptr(om_d), _ptr(omc_d), _ptr(rt), _ptr(dn), _ptr(rk), _ptr(tc),
                                              _ptr(c.status), *c.tail()), "pfml_rtilde_denom_batched")
        torch.cuda.synchronize(c.dp.device)
    st = c.status.cpu().numpy()
    return [dict(r_tilde=rt[b].cpu().numpy(), denom=dn[b].cpu().numpy(), risk=rk[b].cpu().numpy(), tc=tc[b].cpu().numpy(),
                 status=int(st[b])) for b in range(c.B)]


def m_func_batched(w, mu, rf, sigma_gam, gam, lam, iterations, device="cuda:0"):
    """Batched m_func (General_functions.py:919-962): sigma_gam [B,n,n], lam [B,n] (diag of K_Lambda), w/rf [B]."""
    lib = _lib.load()
    dev = torch.device(device)
    sg = _dev(sigma_gam, dev)
    B, n, _ = sg.shape
    lam_d, w_d, rf_d = _dev(lam, dev), _dev(np.broadcast_to(w, (B,)), dev), _dev(np.broadcast_to(rf, (B,)), dev)
    out = torch.empty((B, n, n), dtype=torch.float64, device=dev)
    status = torch.zeros(B, dtype=torch.int32, device=dev)
    nbytes = lib.pfml_mfunc_workspace_bytes(B, n)
    ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        check(lib.pfml_mfunc_batched(B, n, _ptr(sg), _ptr(lam_d), _ptr(w_d), _ptr(rf_d), float(mu), float(gam),
                                     int(iterations), _ptr(out), _ptr(status), _ptr(ws), nbytes, _stream_ptr()),
              "pfml_mfunc_batched")
        torch.cuda.synchronize(dev)
    return out.cpu().numpy(), status.cpu().numpy()


def dgemm(A, B, trans_a=False, trans_b=False, alpha=1.0, device="cuda:0"):
    """D = alpha op(A) op(B) for [batch, ., .] arrays through pfml_dgemm_batched."""
    lib = _lib.load()
    dev = torch.device(device)
    a, b = _dev(A, dev), _dev(B, dev)
    batch = a.shape[0]
    M = a.shape[2] if trans_a else a.shape[1]
    K = a.shape[1] if trans_a else a.shape[2]
    N = b.shape[1] if trans_b else b.shape[2]
    d = torch.empty((batch, M, N), dtype=torch.float64, device=dev)
    with torch.cuda.device(dev):
        check(lib.pfml_dgemm_batched(int(trans_a), int(trans_b), M, N, K, float(alpha), _ptr(a), a.shape[2],
                                     a.shape[1] * a.shape[2], _ptr(b), b.shape[2], b.shape[1] * b.shape[2], 0.0, _ptr(d),
                                     N, M * N, batch, _stream_ptr()), "pfml_dgemm_batched")
        torch.cuda.synchronize(dev)
    return d.cpu().numpy()


def dgemm_emulated(A, B, alpha=1.0, beta=0.0, D0=None, device="cuda:0"):
    """D = alpha A B + beta D0 for [batch, M, K] x [batch, K, N] on the INT8 tensor cores (pfml_dgemm_emulated).
    Returns (D numpy, status int): status bit 3 = internal time-out, bit 4 = NaN/Inf in an operand."""
    lib = _lib.load()
    dev = torch.device(device)
    a, b = _dev(A, dev), _dev(B, dev)
    batch, M, K = a.shape
    N = b.shape[2]
    d = torch.zeros((batch, M, N), dtype=torch.float64, device=dev) if D0 is None else _dev(D0, dev).clone()
    status = torch.zeros(1, dtype=torch.int32, device=dev)
    nbytes = lib.pfml_dgemm_emulated_workspace_bytes(M, N, K, batch)
    ws = torch.empty(max(nbytes, 1), dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        check(lib.pfml_dgemm_emulated(M, N, K, float(alpha), _ptr(a), K, M * K, _ptr(b), N, K * N, float(beta), _ptr(d), N,
                                      M * N, batch, _ptr(status), _ptr(ws), nbytes, _stream_ptr()), "pfml_dgemm_emulated")
        torch.cuda.synchronize(dev)
    return d.cpu().numpy(), int(status.item())


def inverse(A, device="cuda:0"):
    lib = _lib.load()
    dev = torch.device(device)
    a = _dev(A, dev)
    B, n, _ = a.shape
    out = torch.empty_like(a)
    status = torch.zeros(B, dtype=torch.int32, device=dev)
    nbytes = lib.pfml_inverse_workspace_bytes(B, n)
    ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        check(lib.pfml_inverse_batched(B, n, _ptr(a), _ptr(out), _ptr(status), _ptr(ws), nbytes, _stream_ptr()),
              "pfml_inverse_batched")
        torch.cuda.synchronize(dev)
    return out.cpu().numpy(), status.cpu().numpy()
