"""CPU: entry points reject bad arguments with a negative code and a message BEFORE any device work (no GPU needed: every
call here returns from its argument checks)."""
import ctypes as C

from conftest import pkg


def _err(lib):
    return lib.pfml_last_error().decode()


def test_limits_and_null_arguments():
    lib = pkg("_lib").load()
    one = C.c_double(0.0)
    p = C.c_void_p(C.addressof(one))          # a non-null dummy pointer: never dereferenced on these paths
    # percentile rank: more rows per month than the shared-memory sort holds
    assert lib.pfml_percentile_rank(p, 1, 5000, 3, 1, p, None) == -2 and "4096" in _err(lib)
    assert lib.pfml_percentile_rank(None, 1, 10, 3, 1, p, None) == -1
    # daily OLS: too many factors for the 28-wide augmented rows
    assert lib.pfml_xs_ols_daily(p, p, p, 4, 28, p, p, None, None) == -1 and "K >" in _err(lib)
    # factor covariance: window larger than the shared-memory weights; too many factors
    assert lib.pfml_wcov_wcor(p, p, 1, 25, 20000, p, p, 1.0, p, None) == -2 and "window" in _err(lib)
    assert lib.pfml_wcov_wcor(p, p, 1, 40, 100, p, p, 1.0, p, None) == -1
    # INT8 probes
    assert lib.pfml_int8_mma_probe(60, 64, 0, C.byref(one), None) == -1
    assert lib.pfml_int8_pipe_probe(64, 7, C.byref(one), None) == -1
    # emulated GEMM: K must be a multiple of 64
    assert lib.pfml_dgemm_emulated_workspace_bytes(128, 128, 100, 1) == 0
    assert lib.pfml_dgemm_emulated(128, 128, 100, 1.0, p, 100, 0, p, 128, 0, 0.0, p, 128, 0, 1, None, p, 1 << 20, None) == -2
    # ridge grid: null arguments, workspace too small
    assert lib.pfml_ridge_grid(None, p, p, 1, 8, p, 1, p, 1, p, None, p, 1 << 30, None) == -1
    pv = (C.c_int * 1)(8)
    assert lib.pfml_ridge_grid(p, p, p, 1, 8, pv, 1, p, 1, p, None, p, 16, None) == -3 and "workspace" in _err(lib)
    # communicator
    h = C.c_void_p()
    assert lib.pfml_comm_init(None, 0, 1, 0, C.byref(h)) == -1
    assert lib.pfml_comm_init(p, 3, 2, 0, C.byref(h)) == -1 and "rank 3 of 2" in _err(lib)
    assert lib.pfml_comm_world(None) == 1 and lib.pfml_comm_rank(None) == 0
    assert lib.pfml_comm_allreduce_sums(None, p, p, None, 4, 2, None) == 0          # no communicator: the partial sums are the sums
    # multi-g: G out of range, null panel list
    assert lib.pfml_date_steps_multi_g(None, 1, p, 1, 10, p, p, p, None, 0, None, None, p, p, None, None, None, None, p, 0, None) == -1
    # host unpack helpers
    assert lib.pfml_unpack_reals(None, p, 4, p, p, p) == -1


def test_host_entry_point_rejects_before_upload():
    """pfml_date_steps_host validates dates / universes on the host: -2 with a message, nothing is allocated or uploaded."""
    L = pkg("_lib")
    lib = L.load()
    import numpy as np
    T, N, F, p, K, lb = 16, 10, 3, 4, 2, 11
    arr = lambda *s: np.zeros(s)
    a = dict(feats=arr(T * N, F), W=arr(F, p // 2), tr_ld0=arr(T * N), mu_ld0=arr(T * N), lam=arr(T * N) + 1, ret_ld1=arr(T * N),
             fct_load=arr(T * N, K), fct_cov=arr(T, K, K), ivol=arr(T * N), wealth=arr(T) + 1, rf=arr(T))
    blk = L.HostBlockStruct(T=T, N=N, F=F, p=p, K=K, barra_has=None, gamma_rel=10.0, mu=0.007, lb=lb, iterations=10,
                            **{k: v.ctypes.data for k, v in a.items()})
    NP = lib.pfml_padded_n(N)
    P = p + 1
    out = arr(2, P), arr(2, P, P)
    call = lambda date_t, sel, n_sel: lib.pfml_date_steps_host(
        C.byref(blk), 2, N, C.c_void_p(date_t.ctypes.data), C.c_void_p(sel.ctypes.data), C.c_void_p(n_sel.ctypes.data),
        C.c_void_p(out[0].ctypes.data), C.c_void_p(out[1].ctypes.data), None, None, None, None, 0, 0)
    sel = np.tile(np.arange(NP, dtype=np.int32) % N, (2, 1)).copy()
    n_sel = np.array([N, N], dtype=np.int32)
    assert call(np.array([3, 13], dtype=np.int32), sel, n_sel) == -2 and "date_t[0]" in _err(lib)
    assert call(np.array([13, 16], dtype=np.int32), sel, n_sel) == -2 and "date_t[1]" in _err(lib)
    assert call(np.array([13, 14], dtype=np.int32), sel, np.array([N, 0], dtype=np.int32)) == -2 and "n_sel[1]" in _err(lib)
    bad = sel.copy(); bad[1, 2] = N
    assert call(np.array([13, 14], dtype=np.int32), bad, n_sel) == -2 and "sel[1][2]" in _err(lib)
    blk.K = 40
    assert call(np.array([13, 14], dtype=np.int32), sel, n_sel) == -2 and "K = 40" in _err(lib)
