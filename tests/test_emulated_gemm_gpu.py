"""pfml_dgemm_emulated (FP64 products on the INT8 tensor cores) against numpy and against the DMMA kernel."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

TOL = 1e-13   # relative to sum_k |a_ik| |b_kj|: 7 balanced base-256 slices keep 54 bits below the row / column maximum


def _rel(got, A, B, want):
    scale = np.abs(A) @ np.abs(B)
    return float(np.max(np.abs(got - want) / np.maximum(scale, 1e-300)))


@pytest.mark.parametrize("batch,M,N,K", [(3, 128, 64, 64), (2, 200, 130, 128), (5, 512, 1025, 512), (1, 70, 9, 192)])
def test_emulated_gemm_matches_numpy(batch, M, N, K):
    from pfml_b200 import device
    rng = np.random.default_rng(batch * 1000 + M)
    A = rng.standard_normal((batch, M, K)) * np.exp(rng.uniform(-6, 6, (batch, M, 1)))     # rows of very different size
    B = rng.standard_normal((batch, K, N)) * np.exp(rng.uniform(-6, 6, (batch, 1, N)))     # and columns
    D, status = device.dgemm_emulated(A, B)
    assert status == 0
    assert _rel(D, A, B, A @ B) < TOL
    D2 = device.dgemm(A, B)
    assert _rel(D, A, B, D2) < TOL


def test_emulated_gemm_alpha_beta_and_zero_rows():
    from pfml_b200 import device
    rng = np.random.default_rng(7)
    A = rng.standard_normal((2, 128, 64))
    B = rng.standard_normal((2, 64, 64))
    A[:, 5] = 0.0                     # an all-zero row and column must give exact zeros
    B[:, :, 11] = 0.0
    D0 = rng.standard_normal((2, 128, 64))
    D, status = device.dgemm_emulated(A, B, alpha=-0.75, beta=2.0, D0=D0)
    assert status == 0
    want = -0.75 * (A @ B) + 2.0 * D0
    assert np.max(np.abs(D - want)) < 1e-11
    assert np.array_equal(D[:, 5], 2.0 * D0[:, 5])
    assert np.array_equal(D[:, :, 11], 2.0 * D0[:, :, 11])


def test_emulated_gemm_exact_on_small_integers():
    """Integers below 2^20 in magnitude are represented exactly by the slices, so the product is exact."""
    from pfml_b200 import device
    rng = np.random.default_rng(3)
    A = rng.integers(-1000, 1000, (2, 128, 128)).astype(np.float64)
    B = rng.integers(-1000, 1000, (2, 128, 64)).astype(np.float64)
    D, status = device.dgemm_emulated(A, B)
    assert status == 0
    assert np.array_equal(D, A @ B)


def test_emulated_gemm_flags_nan_and_rejects_shapes():
    from pfml_b200 import _lib, device
    A = np.ones((1, 128, 64))
    B = np.ones((1, 64, 64))
    A[0, 3, 7] = np.nan
    _, status = device.dgemm_emulated(A, B)
    assert status & 16
    B[0, 2, 2] = np.inf
    _, status = device.dgemm_emulated(np.ones((1, 128, 64)), B)
    assert status & 16
    lib = _lib.load()
    with pytest.raises(RuntimeError):
        device.dgemm_emulated(np.ones((1, 128, 100)), np.ones((1, 100, 64)))      # K % 64 != 0
    assert b"K" in lib.pfml_last_error()


def test_date_steps_same_with_and_without_emulation():
    """The n x n x P products of the date step on the INT8 path and on the DMMA path agree far inside 1e-8."""
    from conftest import relerr
    from pfml_b200 import _lib, device, synthetic
    lib = _lib.load()
    panel = synthetic.make_panel(N=150, T=15, p=64, seed=31, ragged=0.03)
    months = [12, 13, 14]
    try:
        lib.pfml_set_emulation(1)
        a = device.date_steps(panel, months)
        lib.pfml_set_emulation(0)
        b = device.date_steps(panel, months)
    finally:
        lib.pfml_set_emulation(1)
    for x, y in zip(a, b):
        assert x["status"] == 0 and y["status"] == 0
        for key in ("r_tilde", "denom", "risk", "tc", "signal_t"):
            assert relerr(x[key], y[key]) < 1e-11, key


@pytest.mark.parametrize("N,p,months", [(40, 4, [13]), (70, 2, [12, 13]), (130, 6, [14]), (500, 20, [12])])
def test_date_steps_odd_shapes_both_paths(N, p, months):
    """Tiny universes / feature counts (one k-stage, one tile, P < 64) and n = 500 with few features: the INT8 path
    (direct form, fused standardise + digits, symmetric a9 products at n = 512) against the DMMA path."""
    from conftest import relerr
    from pfml_b200 import _lib, device, synthetic
    lib = _lib.load()
    panel = synthetic.make_panel(N=N, T=15, p=p, seed=40 + N, ragged=0.02)
    try:
        lib.pfml_set_emulation(1)
        a = device.date_steps(panel, months)
        lib.pfml_set_emulation(0)
        b = device.date_steps(panel, months)
    finally:
        lib.pfml_set_emulation(1)
    for x, y in zip(a, b):
        assert x["status"] == 0 and y["status"] == 0
        for key in ("r_tilde", "denom", "risk", "tc", "signal_t"):
            assert relerr(x[key], y[key]) < 1e-10, (key, relerr(x[key], y[key]))
        assert (x["denom"] == x["denom"].T).all()


def test_date_steps_six_digit_mode():
    """pfml_set_emulation(2): 6 of the 7 digits for the products whose errors do not compound (n x P sums, solves, a9).
    Still far inside the 1e-8 tolerance: measured 7e-12 against the DMMA path at N = 500, p = 1024."""
    from conftest import relerr
    from pfml_b200 import _lib, device, synthetic
    import pfml_oracle as O
    lib = _lib.load()
    panel = synthetic.make_panel(N=500, T=14, p=64, seed=52, ragged=0.02)
    rff_x = O.rff(panel.feats.reshape(-1, panel.feats.shape[2]), panel.W).reshape(panel.feats.shape[0], panel.feats.shape[1], -1)
    vs = O.vol_scale_panel(panel)
    try:
        lib.pfml_set_emulation(2)
        a = device.date_steps(panel, [12])
        lib.pfml_set_emulation(0)
        b = device.date_steps(panel, [12])
    finally:
        lib.pfml_set_emulation(1)
    o = O.date_step(panel, 12, rff_x, vs)
    for key in ("r_tilde", "denom", "risk", "tc", "signal_t"):
        assert relerr(a[0][key], b[0][key]) < 1e-9, (key, relerr(a[0][key], b[0][key]))
        assert relerr(a[0][key], o[key]) < 1e-8, key
    assert a[0]["status"] == 0
