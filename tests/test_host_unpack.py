"""CPU: the host-side rebuild of full squares from packed upper triangles (pfml_unpack_upper / pfml_unpack_reals) -- pure
host code of the C-ABI library, no GPU involved: exact copies and one IEEE addition per element."""
import ctypes as C

import numpy as np
import pytest

from conftest import pkg


@pytest.mark.parametrize("P", [1, 2, 31, 32, 33, 65, 257, 1025])
def test_unpack_reals_bit_exact(P):
    lib = pkg("_lib").load()
    rng = np.random.default_rng(P)
    A = rng.standard_normal((P, P)); A = np.triu(A) + np.triu(A, 1).T
    Bm = rng.standard_normal((P, P)); Bm = np.triu(Bm) + np.triu(Bm, 1).T
    iu = np.triu_indices(P)
    a_ut, b_ut = np.ascontiguousarray(A[iu]), np.ascontiguousarray(Bm[iu])      # row-major upper triangle: row r holds columns r..P-1
    risk, tc, denom = (np.full((P, P), np.nan) for _ in range(3))
    p = lambda x: C.c_void_p(x.ctypes.data)
    assert lib.pfml_unpack_reals(p(a_ut), p(b_ut), P, p(risk), p(tc), p(denom)) == 0
    assert np.array_equal(risk, A) and np.array_equal(tc, Bm) and np.array_equal(denom, A + Bm)
    full = np.full((P, P), np.nan)
    assert lib.pfml_unpack_upper(p(a_ut), P, p(full)) == 0
    assert np.array_equal(full, A)
    assert lib.pfml_unpack_upper(None, P, p(full)) == -1
