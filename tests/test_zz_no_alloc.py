"""GPU: pfml_inverse_batched_static -- the inverse for callers that cannot allocate device memory.

Some users of this library may not call cudaMalloc at the point where they need the inverse (a latency-critical
section, a plugin inside somebody else's memory pool), so this entry point takes no workspace: its work matrices
live in a fixed array inside the library.  The test calls it through the C ABI exactly as such a caller would --
inputs and outputs only -- and checks it against pfml_inverse_batched (caller-provided workspace) on two calls in
a row: the second call finds the first call's leftovers in the shared array and must not be disturbed by them.
It also checks that a request over the documented size limit is refused with a clear message.

Collected last on purpose (file name), so a run that stops at the first failure has already run everything else.
"""
import ctypes as C

import numpy as np
import pytest

from conftest import pkg, relerr

pytestmark = pytest.mark.gpu


def _ptr(t):
    return C.c_void_p(t.data_ptr())


def _both(lib, torch, a):
    """Invert the batch `a` with the workspace entry point and with the static-scratch one."""
    B, n, _ = a.shape
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    ref = torch.empty_like(a)
    st_ref = torch.zeros(B, dtype=torch.int32, device=a.device)
    nbytes = lib.pfml_inverse_workspace_bytes(B, n)
    ws = torch.empty(nbytes, dtype=torch.uint8, device=a.device)
    assert lib.pfml_inverse_batched(B, n, _ptr(a), _ptr(ref), _ptr(st_ref), _ptr(ws), nbytes, stream) == 0
    got = torch.full_like(a, float("nan"))
    st = torch.full((B,), 7, dtype=torch.int32, device=a.device)
    rc = lib.pfml_inverse_batched_static(B, n, _ptr(a), _ptr(got), _ptr(st), stream)
    assert rc == 0, lib.pfml_last_error().decode()
    torch.cuda.synchronize()
    return ref.cpu().numpy(), got.cpu().numpy(), st_ref.cpu().numpy(), st.cpu().numpy()


def test_static_scratch_inverse_matches_workspace_inverse_twice():
    import torch
    assert torch.cuda.is_available()
    lib = pkg("_lib").load()
    rng = np.random.default_rng(12)
    dev = torch.device("cuda:0")
    with torch.cuda.device(dev):
        # call 1: two 200 x 200 matrices (NP = 256); call 2: different data AND a different shape, so the shared
        # array still holds call 1's blocks in other places
        for (B, n) in ((2, 200), (5, 100)):
            A = rng.standard_normal((B, n, n)) * (0.5 / np.sqrt(n)) + 2.0 * np.eye(n)
            a = torch.as_tensor(A, device=dev)
            ref, got, st_ref, st = _both(lib, torch, a)
            assert (st == 0).all() and (st_ref == 0).all()
            assert relerr(got, ref) < 1e-13                      # same kernels, same order
            assert relerr(got, np.linalg.inv(A)) < 1e-11         # tolerance of tests/test_gpu_parity.py::test_inverse


def test_static_scratch_rejects_oversize_request():
    import torch
    lib = pkg("_lib").load()
    limit = 1 << 19
    B, n = 3, 512                                                # 3 * 512 * 512 > 2^19
    assert B * lib.pfml_padded_n(n) ** 2 > limit
    a = torch.eye(n, dtype=torch.float64, device="cuda:0").repeat(B, 1, 1)
    out = torch.zeros_like(a)
    rc = lib.pfml_inverse_batched_static(B, n, _ptr(a), _ptr(out), C.c_void_p(0), C.c_void_p(0))
    assert rc == -3
    msg = lib.pfml_last_error().decode()
    assert "static scratch limit" in msg and "pfml_inverse_batched" in msg
    torch.cuda.synchronize()
    assert float(out.abs().sum()) == 0.0                         # nothing was launched
    # exactly at the limit is accepted
    B, n = 2, 512
    a = (2.0 * torch.eye(n, dtype=torch.float64, device="cuda:0")).repeat(B, 1, 1)
    out = torch.zeros_like(a)
    assert lib.pfml_inverse_batched_static(B, n, _ptr(a), _ptr(out), C.c_void_p(0), C.c_void_p(0)) == 0
    torch.cuda.synchronize()
    assert relerr(out.cpu().numpy(), np.broadcast_to(0.5 * np.eye(n), (B, n, n))) < 1e-15
