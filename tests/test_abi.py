"""CPU: the C-ABI shared library loads and exports every symbol include/pfml_b200.h declares (no compute)."""
import os
import re

from conftest import ROOT, pkg


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "pfml_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(pfml_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    L = pkg("_lib")
    lib = L.load()
    names = declared_symbols()
    assert len(names) >= 15
    for n in names:
        assert hasattr(lib, n), n
        assert n in L.SIGNATURES, f"{n} declared in the header but not bound in _lib.SIGNATURES"
    assert set(L.SIGNATURES) <= set(names)
    assert b"sm_100a" in lib.pfml_version()


def test_pure_host_queries():
    lib = pkg("_lib").load()
    assert lib.pfml_padded_n(500) == 512
    assert lib.pfml_padded_n(64) == 64
    assert lib.pfml_padded_n(65) == 128
    assert lib.pfml_mfunc_workspace_bytes(2, 100) > 0
    assert lib.pfml_inverse_workspace_bytes(1, 64) == 3 * 64 * 64 * 8
    assert lib.pfml_prof_ntags() >= 7
    assert lib.pfml_prof_tag_name(0) == b"dmma_gemm"


def test_missing_library_fails_loudly(monkeypatch):
    L = pkg("_lib")
    monkeypatch.setattr(L, "_lib", None)
    monkeypatch.setattr(L, "LIB_PATH", "/nonexistent/libpfml_b200.so")
    import pytest
    with pytest.raises(L.PfmlLibraryError):
        L.load()
