"""ctypes binding of csrc/libpfml_b200.so (the C ABI declared in include/pfml_b200.h).

There is no CPU fallback: if the shared library is missing or a symbol cannot be resolved, importing the
compute modules raises ``PfmlLibraryError`` -- build it with ``python -c "import __graft_entry__ as g; g.build()"``
(or ``make -C <package>/csrc``).
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "csrc", "libpfml_b200.so")


class PfmlLibraryError(RuntimeError):
    pass


class PfmlError(RuntimeError):
    pass


c_double_p = C.c_void_p      # device or host pointers are passed as raw addresses
c_int_p = C.c_void_p


class PanelStruct(C.Structure):
    """pfml_panel_t"""
    _fields_ = [("T", C.c_int), ("N", C.c_int), ("F", C.c_int), ("p", C.c_int), ("K", C.c_int),
                ("rff", C.c_void_p), ("vol_scale", C.c_void_p), ("gt", C.c_void_p), ("lam", C.c_void_p),
                ("ret_ld1", C.c_void_p), ("fct_load", C.c_void_p), ("fct_cov", C.c_void_p), ("ivol", C.c_void_p),
                ("wealth", C.c_void_p), ("rf", C.c_void_p),
                ("gamma_rel", C.c_double), ("mu", C.c_double), ("lb", C.c_int), ("iterations", C.c_int)]


class HostBlockStruct(C.Structure):
    """pfml_host_block_t"""
    _fields_ = [("T", C.c_int), ("N", C.c_int), ("F", C.c_int), ("p", C.c_int), ("K", C.c_int),
                ("feats", C.c_void_p), ("W", C.c_void_p), ("tr_ld0", C.c_void_p), ("mu_ld0", C.c_void_p),
                ("lam", C.c_void_p), ("ret_ld1", C.c_void_p), ("fct_load", C.c_void_p), ("fct_cov", C.c_void_p),
                ("ivol", C.c_void_p), ("wealth", C.c_void_p), ("rf", C.c_void_p), ("barra_has", C.c_void_p),
                ("gamma_rel", C.c_double), ("mu", C.c_double), ("lb", C.c_int), ("iterations", C.c_int)]


# name -> (restype, argtypes); must list every symbol include/pfml_b200.h declares
SIGNATURES = {
    "pfml_version": (C.c_char_p, []),
    "pfml_last_error": (C.c_char_p, []),
    "pfml_rff_sincos": (C.c_int, [C.c_void_p, C.c_void_p, C.c_long, C.c_int, C.c_int, C.c_void_p, C.c_void_p]),
    "pfml_vol_scale": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p,
                                 C.c_void_p]),
    "pfml_growth_ratio": (C.c_int, [C.c_void_p, C.c_void_p, C.c_long, C.c_void_p, C.c_void_p]),
    "pfml_padded_n": (C.c_int, [C.c_int]),
    "pfml_date_steps_workspace_bytes": (C.c_size_t, [C.POINTER(PanelStruct), C.c_int, C.c_int]),
    "pfml_date_steps": (C.c_int, [C.POINTER(PanelStruct), C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p,
                                  C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                  C.c_void_p, C.c_size_t, C.c_void_p]),
    "pfml_date_steps_bucketed": (C.c_int, [C.POINTER(PanelStruct), C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p,
                                           C.c_void_p, C.c_int, C.c_void_p, C.c_void_p,
                                           C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                           C.c_void_p, C.c_size_t, C.c_void_p]),
    "pfml_date_steps_multi_g": (C.c_int, [C.POINTER(PanelStruct), C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                          C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                          C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]),
    "pfml_standardize_scale": (C.c_int, [C.POINTER(PanelStruct), C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p,
                                         C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]),
    "pfml_omega_batched": (C.c_int, [C.POINTER(PanelStruct), C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p,
                                     C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]),
    "pfml_rtilde_denom_batched": (C.c_int, [C.POINTER(PanelStruct), C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p,
                                            C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                            C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]),
    "pfml_mfunc_workspace_bytes": (C.c_size_t, [C.c_int, C.c_int]),
    "pfml_mfunc_batched": (C.c_int, [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_double,
                                     C.c_double, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]),
    "pfml_bucket_accumulate": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p,
                                         C.c_void_p, C.c_void_p]),
    "pfml_bucket_cumsum": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p]),
    "pfml_ridge_grid_workspace_bytes": (C.c_size_t, [C.c_int, C.c_int, C.c_int]),
    "pfml_ridge_grid": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_void_p,
                                  C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]),
    "pfml_comm_unique_id": (C.c_int, [C.c_void_p]),
    "pfml_comm_init": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_void_p)]),
    "pfml_comm_rank": (C.c_int, [C.c_void_p]),
    "pfml_comm_world": (C.c_int, [C.c_void_p]),
    "pfml_comm_nccl_version": (C.c_int, []),
    "pfml_comm_destroy": (C.c_int, [C.c_void_p]),
    "pfml_comm_allreduce_sums": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p]),
    "pfml_ridge_shard_range": (None, [C.c_int, C.c_int, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "pfml_ridge_grid_range": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_void_p,
                                        C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]),
    "pfml_ridge_grid_sharded": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int,
                                          C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]),
    "pfml_xs_ols_daily": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p,
                                    C.c_void_p]),
    "pfml_wcov_wcor": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_double,
                                 C.c_void_p, C.c_void_p]),
    "pfml_ewma_vol": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_double, C.c_int, C.c_void_p, C.c_void_p]),
    "pfml_validation_utility": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int,
                                          C.c_void_p, C.c_void_p]),
    "pfml_aim_weights": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
    "pfml_mfunc_dates_workspace_bytes": (C.c_size_t, [C.POINTER(PanelStruct), C.c_int, C.c_int]),
    "pfml_mfunc_dates": (C.c_int, [C.POINTER(PanelStruct), C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                   C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]),
    "pfml_backtest": (C.c_int, [C.POINTER(PanelStruct), C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p,
                                C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                C.c_void_p]),
    "pfml_dgemm_batched": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_double, C.c_void_p, C.c_long,
                                     C.c_long, C.c_void_p, C.c_long, C.c_long, C.c_double, C.c_void_p, C.c_long,
                                     C.c_long, C.c_int, C.c_void_p]),
    "pfml_dgemm_emulated_workspace_bytes": (C.c_size_t, [C.c_int, C.c_int, C.c_int, C.c_int]),
    "pfml_dgemm_emulated": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_double, C.c_void_p, C.c_long, C.c_long, C.c_void_p, C.c_long,
                                      C.c_long, C.c_double, C.c_void_p, C.c_long, C.c_long, C.c_int, C.c_void_p, C.c_void_p,
                                      C.c_size_t, C.c_void_p]),
    "pfml_set_emulation": (None, [C.c_int]),
    "pfml_inverse_workspace_bytes": (C.c_size_t, [C.c_int, C.c_int]),
    "pfml_inverse_batched": (C.c_int, [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t,
                                       C.c_void_p]),
    "pfml_inverse_batched_static": (C.c_int, [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "pfml_date_steps_host": (C.c_int, [C.POINTER(HostBlockStruct), C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p,
                                       C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                       C.c_int, C.c_int]),
    "pfml_date_steps_host_packed": (C.c_int, [C.POINTER(HostBlockStruct), C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p,
                                              C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int]),
    "pfml_unpack_upper": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p]),
    "pfml_unpack_reals": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
    "pfml_percentile_rank": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]),
    "pfml_int8_pipe_probe": (C.c_int, [C.c_int, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double)]),
    "pfml_int8_mma_probe": (C.c_int, [C.c_int, C.c_int, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double)]),
    "pfml_prof_enable": (None, [C.c_int]),
    "pfml_prof_ntags": (C.c_int, []),
    "pfml_prof_tag_name": (C.c_char_p, [C.c_int]),
    "pfml_prof_collect": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p]),
    "pfml_launch_count": (C.c_long, []),
}

_lib = None


def load():
    """Load the shared library once and bind every declared symbol (no GPU needed for this step)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise PfmlLibraryError(
            f"{LIB_PATH} not found: the CUDA extension has not been built. There is no CPU fallback; "
            "run `python -c 'import __graft_entry__ as g; g.build()'` first.")
    try:
        lib = C.CDLL(LIB_PATH)
    except OSError as e:
        raise PfmlLibraryError(f"cannot load {LIB_PATH}: {e}") from e
    for name, (res, args) in SIGNATURES.items():
        try:
            fn = getattr(lib, name)
        except AttributeError as e:
            raise PfmlLibraryError(f"{LIB_PATH} does not export {name}") from e
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc, what="pfml call"):
    if rc != 0:
        msg = load().pfml_last_error().decode()
        raise PfmlError(f"{what} failed with code {rc}: {msg}")
