"""The one numerical step of Prepare_Data.py that feeds the hot path with a GPU-sized cost: the
cross-sectional percentile rank of the characteristics (Prepare_Data.py:338-345) and the 0.5 imputation (:368).
Everything else in that script is pandas ETL on proprietary data and stays out of scope."""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib
from ._lib import check


def percentile_rank(feats, impute=True, device="cuda:0"):
    """feats [T, N, F] (numpy, NaN = missing) -> ranks in (0, 1], zeros restored, NaN -> 0.5 if ``impute`` (else NaN).
    Equals ``chars.groupby('eom')[features].rank(pct=True)`` + the zero mask + ``fillna(0.5)`` of the reference, bit for bit."""
    lib = _lib.load()
    dev = torch.device(device)
    x = torch.as_tensor(np.ascontiguousarray(feats, dtype=np.float64)).to(dev)
    T, N, F = x.shape
    out = torch.empty_like(x)
    with torch.cuda.device(dev):
        check(lib.pfml_percentile_rank(C.c_void_p(x.data_ptr()), T, N, F, int(bool(impute)), C.c_void_p(out.data_ptr()),
                                       C.c_void_p(torch.cuda.current_stream().cuda_stream)), "pfml_percentile_rank")
        torch.cuda.synchronize(dev)
    return out.cpu().numpy()
