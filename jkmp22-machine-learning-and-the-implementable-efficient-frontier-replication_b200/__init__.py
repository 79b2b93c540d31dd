"""pfml_b200 -- the Portfolio-ML (PFML) hot path of the JKMP22 replication, hand-written for NVIDIA B200.

Host code is Python (same entry-point names as the reference's PFML_Input_Data.py, PFML_Search_Coef.py and
"Estimate Covariance Matrix.py"); all arithmetic runs in sm_100a CUDA kernels behind the C ABI of
csrc/libpfml_b200.so (include/pfml_b200.h).  There is no CPU fallback.

The directory name required by the build contains hyphens, so import it through the root-level alias module
``pfml_b200`` (``import pfml_b200``) or ``importlib.import_module``.
"""
__version__ = "0.1.0"
