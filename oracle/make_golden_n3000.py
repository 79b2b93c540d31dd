"""Oracle outputs at the C5 universe size: tests/golden/oracle_n3000_p64.npz (TEST INFRASTRUCTURE).

Usage:  python oracle/make_golden_n3000.py     (about 3 minutes of CPU: scipy sqrtm and 46 products at n = 3000)

One date-step of oracle/pfml_oracle.py (the numpy restatement, itself pinned to the reference by make_golden.py /
make_golden_mid.py) on the seeded synthetic panel make_panel(N=3000, T=14, p=64, seed=41) at month 13.  The fixture
holds the outputs only; tests/test_baseline_configs_gpu.py regenerates the identical panel from the seed on the GPU box
(same image, same numpy) and compares the CUDA path with these vectors, so the three-minute CPU run is not repeated there.
"""
from __future__ import annotations

import importlib
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
PKG = "jkmp22-machine-learning-and-the-implementable-efficient-frontier-replication_b200"
import pfml_oracle as O   # noqa: E402

N, T, P_FEAT, SEED, MONTH = 3000, 14, 64, 41, 13


def main():
    syn = importlib.import_module(PKG + ".synthetic")
    panel = syn.make_panel(N=N, T=T, p=P_FEAT, seed=SEED, ragged=0.01)
    t0 = time.time()
    o = O.date_step(panel, MONTH)
    print("oracle date-step at n = %d: %.1f s" % (len(o["ids"]), time.time() - t0))
    out = os.path.join(ROOT, "tests", "golden", "oracle_n3000_p64.npz")
    np.savez_compressed(out, r_tilde=o["r_tilde"], denom=o["denom"], risk=o["risk"], tc=o["tc"], signal_t_every8=o["signal_t"][::8],
                        ids=o["ids"], feats_checksum=np.array([panel.feats.sum(), panel.W.sum(), panel.lam.sum()]),
                        params=np.array([N, T, P_FEAT, SEED, MONTH]))
    print("wrote", out, os.path.getsize(out) // 1024, "KiB")


if __name__ == "__main__":
    main()
