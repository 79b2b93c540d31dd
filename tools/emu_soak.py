"""Soak test of the INT8-emulated GEMM: integer-valued operands must give bit-exact products, many seeds and shapes."""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
PKG = "jkmp22-machine-learning-and-the-implementable-efficient-frontier-replication_b200"
dev = importlib.import_module(PKG + ".device")
bad = 0
n = int(sys.argv[1]) if len(sys.argv) > 1 else 40
for it in range(n):
    rng = np.random.default_rng(1000 + it)
    batch = int(rng.integers(1, 40)); M = int(rng.integers(1, 600)); N = int(rng.integers(1, 1100)); K = 64 * int(rng.integers(1, 9))
    A = rng.integers(-2000, 2000, (batch, M, K)).astype(np.float64)
    B = rng.integers(-2000, 2000, (batch, K, N)).astype(np.float64)
    D, st = dev.dgemm_emulated(A, B)
    ok = st == 0 and np.array_equal(D, A @ B)
    if not ok:
        bad += 1
        print("MISMATCH", it, batch, M, N, K, st, float(np.abs(D - A @ B).max()))
print(f"{n} random shapes, {bad} mismatches")
