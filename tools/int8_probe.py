"""tcgen05.mma kind::i8 issue-rate probe over instruction shapes (pfml_int8_mma_probe)."""
import ctypes as C, importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
lib = importlib.import_module("jkmp22-machine-learning-and-the-implementable-efficient-frontier-replication_b200._lib").load()
torch.cuda.init(); torch.zeros(1, device="cuda")
for ts in (0, 1):
    for n in (64, 128, 256):
        for nacc in (1, 2, 4, 8):
            if nacc * n > 512 - 16 * ts:
                continue
            t, c = C.c_double(), C.c_double()
            rc = lib.pfml_int8_mma_probe(n, 8192, ts | (nacc << 8), C.byref(t), C.byref(c))
            print("M=128 N=%3d K=32 %s accumulators=%d: rc=%d %.0f TOPS, %.1f nominal clk/MMA (ideal %.0f)" % (n, "TS" if ts else "SS", nacc, rc, t.value, c.value, 128 * n * 32 * 2 / 16384.0), flush=True)

for mode in (0, 1):
    c, t = C.c_double(), C.c_double()
    rc = lib.pfml_int8_pipe_probe(4096, mode, C.byref(c), C.byref(t))
    print("k-slab stream mode %d (%s): rc=%d %.0f nominal clk per slab, %.0f TOPS" % (mode, "cp mix 18 TS + 10 SS" if mode == 0 else "tcgen05.st stagers 22 TS + 6 SS", rc, c.value, t.value), flush=True)
    if rc: print(lib.pfml_last_error())
