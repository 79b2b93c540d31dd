"""emu_gemm_kernel alone (library profiling slot emu_i8_gemm, slicing excluded) for growing depth K at fixed M x N:
if the time per tile is a fixed overhead plus the k-slab stream, the rate rises with K."""
import ctypes as C, importlib, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
PKG = "jkmp22-machine-learning-and-the-implementable-efficient-frontier-replication_b200"
_L = importlib.import_module(PKG + "._lib")
if os.environ.get("PFML_LIB"): _L.LIB_PATH = os.environ["PFML_LIB"]
lib = _L.load()
dev = torch.device("cuda:0")
nt = lib.pfml_prof_ntags()
names = [lib.pfml_prof_tag_name(i).decode() for i in range(nt)]
k_emu = names.index("emu_i8_gemm")
for (M, N, K, batch) in ((512, 1025, 512, 148), (512, 1025, 1024, 148), (512, 1025, 2048, 74), (512, 1025, 4096, 37), (512, 512, 512, 148), (1025, 1025, 512, 148)):
    A = torch.randn(batch, M, K, dtype=torch.float64, device=dev)
    B = torch.randn(batch, K, N, dtype=torch.float64, device=dev)
    D = torch.empty(batch, M, N, dtype=torch.float64, device=dev)
    wsb = lib.pfml_dgemm_emulated_workspace_bytes(M, N, K, batch)
    ws = torch.empty(wsb, dtype=torch.uint8, device=dev)
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    p = lambda t: C.c_void_p(t.data_ptr())
    def run():
        return lib.pfml_dgemm_emulated(M, N, K, 1.0, p(A), K, M * K, p(B), N, K * N, 0.0, p(D), N, M * N, batch, None, p(ws), wsb, st)
    for _ in range(2): assert run() == 0
    torch.cuda.synchronize()
    ms, fl, ln = (np.zeros(nt) for _ in range(3))
    lib.pfml_prof_enable(1)
    for _ in range(5): run()
    lib.pfml_prof_collect(C.c_void_p(ms.ctypes.data), C.c_void_p(fl.ctypes.data), C.c_void_p(ln.ctypes.data))
    lib.pfml_prof_enable(0)
    t = ms[k_emu] / 5
    MB, NB = (M + 127) // 128, (N + 63) // 64
    tiles = batch * MB * NB
    clk_tile = t * 1e-3 * 1.965e9 / (tiles / 148.0)
    print("M=%d N=%d K=%d batch=%d: %.3f ms, %.1f TFLOP/s-eq executed(padded) %.1f; %.0f clk per tile, %.0f clk per k-slab" % (
        M, N, K, batch, t, 2.0 * M * N * K * batch / t * 1e-9, 2.0 * MB * 128 * NB * 64 * K * batch / t * 1e-9, clk_tile, clk_tile / (K / 32)), flush=True)
    del A, B, D, ws
