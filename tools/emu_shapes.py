"""Time of the INT8-emulated GEMM kernel alone (profiling slot) and with slicing (events) for a few batched shapes."""
import ctypes as C, importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
PKG = "jkmp22-machine-learning-and-the-implementable-efficient-frontier-replication_b200"
lib = importlib.import_module(PKG + "._lib").load()
batch = 148
for (M, N, K) in ((512, 1025, 512), (512, 512, 512), (512, 1024, 512), (256, 256, 256), (512, 64, 512)):
    ldb = (N + 7) // 8 * 8
    A = torch.randn(batch, M, K, dtype=torch.float64, device="cuda")
    Bm = torch.randn(batch, K, ldb, dtype=torch.float64, device="cuda")
    D = torch.zeros(batch, M, ldb, dtype=torch.float64, device="cuda")
    st = torch.zeros(1, dtype=torch.int32, device="cuda")
    nb = lib.pfml_dgemm_emulated_workspace_bytes(M, N, K, batch)
    ws = torch.empty(nb, dtype=torch.uint8, device="cuda")
    def run():
        rc = lib.pfml_dgemm_emulated(M, N, K, 1.0, C.c_void_p(A.data_ptr()), K, M * K, C.c_void_p(Bm.data_ptr()), ldb, K * ldb, 0.0,
                                     C.c_void_p(D.data_ptr()), ldb, M * ldb, batch, C.c_void_p(st.data_ptr()), C.c_void_p(ws.data_ptr()), nb,
                                     C.c_void_p(torch.cuda.current_stream().cuda_stream))
        assert rc == 0
    for _ in range(2): run()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5): run()
    e1.record(); torch.cuda.synchronize()
    nt = lib.pfml_prof_ntags(); ms, fl, ln = (np.zeros(nt) for _ in range(3))
    lib.pfml_prof_enable(1); run()
    lib.pfml_prof_collect(C.c_void_p(ms.ctypes.data), C.c_void_p(fl.ctypes.data), C.c_void_p(ln.ctypes.data)); lib.pfml_prof_enable(0)
    k = [i for i in range(nt) if lib.pfml_prof_tag_name(i).decode() == "emu_i8_gemm"][0]
    tiles = batch * ((M + 127) // 128) * ((N + 63) // 64)
    print(f"M={M} N={N} K={K}: with slicing {e0.elapsed_time(e1)/5:.3f} ms, kernel {ms[k]:.3f} ms, {ms[k]*1e3/ (tiles/148):.1f} us/tile, "
          f"{2.0*batch*M*N*K/ms[k]*1e-9:.1f} TF-eq")
