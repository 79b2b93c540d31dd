"""Per-tag device time of one resident step (library profiling hooks)."""
import ctypes as C, importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
PKG = "jkmp22-machine-learning-and-the-implementable-efficient-frontier-replication_b200"
syn = importlib.import_module(PKG + ".synthetic"); dm = importlib.import_module(PKG + ".device"); lib = importlib.import_module(PKG + "._lib").load()
B = int(sys.argv[1]) if len(sys.argv) > 1 else 148
panel = syn.make_panel(N=500, T=syn.LB + 2 + B, p=1024, seed=7)
months = [int(t) for t in panel.hot_months()]
dp = dm.DevicePanel(panel); sel, n_sel, n_max, NP = dm.universe(panel, months)
r = dm.DateStepRunner(dp, B, n_max); r.set_batch(months, sel, n_sel)
for _ in range(2): r.run()
torch.cuda.synchronize()
nt = lib.pfml_prof_ntags(); ms, fl, ln = (np.zeros(nt) for _ in range(3))
lib.pfml_prof_enable(1); r.run()
lib.pfml_prof_collect(C.c_void_p(ms.ctypes.data), C.c_void_p(fl.ctypes.data), C.c_void_p(ln.ctypes.data)); lib.pfml_prof_enable(0)
for i in range(nt):
    if ln[i]: print(f"{lib.pfml_prof_tag_name(i).decode():14s} {ms[i]:8.2f} ms  launches {int(ln[i]):4d}  {fl[i]/max(ms[i],1e-9)*1e-9:7.1f} TFLOP/s")
