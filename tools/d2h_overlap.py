"""Does a pinned D2H copy keep its bandwidth while the date-step kernels run?"""
import importlib, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
PKG = "jkmp22-machine-learning-and-the-implementable-efficient-frontier-replication_b200"
syn = importlib.import_module(PKG + ".synthetic"); dm = importlib.import_module(PKG + ".device")
B = 74
panel = syn.make_panel(N=500, T=syn.LB + 2 + B, p=1024, seed=7)
months = [int(t) for t in panel.hot_months()]
dp = dm.DevicePanel(panel); sel, n_sel, n_max, NP = dm.universe(panel, months)
r = dm.DateStepRunner(dp, B, n_max); r.set_batch(months, sel, n_sel)
r.run(); torch.cuda.synchronize()
n = 2 * 1024**3 // 8
d = torch.empty(n, dtype=torch.float64, device="cuda"); h = torch.empty(n, dtype=torch.float64).pin_memory()
cp = torch.cuda.Stream()
def t_compute():
    torch.cuda.synchronize(); t0 = time.perf_counter(); r.run(); torch.cuda.synchronize(); return time.perf_counter() - t0
def t_copy():
    torch.cuda.synchronize(); t0 = time.perf_counter()
    with torch.cuda.stream(cp): h.copy_(d, non_blocking=True)
    torch.cuda.synchronize(); return time.perf_counter() - t0
def t_both():
    torch.cuda.synchronize(); t0 = time.perf_counter()
    with torch.cuda.stream(cp): h.copy_(d, non_blocking=True)
    r.run(); torch.cuda.synchronize(); return time.perf_counter() - t0
for _ in range(2): t_compute(); t_copy()
print(f"compute alone {t_compute()*1e3:.1f} ms; 2 GiB D2H alone {t_copy()*1e3:.1f} ms; both concurrently {t_both()*1e3:.1f} ms")
