"""Accuracy and speed of the INT8 modes (1 = 7 digits everywhere, 2 = 6 digits for the non-compounding products)
against the DMMA path, N=500 p=1024 (and N=3000 with argument 3000)."""
import importlib, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
PKG = "jkmp22-machine-learning-and-the-implementable-efficient-frontier-replication_b200"
syn = importlib.import_module(PKG + ".synthetic"); dm = importlib.import_module(PKG + ".device"); lib = importlib.import_module(PKG + "._lib").load()
N = int(sys.argv[1]) if len(sys.argv) > 1 else 500
B = 8 if N > 1000 else 16
panel = syn.make_panel(N=N, T=syn.LB + 2 + B, p=1024, seed=7, ragged=0.02)
months = [int(t) for t in panel.hot_months()]
dp = dm.DevicePanel(panel); sel, n_sel, n_max, NP = dm.universe(panel, months)
r = dm.DateStepRunner(dp, B, n_max); r.set_batch(months, sel, n_sel)
out = {}
for mode in (0, 1, 2):
    lib.pfml_set_emulation(mode); r.run(); torch.cuda.synchronize()
    t0 = time.perf_counter(); r.run(); torch.cuda.synchronize(); dt = time.perf_counter() - t0
    out[mode] = {k: getattr(r, k).cpu().numpy().copy() for k in ("r_tilde", "denom", "risk", "tc")}
    print(f"mode {mode}: {B/dt:.1f} date-steps/s, status {int(r.status.abs().sum())}")
lib.pfml_set_emulation(1)
for mode in (1, 2):
    print(mode, {k: float(np.abs(out[mode][k] - out[0][k]).max() / np.abs(out[0][k]).max()) for k in out[0]})
