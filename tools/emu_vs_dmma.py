"""Date-step outputs with the INT8-emulated products against the DMMA products (same inputs), N=500 p=1024."""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
PKG = "jkmp22-machine-learning-and-the-implementable-efficient-frontier-replication_b200"
syn = importlib.import_module(PKG + ".synthetic"); dm = importlib.import_module(PKG + ".device"); lib = importlib.import_module(PKG + "._lib").load()
B = 8
panel = syn.make_panel(N=500, T=syn.LB + 2 + B, p=1024, seed=7, ragged=0.02)
months = [int(t) for t in panel.hot_months()]
dp = dm.DevicePanel(panel); sel, n_sel, n_max, NP = dm.universe(panel, months)
r = dm.DateStepRunner(dp, B, n_max); r.set_batch(months, sel, n_sel)
out = {}
for emu in (0, 1):
    lib.pfml_set_emulation(emu); r.run(); torch.cuda.synchronize()
    out[emu] = {k: getattr(r, k).cpu().numpy().copy() for k in ("r_tilde", "denom", "risk", "tc")}
lib.pfml_set_emulation(1)
print({k: float(np.abs(out[1][k] - out[0][k]).max() / np.abs(out[0][k]).max()) for k in out[0]})
