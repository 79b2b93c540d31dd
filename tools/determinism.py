"""Run-to-run determinism of the resident date-step path at the bench size: outputs must be bit-identical."""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
PKG = "jkmp22-machine-learning-and-the-implementable-efficient-frontier-replication_b200"
syn = importlib.import_module(PKG + ".synthetic"); dm = importlib.import_module(PKG + ".device")
B = int(sys.argv[1]) if len(sys.argv) > 1 else 148
panel = syn.make_panel(N=500, T=syn.LB + 2 + B, p=1024, seed=7, ragged=0.01)
months = [int(t) for t in panel.hot_months()]
dp = dm.DevicePanel(panel); sel, n_sel, n_max, NP = dm.universe(panel, months)
r = dm.DateStepRunner(dp, B, n_max); r.set_batch(months, sel, n_sel)
r.run(); torch.cuda.synchronize()
ref = {k: getattr(r, k).clone() for k in ("r_tilde", "denom", "risk", "tc", "signal_t")}
diff = 0
for it in range(6):
    r.run(); torch.cuda.synchronize()
    for k, v in ref.items():
        if not torch.equal(getattr(r, k), v):
            diff += 1
            print("run", it, k, "differs, max abs", float((getattr(r, k) - v).abs().max()))
print("status", int(r.status.abs().sum()), "bitwise differences over 6 repeats:", diff)
