"""Stress universe (BASELINE configs[4]): N=3000, p=1024 -- a few date-steps, property checks and timing."""
import importlib, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
PKG = "jkmp22-machine-learning-and-the-implementable-efficient-frontier-replication_b200"
syn = importlib.import_module(PKG + ".synthetic"); dm = importlib.import_module(PKG + ".device")
B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
NS = int(sys.argv[2]) if len(sys.argv) > 2 else 3000
panel = syn.make_panel(N=NS, T=syn.LB + 2 + B, p=1024, seed=11, ragged=0.01)
months = [int(t) for t in panel.hot_months()]
dp = dm.DevicePanel(panel); sel, n_sel, n_max, NP = dm.universe(panel, months)
r = dm.DateStepRunner(dp, B, n_max); r.set_batch(months, sel, n_sel)
print("NP", NP, "workspace GB", r.ws_bytes / 1e9)
r.run(); torch.cuda.synchronize()
t0 = time.perf_counter(); r.run(); torch.cuda.synchronize(); dt = time.perf_counter() - t0
st = r.status.cpu().numpy(); d = r.denom[0].cpu().numpy(); ri = r.risk[0].cpu().numpy(); tc = r.tc[0].cpu().numpy()
ev = np.linalg.eigvalsh(d)
print(f"{B} date-steps in {dt*1e3:.1f} ms = {B/dt:.1f} date-steps/s; status {st.tolist()}; symmetric {bool((d == d.T).all())}; "
      f"risk+tc==denom {float(np.abs(ri + tc - d).max() / np.abs(d).max()):.1e}; min eig / max eig {ev.min() / ev.max():.2e}; finite {bool(np.isfinite(d).all())}")
lib = importlib.import_module(PKG + "._lib").load()
lib.pfml_set_emulation(0)
r.run(); torch.cuda.synchronize()
t0 = time.perf_counter(); r.run(); torch.cuda.synchronize(); dt0 = time.perf_counter() - t0
d0 = r.denom[0].cpu().numpy(); lib.pfml_set_emulation(1)
print(f"DMMA path: {B/dt0:.1f} date-steps/s; INT8-emulated vs DMMA denom: {float(np.abs(d - d0).max() / np.abs(d0).max()):.2e} relative")
