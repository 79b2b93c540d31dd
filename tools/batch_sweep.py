"""Resident-path time per date-step as a function of the batch size (N=500, p=1024)."""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
PKG = "jkmp22-machine-learning-and-the-implementable-efficient-frontier-replication_b200"
syn = importlib.import_module(PKG + ".synthetic"); dm = importlib.import_module(PKG + ".device")
for B in [int(a) for a in sys.argv[1:]] or [25, 74, 148, 296]:
    panel = syn.make_panel(N=500, T=syn.LB + 2 + B, p=1024, seed=7)
    months = [int(t) for t in panel.hot_months()]
    dp = dm.DevicePanel(panel); sel, n_sel, n_max, NP = dm.universe(panel, months)
    r = dm.DateStepRunner(dp, B, n_max); r.set_batch(months, sel, n_sel)
    for _ in range(2): r.run()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3): r.run()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 3
    print(f"B={B}: {ms:.1f} ms/step = {ms/B:.4f} ms/date = {B/ms*1e3:.0f} date-steps/s; workspace {r.ws_bytes/1e9:.1f} GB")
    del r, dp
    torch.cuda.empty_cache()
