"""Resident-path throughput: one batch of 148 on one stream vs two / four sub-batches on concurrent streams."""
import importlib, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
PKG = "jkmp22-machine-learning-and-the-implementable-efficient-frontier-replication_b200"
syn = importlib.import_module(PKG + ".synthetic"); dm = importlib.import_module(PKG + ".device")
B = 148
panel = syn.make_panel(N=500, T=syn.LB + 2 + B, p=1024, seed=7)
months = [int(t) for t in panel.hot_months()]
dp = dm.DevicePanel(panel); sel, n_sel, n_max, NP = dm.universe(panel, months)
for parts in (1, 2, 4):
    runners, streams = [], []
    bounds = np.linspace(0, B, parts + 1).astype(int)
    for k in range(parts):
        a, b = bounds[k], bounds[k + 1]
        r = dm.DateStepRunner(dp, b - a, n_max); r.set_batch(months[a:b], sel[a:b], n_sel[a:b])
        runners.append(r); streams.append(torch.cuda.Stream())
    def step():
        for r, s in zip(runners, streams):
            with torch.cuda.stream(s):
                r.run()
    for _ in range(2): step()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(3): step()
    torch.cuda.synchronize(); dt = (time.perf_counter() - t0) / 3
    print(f"{parts} concurrent sub-batches: {dt*1e3:.1f} ms per 148 dates = {B/dt:.0f} date-steps/s")
    del runners
