"""e2e throughput of pfml_date_steps_host vs chunk size (N=500, p=1024, 148 dates)."""
import importlib, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
PKG = "jkmp22-machine-learning-and-the-implementable-efficient-frontier-replication_b200"
syn = importlib.import_module(PKG + ".synthetic"); host = importlib.import_module(PKG + ".host")
B = 148
panel = syn.make_panel(N=500, T=syn.LB + 2 + B, p=1024, seed=7)
months = [int(t) for t in panel.hot_months()]
blk = host.HostBlock(panel, months, pin=True)
for chunk in [int(c) for c in sys.argv[1:]] or [37]:
    for full in (True, False):
        blk.run(chunk=chunk, device=0, full_outputs=full)
        t0 = time.perf_counter()
        for _ in range(3):
            blk.run(chunk=chunk, device=0, full_outputs=full)
        dt = (time.perf_counter() - t0) / 3
        print(f"chunk {chunk:3d} full_outputs={full}: {B/dt:8.1f} date-steps/s  ({dt*1e3:.1f} ms/step)")
