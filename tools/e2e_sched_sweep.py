"""Chunk-schedule sweep of pfml_date_steps_host_packed at the bench size (PFML_HOST_SCHED is read per call)."""
import importlib, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
PKG = "jkmp22-machine-learning-and-the-implementable-efficient-frontier-replication_b200"
import torch
syn = importlib.import_module(PKG + ".synthetic"); host = importlib.import_module(PKG + ".host")
B = 148
panel = syn.make_panel(N=500, T=syn.LB + 2 + B, p=1024, seed=7, g=float(np.exp(-3.0)))
months = [int(t) for t in panel.hot_months()]
blk = host.HostBlock(panel, months, pin=True)
def timeit(f, n=3):
    f(); f()
    t0 = time.perf_counter()
    for _ in range(n): f()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / n
for sched in sys.argv[1:]:
    if sched == "default":
        os.environ.pop("PFML_HOST_SCHED", None)
    else:
        os.environ["PFML_HOST_SCHED"] = sched
    t = timeit(lambda: blk.run_packed())
    print("packed sched %-24s %.1f ms/step  %.0f date-steps/s" % (sched, t * 1e3, B / t), flush=True)
os.environ.pop("PFML_HOST_SCHED", None)
t = timeit(lambda: blk.run()); print("full squares (download) %.1f ms/step  %.0f date-steps/s" % (t * 1e3, B / t), flush=True)
