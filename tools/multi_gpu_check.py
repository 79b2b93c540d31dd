"""Multi-GPU end-to-end check (run under torchrun, one rank per GPU, NCCL):
dates are sharded in contiguous blocks, every rank runs its own date-steps, the per-window partial sums are combined
by an all-reduce inside search_coef, and rank 0 compares beta with the single-GPU result (summation order only)."""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.distributed as dist
PKG = "jkmp22-machine-learning-and-the-implementable-efficient-frontier-replication_b200"
syn = importlib.import_module(PKG + ".synthetic"); dm = importlib.import_module(PKG + ".device")
sc = importlib.import_module(PKG + ".pfml_search_coef"); sh = importlib.import_module(PKG + ".sharding")
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
dev = f"cuda:{local}"
panel = syn.make_panel(N=120, T=48, p=32, seed=31, ragged=0.03, start="2010-01")
hot = panel.hot_months()
years, p_vec, l_vec = [2013, 2014], [8, 16, 32], [0.0, 1e-4, 1e-2, 1.0]
mine = sh.shard_dates(hot, rank, world)
out = dm.date_steps(panel, [int(t) for t in mine], device=dev) if len(mine) else []
P = panel.p + 1
rt = np.stack([o["r_tilde"] for o in out]) if out else np.zeros((0, P))
dn = np.stack([o["denom"] for o in out]) if out else np.zeros((0, P, P))
coef = sc.search_coef(rt, dn, panel.month_ends[mine], years, p_vec, l_vec, device=dev, group=True)
if rank == 0:
    full = dm.date_steps(panel, [int(t) for t in hot], device=dev)
    ref = sc.search_coef(np.stack([o["r_tilde"] for o in full]), np.stack([o["denom"] for o in full]), panel.month_ends[hot],
                         years, p_vec, l_vec, device=dev)
    worst = 0.0
    for y in years:
        for p in p_vec:
            for j in range(1, len(l_vec)):
                worst = max(worst, float(np.max(np.abs(coef[y][p][j] - ref[y][p][j])) / np.max(np.abs(ref[y][p][j]))))
    print(f"MULTI_GPU_OK world={world} dates/rank={[len(sh.shard_dates(hot, r, world)) for r in range(world)]} max rel diff {worst:.2e}")
    assert worst < 1e-10
dist.barrier()
dist.destroy_process_group()
