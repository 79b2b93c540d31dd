"""Multi-GPU end-to-end check (run under torchrun, one rank per GPU): the sharded hp-grid pipeline over the library's own
NCCL communicator (pfml_comm_*: all-reduce of the window sums, ridge problems dealt over ranks, beta all-gathered) against
the single-GPU pipeline computed on rank 0.  Prints MULTI_GPU_OK on success."""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.distributed as dist
PKG = "jkmp22-machine-learning-and-the-implementable-efficient-frontier-replication_b200"
syn = importlib.import_module(PKG + ".synthetic"); pl = importlib.import_module(PKG + ".pfml_pipeline")
C = importlib.import_module(PKG + ".comm")
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
dev = f"cuda:{local}"
comm = C.Communicator.from_torch_distributed(device=dev)
panel = syn.make_panel(N=120, T=50, p=32, seed=71, ragged=0.03, start="2010-01")
hot = panel.hot_months()
rng = np.random.default_rng(5)
Ws = [panel.W, np.sqrt(np.exp(-2.0)) * rng.standard_normal(panel.W.shape)]
years, p_vec, l_vec = [2013, 2014], [8, 16, 32], [0.0, 1e-4, 1e-2, 1.0]
tm = {}
coef = pl.hp_grid_pipeline(panel, Ws, hot, years, p_vec, l_vec, comm=comm, device=dev, timing=tm)
# every rank must hold the same, complete grid
flat = np.concatenate([coef[g][y][p].ravel() for g in range(len(Ws)) for y in years for p in p_vec])
t = torch.as_tensor(flat).to(dev)
lo, hi = t.clone(), t.clone()
dist.all_reduce(lo, op=dist.ReduceOp.MIN); dist.all_reduce(hi, op=dist.ReduceOp.MAX)
same = bool(torch.equal(lo, hi))
if rank == 0:
    ref = pl.hp_grid_pipeline(panel, Ws, hot, years, p_vec, l_vec, device=dev)
    worst = 0.0
    for g in range(len(Ws)):
        for y in years:
            for p in p_vec:
                for j in range(1, len(l_vec)):
                    worst = max(worst, float(np.max(np.abs(coef[g][y][p][j] - ref[g][y][p][j])) / np.max(np.abs(ref[g][y][p][j]))))
    assert same, "ranks disagree on the gathered beta"
    assert worst < 1e-10, worst
    print(f"MULTI_GPU_OK world={world} nccl={importlib.import_module(PKG + '._lib').load().pfml_comm_nccl_version()} "
          f"allreduce_bytes={tm['allreduce_bytes']} max rel diff vs 1 GPU {worst:.2e} identical_on_all_ranks={same}")
dist.barrier()
comm.close()
dist.destroy_process_group()
