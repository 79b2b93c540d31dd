"""GPU (one device is enough): the sharded hp-grid pipeline -- date-steps with fused window sums, exchange, running sums,
ridge problems dealt over ranks, beta gathered -- against the unsharded path and the oracle.

The ranks of a 2-, 3- and 8-rank world run one after the other on the same GPU (``virtual_world``): same date -> rank map,
same kernels, same ridge ranges as the N-GPU run, only the NCCL transport is replaced by in-order adds.  The real NCCL
path is exercised with a single-rank communicator here and with torchrun on >= 2 GPUs in tests/test_multi_gpu.py.
Bound: N-rank vs 1-rank differ by FP64 summation order only (<= 1e-10 relative on beta for lambda > 0)."""
import numpy as np
import pytest

import pfml_oracle as O
from conftest import pkg, relerr

pytestmark = pytest.mark.gpu


def _setup(syn):
    panel = syn.make_panel(N=120, T=50, p=32, seed=71, ragged=0.03, start="2010-01")
    hot = panel.hot_months()
    rng = np.random.default_rng(5)
    Ws = [panel.W, np.sqrt(np.exp(-2.0)) * rng.standard_normal(panel.W.shape)]      # two bandwidths
    return panel, hot, Ws, [2013, 2014], [8, 16, 32], [0.0, 1e-4, 1e-2, 1.0]


def test_virtual_ranks_match_single_rank_and_oracle(syn):
    import dataclasses
    pl = pkg("pfml_pipeline")
    panel, hot, Ws, years, p_vec, l_vec = _setup(syn)
    one = pl.hp_grid_pipeline(panel, Ws, hot, years, p_vec, l_vec)
    for world in (2, 3, 8):
        got = pl.hp_grid_pipeline(panel, Ws, hot, years, p_vec, l_vec, virtual_world=world, batch=7)
        for g in range(len(Ws)):
            for y in years:
                for p in p_vec:
                    for j in range(1, len(l_vec)):
                        assert relerr(got[g][y][p][j], one[g][y][p][j]) < 1e-10, (world, g, y, p, j)
    # and the single-rank pipeline against the oracle's date-steps + search_coef, g by g
    for g, W in enumerate(Ws):
        pg = dataclasses.replace(panel, W=W)
        T, N, F = pg.feats.shape
        rff_x = O.rff(pg.feats.reshape(T * N, F), W).reshape(T, N, -1)
        vs = O.vol_scale_panel(pg)
        outs = [O.date_step(pg, int(t), rff_x, vs) for t in hot]
        ref = O.search_coef(np.stack([o["r_tilde"] for o in outs]), np.stack([o["denom"] for o in outs]), panel.month_ends[hot],
                            years, p_vec, l_vec)
        for y in years:
            for p in p_vec:
                for j in range(1, len(l_vec)):
                    assert relerr(one[g][y][p][j], ref[y][p][j]) < 1e-8, (g, y, p, j)


def test_single_rank_nccl_communicator(syn):
    """pfml_comm_init / allreduce_sums / ridge_grid_sharded through a real (one-rank) NCCL communicator."""
    import torch
    pl = pkg("pfml_pipeline")
    C = pkg("comm")
    assert pkg("_lib").load().pfml_comm_nccl_version() > 20000
    comm = C.Communicator(C.Communicator.unique_id(), 0, 1, torch.cuda.current_device())
    try:
        panel, hot, Ws, years, p_vec, l_vec = _setup(syn)
        a = pl.hp_grid_pipeline(panel, Ws[:1], hot, years, p_vec, l_vec, comm=comm)
        b = pl.hp_grid_pipeline(panel, Ws[:1], hot, years, p_vec, l_vec)
        for y in years:
            for p in p_vec:
                assert np.array_equal(a[0][y][p], b[0][y][p])
        with pytest.raises(ValueError):
            pl.hp_grid_pipeline(panel, Ws[:1], hot, [2030], p_vec, l_vec, comm=comm)
    finally:
        comm.close()
