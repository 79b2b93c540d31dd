"""CPU, world_size 2 over gloo: the host logic of the multi-GPU path -- date partition, per-rank window buckets,
all-reduce of partial window sums -- reproduces the single-process sums (summation order only: <= 1e-12)."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import PKG, ROOT


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _worker(rank, world, port, out_path):
    import importlib, sys
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sh = importlib.import_module(PKG + ".sharding")
    sc = importlib.import_module(PKG + ".pfml_search_coef")
    syn = importlib.import_module(PKG + ".synthetic")
    rng = np.random.default_rng(0)                       # same data on every rank; each uses its own slice
    T, P = 47, 9
    month_ends = syn.month_end_range("2011-01", T)
    rt = rng.standard_normal((T, P)); dn = rng.standard_normal((T, P, P))
    years = [2013, 2014, 2015]
    mine = sh.shard_dates(np.arange(T), rank, world)
    bucket = sc._rank_buckets(month_ends[mine], sorted(years))
    nb = len(years) + 1
    r_sum = torch.zeros((nb, P), dtype=torch.float64); d_sum = torch.zeros((nb, P, P), dtype=torch.float64)
    cnt = torch.zeros(nb, dtype=torch.float64)
    for i, k in zip(mine, bucket):
        if k >= 0:
            r_sum[k] += torch.as_tensor(rt[i]); d_sum[k] += torch.as_tensor(dn[i]); cnt[k] += 1
    for t in (r_sum, d_sum, cnt):
        dist.all_reduce(t)
    if rank == 0:
        full, _ = sc.window_buckets(month_ends, years)
        ok = True
        for k in range(nb):
            ok &= bool(np.allclose(d_sum[k].numpy(), dn[full == k].sum(0), rtol=1e-12, atol=1e-13))
            ok &= bool(np.allclose(r_sum[k].numpy(), rt[full == k].sum(0), rtol=1e-12, atol=1e-13))
            ok &= int(cnt[k]) == int((full == k).sum())
        np.save(out_path, np.array([ok]))
    dist.destroy_process_group()


def test_two_rank_partial_sums(tmp_path):
    out = str(tmp_path / "ok.npy")
    mp.spawn(_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    assert bool(np.load(out)[0])


def test_shard_dates_partition():
    import importlib
    sh = importlib.import_module(PKG + ".sharding")
    months = np.arange(12, 372)
    for world in (1, 2, 3, 8):
        parts = [sh.shard_dates(months, r, world) for r in range(world)]
        assert (np.concatenate(parts) == months).all()
        assert max(len(p) for p in parts) - min(len(p) for p in parts) <= 1
    assert sh.panel_slice_for([20, 21, 22], 11) == (8, 22)
    assert len(sh.shard_dates([5], 1, 2)) + len(sh.shard_dates([5], 0, 2)) == 1
