"""CPU: host logic of the sharded hp-grid pipeline that needs no GPU -- the ridge problem ranges (pfml_ridge_shard_range,
pure host code of the C-ABI library) and the panel slice a rank keeps resident."""
import numpy as np

from conftest import pkg


def test_ridge_shard_ranges_partition_and_balance():
    comm = pkg("comm")
    for nprob in (1, 7, 15, 420, 435):
        for world in (1, 2, 3, 4, 8, 16):
            ranges = [comm.shard_range(nprob, r, world) for r in range(world)]
            assert ranges[0][0] == 0 and ranges[-1][1] == nprob
            assert all(a[1] == b[0] for a, b in zip(ranges, ranges[1:]))            # contiguous, disjoint, complete
            sizes = [hi - lo for lo, hi in ranges]
            assert max(sizes) - min(sizes) <= 1                                       # balanced to one problem


def test_slice_panel_keeps_lookback(syn):
    pl = pkg("pfml_pipeline")
    sh = pkg("sharding")
    panel = syn.make_panel(N=12, T=40, p=4, seed=1)
    hot = panel.hot_months()
    mine = sh.shard_dates(hot, 1, 3)
    t0, t1 = sh.panel_slice_for(mine, panel.lb)
    sub = pl.slice_panel(panel, t0, t1, W=panel.W * 2.0)
    assert sub.T == t1 - t0 + 1 and sub.N == panel.N
    assert (sub.feats == panel.feats[t0:t1 + 1]).all() and (sub.month_ends == panel.month_ends[t0:t1 + 1]).all()
    assert (sub.W == panel.W * 2.0).all() and (sub.ids == panel.ids).all()
    assert int(mine.min()) - t0 == panel.lb + 1
