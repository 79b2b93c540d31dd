"""Date sharding over ranks: date-steps are independent, so each rank takes one contiguous
block of dates (contiguous blocks share look-back months, which keeps the per-rank panel slice small)."""
from __future__ import annotations

import numpy as np


def shard_dates(months, rank, world):
    """Contiguous, balanced partition of ``months`` (any sequence); returns this rank's slice (may be empty)."""
    months = np.asarray(list(months))
    bounds = np.linspace(0, len(months), world + 1).round().astype(int)
    return months[bounds[rank]:bounds[rank + 1]]


def panel_slice_for(months, lb):
    """First and last month index a rank needs resident: its dates plus lb+1 look-back months."""
    months = np.asarray(list(months))
    if len(months) == 0:
        return 0, -1
    return int(months.min()) - lb - 1, int(months.max())
