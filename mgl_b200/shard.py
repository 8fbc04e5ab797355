"""How a batch is spread over the GPUs of one box: pairs are independent, so
each rank takes a contiguous range and nothing is exchanged."""
from __future__ import annotations

from typing import Sequence, Tuple

import numpy as np


def weak_shard(pairs_per_gpu: int, rank: int) -> Tuple[int, int]:
    """bench.py's layout: every rank aligns its own `pairs_per_gpu` of the synthetic stream."""
    return rank * pairs_per_gpu, pairs_per_gpu


def strong_shard_by_cells(ref_len: Sequence[int], alt_len: Sequence[int], world: int) -> np.ndarray:
    """Cut [0, n) into `world` contiguous ranges of near-equal DP cells; returns world + 1 boundaries."""
    cells = np.asarray(ref_len, dtype=np.uint64) * np.asarray(alt_len, dtype=np.uint64)
    csum = np.concatenate([[0], np.cumsum(cells)])
    targets = csum[-1] * np.arange(1, world, dtype=np.float64) / world
    cuts = np.searchsorted(csum, targets, side="left")
    return np.concatenate([[0], cuts, [len(cells)]]).astype(np.int64)
