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


def pairhmm_shard_reads(lengths: np.ndarray, reads: np.ndarray, world: int, rank: int):
    """PairHMM over several GPUs: every (read, haplotype) pair is independent, so rank r takes a contiguous range of
    READS cut by read bases (its cells are read bases x the sum of the haplotype lengths, which every rank shares)
    and all haplotypes; nothing is exchanged.  Returns (first_read, sub_lengths, sub_reads): the length and read
    buffers of the rank's own computeLikelihoodsNative-shaped job, whose output rows are rows
    [first_read, first_read + sub_lengths[0]) of the whole job's."""
    R = int(lengths[0])
    rl = np.asarray(lengths[1:1 + R], dtype=np.int64)
    csum = np.concatenate([[0], np.cumsum(rl)])
    targets = csum[-1] * np.arange(1, world, dtype=np.float64) / world
    cuts = np.concatenate([[0], np.searchsorted(csum, targets, side="left"), [R]]).astype(np.int64)
    lo, hi = int(cuts[rank]), int(cuts[rank + 1])
    sub_lengths = np.concatenate([[hi - lo], lengths[1 + lo:1 + hi], lengths[1 + R:]]).astype(np.int32)
    return lo, sub_lengths, reads[5 * int(csum[lo]): 5 * int(csum[hi])]
