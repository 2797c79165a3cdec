"""Per-coordinate "which ranks selected me" bitmaps (the idea sketched in the reference's test/tensor_test.py:
1-26: set / test bit ``rank`` of a uint32 flag word).  Used by the analysis tools to measure how much the
workers' selections overlap."""
import numpy as np


def set_rank_bit(flags: np.ndarray, indexes: np.ndarray, rank: int) -> np.ndarray:
    """flags[indexes] |= 1 << rank (in place); ranks 0..31."""
    assert 0 <= rank < 32 and flags.dtype == np.uint32
    np.bitwise_or.at(flags, indexes, np.uint32(1 << rank))
    return flags


def has_rank_bit(flags: np.ndarray, rank: int) -> np.ndarray:
    return (flags >> np.uint32(rank)) & np.uint32(1) == 1


def selection_overlap(index_lists, n: int):
    """Histogram h[c] = number of coordinates selected by exactly c workers (c = 0..P)."""
    flags = np.zeros(n, dtype=np.uint32)
    for r, idx in enumerate(index_lists):
        set_rank_bit(flags, np.asarray(idx, dtype=np.int64), r)
    counts = np.zeros(n, dtype=np.int64)
    for r in range(len(index_lists)):
        counts += has_rank_bit(flags, r)
    return np.bincount(counts, minlength=len(index_lists) + 1)
