"""Experimental, ordering-based selectors of the reference fork (compression.py:52-177).

These are research knobs that the reference swaps in by editing one line (compression.py:191).  They
need a full ordering of the tensor, are not on any benchmarked path, and are expressed here with
vectorised torch ops (``argsort``) — the per-bin Python loops of the reference are replaced by
cumulative-sum arithmetic.  Each returns ``(values, indexes)`` like the reference's static methods.
"""
from __future__ import annotations

import torch

UNIFORM_INTERVAL = 101  # compression.py:98


def mink(tensor: torch.Tensor, k: int):
    """k smallest |x| (compression.py:52-56)."""
    idx = torch.abs(tensor).argsort()[:k]
    return tensor[idx], idx


def uniform_topk(tensor: torch.Tensor, k: int):
    """Every 101st element of the ascending order, last k of those (compression.py:96-100)."""
    order = tensor.argsort()
    idx = order[::UNIFORM_INTERVAL][-k:]
    return tensor[idx], idx


def uniform_abs_topk(tensor: torch.Tensor, k: int):
    """What ``compressors['topk']`` actually runs at the reference's HEAD (compression.py:103-104,191;
    SURVEY.md Q1): uniform sampling of the |x| ordering.  Values are those of |x|, like the reference."""
    return uniform_topk(torch.abs(tensor), k)


def bucketized_topk(tensor: torch.Tensor, k: int):
    """compression.py:60-93: order by |x| descending, bucket by int(100*x), take round(count*k/n) from
    every bucket (singletons always), pad from the leftovers, trim to k.  The reference walks the
    buckets in ascending bucket-value order over the |x|-descending permutation; that literal behaviour
    is kept."""
    n = tensor.numel()
    order = torch.abs(tensor).argsort(descending=True)
    q = (tensor[order] * 100).int()
    _bins, counts = torch.unique(q, return_counts=True)
    take = torch.where(counts == 1, torch.ones_like(counts), torch.round(counts.double() * k / n).long())
    take = torch.minimum(take, counts)
    starts = torch.cumsum(counts, 0) - counts
    bin_of_pos = torch.repeat_interleave(torch.arange(counts.numel(), device=tensor.device), counts)
    pos_in_bin = torch.arange(n, device=tensor.device) - starts[bin_of_pos]
    chosen = pos_in_bin < take[bin_of_pos]
    idx, rest = order[chosen], order[~chosen]
    if k > idx.numel():
        idx = torch.cat([idx, rest[:k - idx.numel()]])
    idx = idx[:k]
    return tensor[idx], idx


def bellk(tensor: torch.Tensor, k: int):
    """Best-effort restatement of compression.py:107-153 (the original does float index arithmetic that
    fails under Python 3): keep the element(s) at the median of the value ordering and sample both
    halves uniformly outwards, (k-1)//2 per side."""
    order = tensor.argsort()
    n = order.numel()
    if n % 2 == 0:
        mids = order[n // 2 - 1: n // 2 + 1]
        lo, hi = order[: n // 2 - 1], order[n // 2 + 1:]
        side = max((k - 2) // 2, 0)
    else:
        mids = order[n // 2: n // 2 + 1]
        lo, hi = order[: n // 2], order[n // 2 + 1:]
        side = max((k - 1) // 2, 0)
    lo_pick = lo[::UNIFORM_INTERVAL][-side:] if side and lo.numel() else lo[:0]
    hi_pick = hi[::UNIFORM_INTERVAL][-side:] if side and hi.numel() else hi[:0]
    idx = torch.cat([lo_pick, mids, hi_pick])[:max(k, 1)]
    return tensor[idx], idx


def median_topk(tensor: torch.Tensor, k: int):
    """Best-effort restatement of compression.py:156-177 (its loop never advances): split the value
    ordering into k windows and keep the median element of each; the extremes are always kept."""
    order = tensor.argsort()
    n = order.numel()
    k = max(min(k, n), 1)
    edges = torch.linspace(0, n, k + 1, device=tensor.device).long()
    mid = ((edges[:-1] + edges[1:]) // 2).clamp(max=n - 1)
    idx = order[mid]
    if k >= 2:
        idx[0], idx[-1] = order[0], order[-1]
    idx = torch.unique(idx)
    return tensor[idx], idx


SELECTORS = {"mink": mink, "uniform_topk": uniform_topk, "uniform_abs_topk": uniform_abs_topk,
             "bucketized_topk": bucketized_topk, "bellk": bellk, "median_topk": median_topk}
