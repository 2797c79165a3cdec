"""Small helpers + re-exports of the cost models (the reference's utils.py:1-139)."""
from __future__ import annotations

import os

import numpy as np

from .perf_model import (compress_time, dense_allreduce_time, get_approximate_sigma_scale,  # noqa: F401
                         legacy_allgather_perf_model as allgather_perf_model,
                         legacy_allreduce_time, legacy_topk_perf_model as topk_perf_model,
                         predict_density_with_size_and_computation, sparse_exchange_time)


def create_path(relative_path: str) -> None:
    """mkdir -p (utils.py:13)."""
    os.makedirs(relative_path, exist_ok=True)


def topk(tensor: np.ndarray, k: int):
    """numpy top-k by magnitude (utils.py:38-40): (indexes, values)."""
    indexes = np.abs(tensor).argsort()[-k:]
    return indexes, tensor[indexes]


def force_insert_item(d: dict, key, val) -> None:
    d.setdefault(key, []).append(val)


def predict_allreduce_time_with_size(alpha, beta, size, P):
    return legacy_allreduce_time(alpha, beta, size)


def gen_threshold_from_normal_distribution(p_value, mu, sigma):
    """(left, right) thresholds enclosing probability mass ``p_value`` of N(mu, sigma^2) (utils.py:136-138)."""
    from scipy import stats
    zvalue = stats.norm.ppf((1 - p_value) / 2)
    return mu + zvalue * sigma, mu - zvalue * sigma
