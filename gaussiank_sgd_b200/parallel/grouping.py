"""Tensor-fusion planners: which gradients are sparsified / communicated together.

Three policies of the reference are provided (all operate on tensors in *backward* order, i.e. the
reverse of ``named_parameters()`` unless a profiled order is supplied):

* size threshold (distributed_optimizer.py:112-134) — the default; with the CLI default of
  524 288 000 elements the whole model is one group,
* MGS, merged-gradient sparsification (:230-310) — merge layer l into l-1 when the extra wait is
  smaller than the saved communication start-up,
* MG-WFBP for the dense path (:162-228).

The planners take the cost models as callables so that they can run either on the B200 models of
``utils.perf_model`` or on the reference's legacy tables.
"""
from __future__ import annotations

from typing import Callable, Dict, List, Sequence

from ..utils import perf_model as pm


def groups_by_threshold(keys_forward: Sequence[str], sizes: Dict[str, int], threshold: int):
    """Walk the tensors in reverse-forward order, closing a group as soon as its element count
    reaches ``threshold`` (threshold <= 0: one group per tensor)."""
    groups: List[List[str]] = []
    group_of: Dict[str, int] = {}
    cur: List[str] = []
    acc = 0
    for k in reversed(list(keys_forward)):
        acc += sizes[k]
        cur.append(k)
        group_of[k] = len(groups)
        if not acc < threshold:
            groups.append(cur)
            cur, acc = [], 0
    if cur:
        groups.append(cur)
    return groups, group_of


def _emit(seq: Sequence[str], cut_after: Sequence[bool]):
    groups, group_of, cur = [], {}, []
    for l in range(len(seq) - 1, -1, -1):
        cur.append(seq[l])
        group_of[seq[l]] = len(groups)
        if cut_after[l]:
            groups.append(cur)
            cur = []
    if cur:
        groups.append(cur)
    return groups, group_of


def groups_mgs(seq_layernames: Sequence[str], sizes: Dict[str, int], layerwise_times: Sequence[float],
               P: int, density: float,
               t_select: Callable[[float], float] = None,
               t_comm: Callable[[float, int, float], float] = None):
    """Merged-gradient sparsification planner.  Layers are indexed in forward order; backward visits
    L-1 .. 0.  Layer l joins layer l-1's group when waiting for l-1's gradient costs less than the
    start-up of a separate sparse exchange."""
    t_select = t_select or pm.compress_time
    t_comm = t_comm or pm.sparse_exchange_time
    L = len(seq_layernames)
    p = [float(sizes[k]) for k in seq_layernames]
    tb = [float(t) for t in layerwise_times]

    def schedule(tb_, p_, upto, start=0.0):
        taob, taos = [start] * upto, [start] * upto
        ts = [t_select(x) for x in p_[:upto]]
        if upto == 0:
            return taob, taos, ts
        taos[upto - 1] = taob[upto - 1] + tb_[upto - 1]
        for l in range(upto - 2, -1, -1):
            taob[l] = taos[l + 1] + ts[l + 1]
            taos[l] = taob[l] + tb_[l]
        return taob, taos, ts

    def comm_start(ts, taos, p_):
        tc = [t_comm(x, P, density) for x in p_]
        taoc = [0.0] * L
        taoc[L - 1] = taos[L - 1] + ts[L - 1]
        for l in range(L - 2, -1, -1):
            taoc[l] = max(taoc[l + 1] + tc[l + 1], taos[l] + ts[l])
        return taoc, tc

    taob, taos, ts = schedule(tb, p, L)
    taoc, tc = comm_start(ts, taos, p)
    cut = [False] * L          # cut[l]: a group boundary right after visiting layer l
    for l in range(L - 2, 0, -1):
        merged = p[l] + p[l - 1]
        wait = tb[l - 1] + t_select(merged) - t_select(p[l]) - t_select(p[l - 1]) - (taoc[l] - (taos[l] + ts[l]))
        saved = t_comm(p[l], P, density) + t_comm(p[l - 1], P, density) - t_comm(merged, P, density)
        if wait < saved:
            tb[l - 1] += tb[l]; tb[l] = 0.0
            p[l - 1] = merged; p[l] = 0.0
            tc[l - 1] = t_comm(merged, P, density); tc[l] = 0.0
            ts[l - 1] = t_select(merged); ts[l] = 0.0
            b2, s2, _ = schedule(tb, p, l, start=taob[l] + tb[l])
            taob[:l], taos[:l] = b2, s2
            taoc, tc = comm_start(ts, taos, p)
        else:
            cut[l] = True
    groups, group_of = _emit(seq_layernames, cut)
    info = {"non_overlapped": taoc[0] + tc[0] - (taos[0] + ts[0]) if L else 0.0, "compress_total": sum(ts)}
    return groups, group_of, info


def groups_mgwfbp(seq_layernames: Sequence[str], sizes: Dict[str, int], layerwise_times: Sequence[float],
                  P: int, t_allreduce: Callable[[float], float] = None, alpha: float = None):
    """MG-WFBP planner for the dense allreduce path."""
    if t_allreduce is None:
        t_allreduce = lambda nbytes: pm.dense_allreduce_time(nbytes, P)  # noqa: E731
    if alpha is None:
        alpha = t_allreduce(4.0)
    L = len(seq_layernames)
    p = [float(sizes[k]) for k in seq_layernames]
    tb = [float(t) for t in layerwise_times]
    tc = [t_allreduce(s * 4) for s in p]
    taob = [0.0] * L
    for l in range(L - 2, -1, -1):
        taob[l] = taob[l + 1] + tb[l + 1]

    def comm_start():
        taoc = [0.0] * L
        taoc[L - 1] = taob[L - 1] + tb[L - 1]
        for l in range(L - 2, -1, -1):
            taoc[l] = max(taoc[l + 1] + tc[l + 1], taob[l] + tb[l])
        return taoc

    taoc = comm_start()
    cut = [False] * L
    for l in range(L - 2, 0, -1):
        ready = taob[l - 1] + tb[l - 1]
        prev_done = taoc[l + 1] + tc[l + 1]
        merge = ready < prev_done or (prev_done < ready < taoc[l] + tc[l] and taoc[l] + alpha > ready)
        if merge:
            tc[l] = 0.0
            p[l - 1] += p[l]; p[l] = 0.0
            tc[l - 1] = t_allreduce(p[l - 1] * 4)
            taoc = comm_start()
        else:
            cut[l] = True
    groups, group_of = _emit(seq_layernames, cut)
    info = {"non_overlapped": taoc[0] + tc[0] - (taob[0] + tb[0]) if L else 0.0}
    return groups, group_of, info
