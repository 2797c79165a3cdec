"""Pure-PyTorch oracle of the selection rules (fp32, any device).

Every function restates the *intended* semantics of one reference compressor
 with plain torch ops and host control flow.  It is the numerics reference of
the CUDA kernels and of the C++ CPU twin, and the fallback body of the public
``compression.compressors[...]`` classes for selectors that have no kernel.

All selectors return **ascending int64 indices** into the flattened input.
"""
from __future__ import annotations

import math
from typing import Tuple

import torch

from .chain import compute_k, gaussian_zabs


def moments(acc: torch.Tensor) -> Tuple[float, float]:
    """(mean, unbiased std) as the reference obtains them: fp32 tensors converted to python floats
    (compression.py:367-369)."""
    a = acc.double()
    n = a.numel()
    mean = a.sum() / n
    var = ((a * a).sum() - a.sum() * mean) / (n - 1) if n > 1 else torch.zeros((), dtype=torch.float64)
    var = torch.clamp(var, min=0)
    return float(mean.float()), float(var.sqrt().float())


def gaussian_threshold_candidates(mean: float, std: float, density: float, nloops: int):
    """All thresholds the <=nloops-step loop of compression.py:372-381 can evaluate, as fp32 values
    keyed by (halvings a, x1.5 steps b)."""
    t0 = mean + gaussian_zabs(density) * std
    out = {}
    for b in range(nloops):
        for a in range(nloops - b):
            t = t0
            for _ in range(b):
                t *= 1.5
            for _ in range(a):
                t *= 0.5
            out[(a, b)] = float(torch.tensor(t, dtype=torch.float64).float())
    return out


def gaussian_select(acc: torch.Tensor, density: float, nloops: int = 3, mean=None, std=None):
    """compression.py:358-389 (nloops=3) / :405-435 (nloops=5).  ``mean``/``std`` may be injected
    (e.g. the kernel's own moments) so that selection can be compared bit-exactly."""
    n = acc.numel()
    k = compute_k(n, density)
    if mean is None or std is None:
        mean, std = moments(acc)
    cands = gaussian_threshold_candidates(mean, std, density, nloops)
    absx = acc.abs()
    a = b = 0
    thr = cands[(0, 0)]
    mask = None
    for _ in range(nloops):
        thr = cands[(a, b)]
        mask = absx > thr
        cnt = int(mask.sum())
        if cnt < 2 * k / 3:
            a += 1
        elif cnt > 4 * k / 3:
            b += 1
        else:
            break
    idx = mask.nonzero().view(-1)
    return idx, thr


def topk_select(acc: torch.Tensor, k: int, smallest: bool = False) -> torch.Tensor:
    """Exact k largest |x| (compression.py:45-49), ties broken towards the lowest index, returned in
    ascending index order.  ``smallest`` gives min-k (compression.py:52-56)."""
    bits = acc.contiguous().view(torch.int32) & 0x7FFFFFFF
    key = (0x7FFFFFFF - bits) if smallest else bits
    key = key.to(torch.int64)
    n = key.numel()
    k = min(k, n)
    # composite key: magnitude first, then lower index wins
    comp = key * (2 ** 31) + (n - 1 - torch.arange(n, device=acc.device, dtype=torch.int64))
    idx = torch.topk(comp, k, largest=True, sorted=False).indices
    return idx.sort().values


def redsync_select(acc: torch.Tensor, density: float):
    """compression.py:644-677."""
    n = acc.numel()
    k = compute_k(n, density)
    absx = acc.abs()
    mean_val = absx.double().sum().div(n).float()
    max_val = absx.max()
    l, r = 0.0, 1.0
    idx = None
    thr = None
    while r - l > 0.2:
        tmp = l + (r - l) / 2
        thr = mean_val + torch.tensor(tmp, dtype=torch.float32, device=acc.device) * (max_val - mean_val)
        mask = absx > thr
        nnz = int(mask.sum())
        idx = mask.nonzero().view(-1)
        if nnz > k and 2 * k > nnz:
            break
        elif nnz < k / 2:
            r = tmp
        else:
            l = tmp
    return idx, float(thr)


def redsynctrim_select(acc: torch.Tensor, density: float, max_steps: int = 16):
    """compression.py:707-738 (the ratio ladder is cut after ``max_steps`` = GK_NC rungs)."""
    n = acc.numel()
    k = compute_k(n, density)
    absx = acc.abs()
    mean_val = absx.double().sum().div(n).float()
    max_val = absx.max()
    tmp = 1 - 0.2
    idx, thr = None, None
    for _ in range(max_steps):
        thr = mean_val + torch.tensor(tmp, dtype=torch.float32, device=acc.device) * (max_val - mean_val)
        mask = absx > thr
        idx = mask.nonzero().view(-1)
        if idx.numel() >= k:
            break
        tmp = tmp - 0.2
    return idx, float(thr)


def select(compressor: str, acc: torch.Tensor, density: float):
    k = compute_k(acc.numel(), density)
    if compressor in ("gaussian", "gaussiank"):
        return gaussian_select(acc, density, 3)[0]
    if compressor == "gaussian2":
        return gaussian_select(acc, density, 5)[0]
    if compressor in ("topk", "topk2", "dgcsampling"):
        return topk_select(acc, k)
    if compressor == "mink":
        return topk_select(acc, k, smallest=True)
    if compressor == "redsync":
        return redsync_select(acc, density)[0]
    if compressor == "redsynctrim":
        return redsynctrim_select(acc, density)[0]
    raise KeyError(compressor)


def compress(compressor: str, grad: torch.Tensor, resid: torch.Tensor, density: float,
             use_residual: bool = True, cap: int = None):
    """One error-feedback step: returns (acc, idx, values, new_resid).  Pairs beyond ``cap`` (in index
    order) are withheld and stay in the residual — the framework's one deliberate deviation
    (SURVEY.md §7.4-2)."""
    acc = grad + resid if use_residual else grad.clone()
    idx = select(compressor, acc, density)
    if cap is not None and idx.numel() > cap:
        idx = idx[:cap]
    vals = acc[idx]
    new_resid = acc.clone()
    new_resid[idx] = 0
    return acc, idx, vals, new_resid


def decompress_mean(pairs, n: int, nworkers: int) -> torch.Tensor:
    """Mean over workers of their sparse vectors — the intended semantics of
    distributed_optimizer.py:468-482 (coordinates picked by several workers are summed)."""
    device = pairs[0][0].device if pairs else "cpu"
    g = torch.zeros(n, dtype=torch.float32, device=device)
    scale = 1.0 / nworkers
    for idx, val in pairs:
        g.index_add_(0, idx.long(), val.float() * scale)
    return g


def sgd_reference(p, g, m, lr, momentum, dampening, wd, nesterov, first):
    """torch.optim.SGD single-tensor arithmetic (dl_trainer.py:848-872 restates it)."""
    if wd != 0:
        g = g.add(p, alpha=wd)
    if momentum != 0:
        if first:
            m = g.clone()
        else:
            m = m.mul(momentum).add(g, alpha=1 - dampening)
        u = g.add(m, alpha=momentum) if nesterov else m
    else:
        u = g
    return p.add(u, alpha=-lr), m


def lars_reference(w, g, acc, lr, momentum, wd, eeta, eps):
    """lars.py:83-132."""
    wn, gn = w.norm(2), g.norm(2)
    tr = torch.where(wn > 0, torch.where(gn > 0, eeta * wn / (gn + wd * wn + eps), torch.ones_like(wn)),
                     torch.ones_like(wn)).clamp(0.0, 50.0)
    slr = lr * tr
    dg = (g + wd * w).clamp(-10.0, 10.0)
    acc = acc * momentum + slr * dg
    return w - acc, acc
