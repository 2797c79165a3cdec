"""Compressor registry with the reference's calling convention (compression.py:12-754).

``compressors[name].compress(tensor, name, sigma_scale, ratio) -> (tensor, indexes, values)``:
``tensor`` is modified in place to ``grad + residual``, the class-level ``residuals[name]`` is replaced
by the un-sent part, ``indexes`` are int64 positions in the flattened tensor, ``k = max(int(n*ratio), 1)``.
``clear()`` drops the per-name state; ``decompress`` is the identity for the sparse compressors.

This module is the *compatibility* surface for code that drives compressors tensor by tensor.  The
training hot path does not go through it: :mod:`distributed_optimizer` runs the same selection rules
as a fused kernel chain over the gradient arena.  On CUDA tensors ``compress`` itself uses that chain
(one cached :class:`ops.chain.CompressChain` per tensor name); on CPU tensors it uses the C++ twin.

Differences from the reference's HEAD, on purpose: ``topk`` is exact top-k
(HEAD calls ``uniform_abs_topk``, :191 — available here as its own key), ``gaussiank`` is an alias of
``gaussian``, the experimental selectors are separate registry keys, ``randomksame[ec]`` are registered.
"""
from __future__ import annotations

from typing import Dict, Tuple

import torch

from .ops import chain as ch
from .ops import oracle, selectors
from .ops.chain import compute_k


class NoneCompressor:
    """Identity (compression.py:12-20); with >1 worker the optimizer allreduce-averages densely."""
    name = 'none'
    key = 'none'

    @staticmethod
    def compress(tensor, name=None, sigma_scale=None, ratio=None):
        return tensor, tensor.dtype

    @staticmethod
    def decompress(tensor, ctc=None, name=None):
        return tensor

    @staticmethod
    def clear():
        pass


class _SparseCompressor:
    """Shared machinery: class-level per-name state like the reference (compression.py:27-43)."""
    residuals: Dict[str, torch.Tensor] = {}
    values: Dict[str, torch.Tensor] = {}
    indexes: Dict[str, torch.Tensor] = {}
    sparsities: list = []
    zero_conditions: dict = {}
    name = 'sparse'
    key = 'sparse'
    error_feedback = True     # add the residual back before selecting
    kernel = None             # ops.chain mode name when a fused kernel exists
    _chains: Dict[Tuple, tuple] = {}

    @classmethod
    def clear(cls):
        cls.residuals = {}
        cls.values = {}
        cls.indexes = {}
        cls.sparsities = []
        cls.zero_conditions = {}
        for k in [k for k in _SparseCompressor._chains if k[0] is cls]:
            del _SparseCompressor._chains[k]

    @classmethod
    def _select(cls, acc: torch.Tensor, k: int, ratio: float) -> torch.Tensor:
        raise NotImplementedError

    @classmethod
    def _kernel_compress(cls, flat: torch.Tensor, name, ratio):
        """Run the native chain in place on (flat, residuals[name]); returns (idx, values)."""
        key = (cls, name, flat.numel(), flat.device, flat.data_ptr())
        ent = _SparseCompressor._chains.get(key)
        resid = cls.residuals[name]
        if ent is None or ent[0].resid.data_ptr() != resid.data_ptr():
            n = flat.numel()
            chain = ch.CompressChain(flat, resid.view(-1), [[(0, n)]], compressor=cls.kernel, density=ratio,
                                     use_residual=cls.error_feedback, zero_grad=False, cap=n)
            slot = ch.SlotBuffer(1, chain.cap, flat.device)
            ent = (chain, slot, ch.make_dst([slot.slot_ptr(0)], 0))
            _SparseCompressor._chains[key] = ent
        chain, slot, dst = ent
        chain.set_density(ratio)
        chain.compress(dst)
        cnt = int(slot.t[0].item()) & 0xFFFFFFFF
        words = slot.t[ch.N.GK_SLOT_HDR: ch.N.GK_SLOT_HDR + cnt]
        idx = words & 0xFFFFFFFF
        vals = (words >> 32).to(torch.int32).view(torch.float32)
        return idx, vals

    @classmethod
    def compress(cls, tensor, name=None, sigma_scale=3, ratio=0.05):
        with torch.no_grad():
            if name not in cls.residuals or cls.residuals[name].shape != tensor.shape:
                cls.residuals[name] = torch.zeros_like(tensor.data)
            flat = tensor.data.view(-1) if tensor.is_contiguous() else None
            usable = (flat is not None and cls.kernel is not None and flat.dtype == torch.float32
                      and flat.data_ptr() % 16 == 0 and cls.residuals[name].is_contiguous()
                      and cls.residuals[name].data_ptr() % 16 == 0 and 0 < ratio < 1)
            if usable:
                idx, vals = cls._kernel_compress(flat, name, ratio)
                # the chain left the new residual in residuals[name]; the caller's tensor becomes acc
                flat.copy_(cls.residuals[name].view(-1))
                flat[idx] = vals
            else:
                numel = tensor.numel()
                k = compute_k(numel, ratio)
                if cls.error_feedback:
                    tensor.add_(cls.residuals[name].data)
                acc = tensor.data.reshape(-1)
                idx = cls._select(acc, k, ratio)
                vals = acc[idx]
                res = tensor.data.clone()
                res.view(-1)[idx] = 0.0
                cls.residuals[name] = res
            cls.values[name] = vals
            cls.indexes[name] = idx
            return tensor, idx, vals

    @classmethod
    def get_residuals(cls, name, like_tensor):
        if name not in cls.residuals:
            cls.residuals[name] = torch.zeros_like(like_tensor.data)
        return cls.residuals[name]

    @classmethod
    def add_residuals(cls, included_indexes, name):
        """Put back into the residual the values a *global* second-stage selection rejected
        (gTop-k leftovers; compression.py:207-219): ``included_indexes`` index into values[name]."""
        with torch.no_grad():
            residuals = cls.residuals[name]
            if not torch.is_tensor(included_indexes):
                included_indexes = torch.as_tensor(included_indexes)
            inc = included_indexes.to(device=residuals.device).long()
            values = cls.values[name].clone()
            values[inc] = 0.0
            residuals.view(-1)[cls.indexes[name]] += values

    @staticmethod
    def decompress(tensor, ctc=None, name=None):
        return tensor


def _make(key_, selector=None, ef=True, kernel=None, doc=""):
    body = {"name": key_, "key": key_, "error_feedback": ef, "kernel": kernel, "__doc__": doc,
            "residuals": {}, "values": {}, "indexes": {}, "sparsities": [], "zero_conditions": {}}
    if selector is not None:
        body["_select"] = classmethod(lambda cls, acc, k, ratio, _s=selector: _s(acc, k, ratio))
    return type(key_.capitalize() + "Compressor", (_SparseCompressor,), body)


def _randperm_select(acc, k, ratio):
    return torch.randperm(acc.numel(), device=acc.device)[:k]


class _SameSeed:
    counter = 0


def _randperm_same_select(acc, k, ratio):
    # all ranks draw the same coordinates (compression.py:526-527)
    gen = torch.Generator(device="cpu").manual_seed(_SameSeed.counter)
    _SameSeed.counter += 1
    return torch.randperm(acc.numel(), generator=gen)[:k].to(acc.device)


def _dgc_select(acc, k, ratio):
    """compression.py:576-606: threshold from a 1 % sample, then trim to exactly k when more than 4k/3
    pass it (which makes it an exact top-k whenever the sample threshold is loose enough)."""
    n = acc.numel()
    absx = acc.abs()
    fk = int(n * 0.01)
    if fk < k:
        return oracle.topk_select(acc, k)
    sample = absx[torch.randperm(n, device=acc.device)[:fk]]
    thres = torch.topk(sample, k=k).values[k - 1]
    idx = (absx > thres).nonzero().view(-1)
    if idx.numel() > 4 * k / 3:
        idx = oracle.topk_select(acc, k)
    return idx


TopKCompressor = _make("topk", lambda a, k, r: oracle.topk_select(a, k), True, "topk",
                       "Exact top-k with error feedback (Aji & Heafield 2017; compression.py:23-224, selector :45-49).")
TopKCompressor2 = _make("topk2", lambda a, k, r: oracle.topk_select(a, k), False, "topk",
                        "Top-k without adding the residual back (compression.py:315-334).")
GaussianCompressor = _make("gaussian", lambda a, k, r: oracle.gaussian_select(a, r, 3)[0], True, "gaussian",
                           "Gaussian-k: threshold from a fitted normal + <=3 refinements (compression.py:337-403).")
GaussianCompressor2 = _make("gaussian2", lambda a, k, r: oracle.gaussian_select(a, r, 5)[0], False, "gaussian2",
                            "Gaussian-k with <=5 refinements and no add-back (compression.py:405-435).")
RandomKCompressor = _make("randomk", _randperm_select, False, None, "k random coordinates, no add-back (:439-489).")
RandomKECCompressor = _make("randomkec", _randperm_select, True, None, "random-k with error feedback (:492-509).")
RandomKSameCompressor = _make("randomksame", _randperm_same_select, False, None,
                              "random-k, same coordinates on every rank (:512-533).")
RandomKSameECCompressor = _make("randomksameec", _randperm_same_select, True, None,
                                "random-k same coordinates + error feedback (:536-553).")
DGCSamplingCompressor = _make("dgcsampling", _dgc_select, True, None, "DGC-style sampled threshold (:555-620).")
RedSyncCompressor = _make("redsync", lambda a, k, r: oracle.redsync_select(a, r)[0], True, "redsync",
                          "RedSync bisection between mean|x| and max|x| (:623-691).")
RedSyncTrimCompressor = _make("redsynctrim", lambda a, k, r: oracle.redsynctrim_select(a, r)[0], True,
                              "redsynctrim", "RedSync trimmed search (:694-738).")
MinKCompressor = _make("mink", lambda a, k, r: oracle.topk_select(a, k, smallest=True), True, "mink",
                       "k smallest |x| (selector compression.py:52-56).")
UniformTopKCompressor = _make("uniform_topk", lambda a, k, r: selectors.uniform_topk(a, k)[1], True)
UniformAbsTopKCompressor = _make("uniform_abs_topk", lambda a, k, r: selectors.uniform_abs_topk(a, k)[1], True,
                                 doc="The selector the reference's 'topk' key runs at HEAD (:103-104, :191).")
BucketizedTopKCompressor = _make("bucketized_topk", lambda a, k, r: selectors.bucketized_topk(a, k)[1], True)
BellKCompressor = _make("bellk", lambda a, k, r: selectors.bellk(a, k)[1], True)
MedianTopKCompressor = _make("median_topk", lambda a, k, r: selectors.median_topk(a, k)[1], True)


class BucketCompressor:
    """Sign-bucket mean compressor of the fork (compression.py:227-312): not a sparsifier.  Each worker
    subtracts the mean of its non-negative and of its negative entries, the two means are allreduce-
    averaged, and the averaged means are added back."""
    name = 'bucket'
    key = 'bucket'
    buckets = []
    means = []
    residuals = []

    @staticmethod
    def clear():
        BucketCompressor.buckets = []
        BucketCompressor.means = []
        BucketCompressor.residuals = []

    @staticmethod
    def bucket_mean_2(tensor, l=1):
        pos = tensor >= 0
        means = torch.zeros(2, dtype=tensor.dtype, device=tensor.device)
        if bool(pos.any()):
            means[0] = tensor[pos].mean()
        if bool((~pos).any()):
            means[1] = tensor[~pos].mean()
        tensor.sub_(torch.where(pos, means[0], means[1]))
        BucketCompressor.buckets = [pos, ~pos]
        BucketCompressor.means = means
        BucketCompressor.residuals = tensor
        return means

    @staticmethod
    def compress(tensor, name=None, sigma_scale=None, ratio=None):
        with torch.no_grad():
            means = BucketCompressor.bucket_mean_2(tensor.data)
            return tensor, None, means

    @staticmethod
    def decompress(tensor, ctc=None, name=None):
        """``tensor`` = the allreduce-averaged bucket means."""
        res = BucketCompressor.residuals
        pos = BucketCompressor.buckets[0]
        res.add_(torch.where(pos, tensor[0], tensor[1]))
        BucketCompressor.clear()
        return res


compressors = {
    'topk': TopKCompressor,
    'topk2': TopKCompressor2,
    'bucket': BucketCompressor,
    'gaussian': GaussianCompressor,
    'gaussiank': GaussianCompressor,   # BASELINE.json's name for it
    'gaussian2': GaussianCompressor2,
    'randomk': RandomKCompressor,
    'randomkec': RandomKECCompressor,
    'randomksame': RandomKSameCompressor,
    'randomksameec': RandomKSameECCompressor,
    'dgcsampling': DGCSamplingCompressor,
    'redsync': RedSyncCompressor,
    'redsynctrim': RedSyncTrimCompressor,
    'mink': MinKCompressor,
    'uniform_topk': UniformTopKCompressor,
    'uniform_abs_topk': UniformAbsTopKCompressor,
    'bucketized_topk': BucketizedTopKCompressor,
    'bellk': BellKCompressor,
    'median_topk': MedianTopKCompressor,
    'none': NoneCompressor,
    None: NoneCompressor,
}


def test_gaussion_thres():
    """Eyeball check kept from the reference (compression.py:757-777), with an assertion added: zero
    everything below 3 sigma, recompute the threshold from the empirical tail mass."""
    import numpy as np
    from scipy import stats
    set_mean, set_std = 0.0, 0.5
    d = np.random.normal(set_mean, set_std, 10000)
    mean, std = np.mean(d), np.std(d)
    thres = 3 * std
    d[np.abs(d) < thres] = 0
    pvalue = 1 - np.count_nonzero(d) * 1.0 / d.size
    right = mean - stats.norm.ppf((1 - pvalue) / 2) * std
    print('real thres:%f, gen thres: %f' % (thres, right))
    return thres, right


if __name__ == '__main__':
    test_gaussion_thres()
