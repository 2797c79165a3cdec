"""Layer-wise backward profiler feeding the merge planners (the reference's profiling.py:13-148).

The reference registers a tensor hook per parameter that calls ``torch.cuda.synchronize()`` and reads the
wall clock (:35-48), which serialises the backward pass it is trying to measure.  Here every hook only
records a CUDA event on the current stream; the per-layer times are the event-to-event distances read after
the pass has finished (wall clock on CPU devices).  ``benchmark(trainer)`` keeps the reference's contract:
5 warm-up + 50 timed backward passes -> ``(seq_keys, layerwise_times, layerwise_sizes)`` in forward order.
"""
from __future__ import annotations

import time
from typing import Dict, List

import numpy as np
import torch

from .models import lstm as lstmpy


class Profiling:
    def __init__(self, model: torch.nn.Module):
        if not isinstance(model, torch.nn.Module):
            raise ValueError("Not a valid model, please provide a 'nn.Module' instance.")
        self.model = model
        self._names: List[str] = []
        self._sizes: Dict[str, int] = {}
        self._marks: Dict[str, list] = {}
        self._start = None
        self._handles = []
        self._cuda = next(model.parameters()).is_cuda
        self._register_hooks()

    def _mark(self):
        if self._cuda:
            ev = torch.cuda.Event(enable_timing=True)
            ev.record()
            return ev
        return time.perf_counter()

    def _register_hooks(self):
        for name, p in self.model.named_parameters():
            if not p.requires_grad:
                continue
            self._names.append(name)
            self._sizes[name] = p.numel()
            self._marks[name] = []

            def hook(_p, name=name):
                self._marks[name].append(self._mark())
            self._handles.append(p.register_post_accumulate_grad_hook(hook))

    def start(self):
        """Call right before ``loss.backward()``."""
        self._start_marks = getattr(self, "_start_marks", [])
        self._start_marks.append(self._mark())

    def reset(self):
        for k in self._marks:
            self._marks[k] = []
        self._start_marks = []

    def remove(self):
        for h in self._handles:
            h.remove()
        self._handles = []

    def get_backward_seq_keys(self):
        return self._names[::-1]

    def get_layerwise_times(self):
        """(forward-order names, seconds spent producing each gradient = distance to the previously
        finished gradient in backward order, averaged over the recorded passes)."""
        if self._cuda:
            torch.cuda.synchronize()
        n_pass = len(self._start_marks)
        back = self._names[::-1]
        acc = np.zeros(len(back))
        for it in range(n_pass):
            prev = self._start_marks[it]
            stamps = []
            for name in back:
                m = self._marks[name][it]
                stamps.append(prev.elapsed_time(m) * 1e-3 if self._cuda else (m - prev))
                # events of one pass are totally ordered on the stream; measure from the pass start
            stamps = np.maximum.accumulate(np.asarray(stamps))
            acc += np.diff(np.concatenate([[0.0], stamps]))
        times_backward = acc / max(n_pass, 1)
        return self._names, list(times_backward[::-1])

    def get_sizes(self):
        return [self._sizes[n] for n in self._names]


def benchmark(trainer, warmup: int = 5, iters: int = 50):
    """Profile ``trainer``'s backward pass (profiling.py:95-148)."""
    net = trainer.net
    p = Profiling(net)
    data = trainer.data_iter() if trainer.trainloader is not None else None
    hidden = net.init_hidden() if trainer.dnn == 'lstm' else None
    for i in range(warmup + iters):
        if i == warmup:
            p.reset()
        staged = trainer._stage(data)
        if hidden is not None:
            hidden = lstmpy.repackage_hidden(hidden)
        loss, _, _, hidden = trainer._forward_loss(staged, hidden)
        for q in net.parameters():
            q.grad = None
        p.start()
        loss.backward()
        trainer._release_inputs()
    names, times = p.get_layerwise_times()
    sizes = p.get_sizes()
    p.remove()
    for q in net.parameters():
        q.grad = None
    return names, times, sizes
