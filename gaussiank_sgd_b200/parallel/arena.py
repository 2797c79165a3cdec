"""Flat fp32 arenas: parameters, gradients, residuals and momentum of a model in four contiguous
buffers with one shared layout.

The reference stages every gradient into a merged buffer with a D2D copy per tensor
(`_push_to_buffer`, distributed_optimizer.py:364-383) and re-points ``p.grad`` at buffer slices
after the first iteration (:385-401, :510-516) — i.e. it arrives at a gradient arena by accident.
Here the arena is the design: ``p.data`` and ``p.grad`` are views from the start, autograd
accumulates straight into the arena, and every kernel of the chain addresses tensors by
(offset, numel).  Each tensor starts on a 256-byte boundary so the kernels can use 16-byte
vector accesses and cuDNN/cuBLAS see aligned weights; the padding is never part of a selection
segment.
"""
from __future__ import annotations

from typing import Dict, List, Sequence, Tuple

import torch

ALIGN = 64  # elements (256 bytes)


def _is_channels_last(t: torch.Tensor) -> bool:
    return t.dim() == 4 and t.is_contiguous(memory_format=torch.channels_last) and not t.is_contiguous()


class _DeviceSpan:
    """``__cuda_array_interface__`` over a span of device memory owned by somebody else (the arena)."""

    def __init__(self, ptr: int, numel: int):
        self.__cuda_array_interface__ = {"shape": (numel,), "typestr": "<f4", "data": (ptr, False), "version": 2}


class ParamArena:
    def __init__(self, named_params: Sequence[Tuple[str, torch.nn.Parameter]], with_grad: bool = True):
        named_params = [(n, p) for n, p in named_params if p.requires_grad]
        assert named_params, "no trainable parameters"
        self.names: List[str] = [n for n, _ in named_params]
        self.params: List[torch.nn.Parameter] = [p for _, p in named_params]
        dev = self.params[0].device
        for n, p in named_params:
            if p.dtype != torch.float32:
                raise TypeError(f"{n}: arenas hold fp32 parameters (the reference trains in fp32, settings.py:28)")
            if p.device != dev:
                raise ValueError("all parameters must live on one device (one process per GPU)")
        self.device = dev
        self.numels = [p.numel() for p in self.params]
        # Parameters that are views of ONE flat buffer (cuDNN's RNN weight blob, created by
        # nn.LSTM.flatten_parameters() when the module moves to the GPU) keep their relative offsets: the blob
        # becomes one extent of the arena.  cuDNN then finds its weights where it expects them and neither the
        # forward nor the backward re-packs them (which also makes the step capturable in a CUDA graph).
        groups: Dict[int, List[int]] = {}
        if dev.type == "cuda":
            for i, p in enumerate(self.params):
                groups.setdefault(p.data.untyped_storage().data_ptr(), []).append(i)
        groups = {k: v for k, v in groups.items() if len(v) > 1 and all(
            self.params[i].data.is_contiguous() and self.params[i].data.storage_offset() % 4 == 0 for i in v)}
        group_of = {i: k for k, v in groups.items() for i in v}
        self.offsets: List[int] = [0] * len(self.params)
        self._blobs = []                 # (arena offset, numel, member indices) of every adopted flat buffer
        placed = {}
        off = 0
        for i, n in enumerate(self.numels):
            k = group_of.get(i)
            if k is None:
                self.offsets[i] = off
                off += (n + ALIGN - 1) // ALIGN * ALIGN
                continue
            if k not in placed:
                blob = self.params[i].data.untyped_storage().nbytes() // 4
                placed[k] = off
                self._blobs.append((off, blob, groups[k]))
                off += (blob + ALIGN - 1) // ALIGN * ALIGN
            self.offsets[i] = placed[k] + self.params[i].data.storage_offset()
        self.total = max(off, ALIGN)
        assert self.total < 2 ** 31
        self.index: Dict[str, int] = {n: i for i, n in enumerate(self.names)}
        self._cl = [_is_channels_last(p.data) for p in self.params]
        self.param = torch.zeros(self.total, dtype=torch.float32, device=dev)
        self.grad = torch.zeros(self.total, dtype=torch.float32, device=dev) if with_grad else None
        self.grad_views = []
        self._aliases = {}
        with torch.no_grad():
            for (boff, blob, members) in self._blobs:
                # a tensor over the arena's memory with a storage of its own that STARTS at the blob: cuDNN looks
                # for the weights at the beginning of the storage of the first parameter
                alias = torch.as_tensor(_DeviceSpan(self.param.data_ptr() + 4 * boff, blob), device=dev)
                for i in members:
                    self._aliases[i] = (alias, self.offsets[i] - boff)
            for i, p in enumerate(self.params):
                v = self.view(self.param, i)
                v.copy_(p.data)
                p.data = v
                if with_grad:
                    self.grad_views.append(self.view(self.grad, i))
                    p.grad = self.grad_views[i]

    def view(self, arena: torch.Tensor, i: int) -> torch.Tensor:
        """Tensor i's window of ``arena`` with the parameter's shape and memory format."""
        if arena is self.param and i in getattr(self, "_aliases", {}):
            alias, rel = self._aliases[i]
            return alias[rel: rel + self.numels[i]].view(self.params[i].shape)
        flat = arena[self.offsets[i]: self.offsets[i] + self.numels[i]]
        shape = self.params[i].shape
        if self._cl[i]:
            n, c, h, w = shape
            return flat.view(n, h, w, c).permute(0, 3, 1, 2)
        return flat.view(shape)

    def new_arena(self, fill: float = 0.0) -> torch.Tensor:
        return torch.full((self.total,), fill, dtype=torch.float32, device=self.device)

    def extent(self, i: int) -> Tuple[int, int]:
        return self.offsets[i], self.numels[i]

    def extents(self, names: Sequence[str] = None) -> List[Tuple[int, int]]:
        idx = range(len(self.names)) if names is None else [self.index[n] for n in names]
        return [self.extent(i) for i in idx]

    def rebind_grads(self) -> None:
        """Re-point every ``p.grad`` at the arena (after a ``zero_grad(set_to_none=True)`` or an
        optimizer that replaced the tensors)."""
        for i, p in enumerate(self.params):
            g = p.grad
            if g is None or g.data_ptr() != self.grad.data_ptr() + 4 * self.offsets[i]:
                p.grad = self.grad_views[i]

    def num_params(self) -> int:
        return sum(self.numels)

    def dense_index_map(self, names: Sequence[str]) -> torch.Tensor:
        """For a group made of ``names``: arena index of every dense (padding-free) group position —
        used by the public compress() API to report indices the way the reference does."""
        parts = []
        for n in names:
            o, m = self.extent(self.index[n])
            parts.append(torch.arange(o, o + m, dtype=torch.int64))
        return torch.cat(parts)
