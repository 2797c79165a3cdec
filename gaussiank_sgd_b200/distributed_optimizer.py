"""Distributed optimizer: sparsified synchronous data-parallel SGD on flat arenas.

Public surface of the reference module, used ``as hvd`` by the training loop:
``init() rank() size() local_rank() local_size() broadcast(t, root_rank) DistributedOptimizer(...)
broadcast_parameters(state_dict, root_rank) broadcast_optimizer_state(opt, root_rank)`` and the
optimizer methods ``step() synchronize() zero_grad() increase_one_epoch() get_current_density()``
plus the ``local`` flag (distributed_optimizer.py:38-585, 588-736).

What happens underneath is different by design:

* parameters, gradients, residuals and momentum live in four flat fp32 arenas
  (:mod:`parallel.arena`); ``p.grad`` are arena views, so autograd accumulates in place and the
  reference's per-tensor staging copy (`_push_to_buffer`, :364-383) does not exist;
* a group's send side is one kernel chain (:class:`ops.chain.CompressChain`) without host syncs; its
  compaction kernel stores the (index,value) pairs into every peer's receive slot over NVLink
  (:mod:`parallel.transport`) — the two ``hvd.allgather_async`` calls of :426-427 are gone;
* ``synchronize()`` launches the receive kernel (bounded spin on the peers' headers, scatter-add of
  value/P into the zeroed gradient arena: the *exact* mean of the workers' sparse vectors, also when
  the per-rank counts differ — the reference's equal-chunk split at :471 mis-partitions them);
* ``step()`` is one fused momentum-SGD (or LARS) kernel over the arenas with torch.optim.SGD
  arithmetic instead of 3-4 launches per tensor.
"""
from __future__ import annotations

import os
from typing import Dict, List, Optional

import numpy as np
import torch
import torch.distributed as dist

from . import settings
from .ops import chain as ch
from .parallel import comm, grouping
from .parallel.arena import ParamArena
from .parallel.comm import (allreduce_, barrier, broadcast, init, local_rank, local_size, rank,  # noqa: F401
                            shutdown, size)
from .parallel.transport import Transport, make_transport
from .settings import logger

DEFAULT_DYNAMIC_DENSITIES = [0.015625, 0.004, 0.001]  # distributed_optimizer.py:60
# compressors whose selection runs in the sm_100a / C++ kernels; every other registered key (the fork's
# ordering selectors: bucketized_topk, uniform_topk, uniform_abs_topk, bellk, median_topk) runs on the
# torch-op path of _launch_group_torch — same algorithm as compression.py, library sort, never a substitute
_KERNEL_COMPRESSORS = set(ch.MODES) | {"randomk", "randomkec", "randomksame", "randomksameec"}


def synchronize(handle=None):
    """Module-level ``hvd.synchronize``: wait for an async collective handle (None = no-op)."""
    if handle is not None and hasattr(handle, "wait"):
        handle.wait()
    return handle


def _compressor_name(compression) -> str:
    if compression is None:
        return "none"
    if isinstance(compression, str):
        return compression
    return getattr(compression, "key", None) or getattr(compression, "name", "none")


class _Group:
    """One communication group: tensors that are compressed and exchanged together."""

    def __init__(self, names: List[str], idx: int):
        self.names = names
        self.idx = idx
        self.chain: Optional[ch.CompressChain] = None
        self.transport: Optional[Transport] = None
        self.pending = 0          # gradients still missing in this backward pass
        self.launched = False
        self.extents = []
        self.lo = self.hi = 0     # arena span (dense path)
        self.work = None
        self.means = None
        self.gmeans = None
        self.ready_event = None


class _Phase:
    """NVTX range + optional CUDA-event pair around one phase (see _DistributedOptimizer._phase)."""

    def __init__(self, opt, name):
        self.opt, self.name = opt, name
        self.cuda = opt.device.type == "cuda"
        self.timed = opt._profiling and not (self.cuda and torch.cuda.is_current_stream_capturing())

    def __enter__(self):
        if self.cuda and settings.NVTX:
            torch.cuda.nvtx.range_push("gk:" + self.name)
        if self.timed:
            if self.cuda:
                self.e0 = torch.cuda.Event(enable_timing=True)
                self.e0.record()
            else:
                import time
                self.t0 = time.perf_counter()
        return self

    def __exit__(self, *exc):
        if self.timed:
            if self.cuda:
                e1 = torch.cuda.Event(enable_timing=True)
                e1.record()
                self.opt._phase_events.append((self.name, self.e0, e1))
            else:
                import time
                self.opt._phase_events.append((self.name, None, time.perf_counter() - self.t0))
        if self.cuda and settings.NVTX:
            torch.cuda.nvtx.range_pop()
        return False


class _DistributedOptimizer(torch.optim.Optimizer):
    def __init__(self, params, named_parameters, compression, is_sparse=False, density=0.001,
                 seq_layernames=None, layerwise_times=None, norm_clip=None, threshold=0, writer=None,
                 gradient_path=None, fused=True, transport=None, dynamic_densities="default",
                 force_compress=False, chunk=ch.DEFAULT_CHUNK, overlap=None, steal_grads=None):
        super(self.__class__, self).__init__(params)
        self._compression = compression
        self._compressor = _compressor_name(compression)
        self._density = float(density)
        self._sparse = bool(is_sparse) and self._compressor not in ("none", "bucket")
        self._seq_layernames = seq_layernames
        self._layerwise_times = layerwise_times
        self._norm_clip = norm_clip
        self._threshold = threshold
        self._writer = writer
        self._gradient_path = gradient_path
        self._profiling = settings.PROFILE_PHASES
        self._timers = {"compress": [], "comm": [], "update": []}
        self._phase_events = []
        self._verify_every = settings.VERIFY_EVERY
        self._verify = None
        self.train_epoch = 0
        self.train_iter = 0
        if dynamic_densities == "default":
            dynamic_densities = settings.DYNAMIC_DENSITIES
        self._dynamic_densities = list(dynamic_densities) if dynamic_densities else None
        logger.info('_dynamic_densities: %s', self._dynamic_densities)
        self._selected_num_gradients: List[int] = []
        self.local = False
        self._force_compress = force_compress or settings.COMPRESS_SINGLE
        self._chunk = chunk
        self._transport_kind = transport
        self._random_seed = 0
        self.loss_scale = 1.0         # gradients arrive multiplied by this (fp16 training); divided out on receipt
        self.graph_version = 0        # bumped whenever device pointers a captured step depends on change
        self._torch_select = False
        self._steal_req = settings.STEAL_GRADS if steal_grads is None else bool(steal_grads)

        named_parameters = list(named_parameters) if named_parameters is not None else []
        if any(not isinstance(p, tuple) for p in named_parameters):
            raise ValueError('named_parameters should be a sequence of tuples (name, parameter), '
                             'usually produced by model.named_parameters().')
        if not named_parameters:
            named_parameters = [('allreduce.noname.%d.%d' % (gi, i), v)
                                for gi, g in enumerate(self.param_groups) for i, v in enumerate(g['params'])]
        in_groups = {id(p) for g in self.param_groups for p in g['params']}
        named_parameters = [(n, p) for n, p in named_parameters if p.requires_grad and id(p) in in_groups]
        self._named_parameters: Dict[str, torch.nn.Parameter] = dict(named_parameters)
        self._parameter_names = {p: n for n, p in named_parameters}
        self._sequential_keys = list(seq_layernames) if seq_layernames is not None else [n for n, _ in named_parameters]

        # ---- arenas ------------------------------------------------------------------------
        self._arena = ParamArena(named_parameters)
        self.device = self._arena.device
        self._resid = self._arena.new_arena() if self._sparse or self._compressor == "bucket" else None
        self._active = size() > 1 or self._force_compress   # distributed_optimizer.py:97: 1 worker => plain optimizer
        self._steal = (self._steal_req and self._active and self._sparse
                       and self._compressor in ch.MODES)       # random-k / torch-op selectors read the arena
        group_of_tensor = self._param_group_index()
        self._tables = ch.FlatTables(self._arena.extents(), group_of_tensor, self.device, chunk)
        self._fused_kind = self._detect_fused(fused)
        self._mom = None
        self._hyper_host = None
        self._hyper_dev = None
        self._hyper_key = None
        self._steps_done = 0
        if self._fused_kind is not None:
            self._init_fused_state()
        self._clip_coef = None     # device scalar consumed by the next fused step

        # ---- groups / chains ----------------------------------------------------------------
        self._generate_merged_parameters()
        self._grad_state = "zero"
        self._hooks = []
        self._overlap = (len(self._groups) > 1) if overlap is None else overlap
        self._comm_stream = None
        if self._active:
            self._register_hooks()

    # ------------------------------------------------------------------ set-up helpers
    def _param_group_index(self) -> List[int]:
        gid = {}
        for gi, g in enumerate(self.param_groups):
            for p in g['params']:
                gid[id(p)] = gi
        return [gid[id(p)] for p in self._arena.params]

    def _detect_fused(self, fused):
        if not fused:
            return None
        base = type(self).__mro__[1]
        if base.__name__ == "LARS":
            return "lars"
        if base is torch.optim.SGD or base.__name__ == "SGD":
            g0 = self.param_groups[0]
            if g0.get("maximize", False):
                return None
            return "sgd"
        return None

    def _init_fused_state(self):
        """Momentum (or LARS acceleration) arena with views installed as the optimizer state, so
        ``state_dict()`` keeps torch's format (checkpointed, unlike the reference)."""
        if self._fused_kind == "lars":
            self._mom = self._arena.new_arena(1.0)      # lars.py:116-119: acceleration starts at ones
            key = "acceleration"
        else:
            self._mom = self._arena.new_arena(0.0)
            key = "momentum_buffer"
        need = self._fused_kind == "lars" or any(g.get("momentum", 0) != 0 for g in self.param_groups)
        if need:
            for i, p in enumerate(self._arena.params):
                self.state[p][key] = self._arena.view(self._mom, i)
        else:
            self._mom = None

    def _init_fused_state_from_loaded(self):
        """After ``load_state_dict``: move the loaded momentum tensors back into the arena views."""
        if self._mom is None:
            return
        key = "acceleration" if self._fused_kind == "lars" else "momentum_buffer"
        with torch.no_grad():
            for i, p in enumerate(self._arena.params):
                view = self._arena.view(self._mom, i)
                buf = self.state[p].get(key)
                if buf is not None and buf.data_ptr() != view.data_ptr():
                    view.copy_(buf.to(view.device))
                self.state[p][key] = view
        self._hyper_key = None

    def _generate_merged_parameters(self):
        sizes = {k: self._named_parameters[k].numel() for k in self._sequential_keys}
        self._sizes = [sizes[k] for k in self._sequential_keys][::-1]
        P = size()
        info = {}
        if settings.ADAPTIVE_MERGE and self._layerwise_times is not None and self._seq_layernames is not None:
            if self._density < 1:
                groups, gof, info = grouping.groups_mgs(self._seq_layernames, sizes, self._layerwise_times,
                                                        max(P, 2), self._density)
            else:
                groups, gof, info = grouping.groups_mgwfbp(self._seq_layernames, sizes, self._layerwise_times,
                                                           max(P, 2))
        else:
            groups, gof = grouping.groups_by_threshold(self._sequential_keys, sizes, self._threshold)
        logger.info('# of parameters: %d', int(np.sum(self._sizes)))
        logger.info('Total number of tensors: %s', len(self._sizes))
        logger.info('Merged Number of groups: %s', len(groups))
        if info:
            logger.info('Merge planner: %s', info)
        self._key_groupidx_maps = gof
        self._per_tensor = len(groups) == len(self._sequential_keys) and len(groups) > 1
        arena = self._arena
        if self._per_tensor:
            # threshold=0: every tensor is its own selection segment, but all of them go through ONE
            # segmented launch instead of 161 chains.
            # Inside a group the tensors are taken in ARENA order (the planners work in backward order): the
            # compaction kernel emits pairs in table order, so every rank's list comes out sorted by index,
            # which the folded receive relies on.  Selection does not depend on the order.
            g = _Group(sorted((n for grp in groups for n in grp), key=lambda n: arena.offsets[arena.index[n]]), 0)
            g.segments = [[arena.extent(arena.index[n])] for n in g.names]
            self._groups = [g]
        else:
            self._groups = []
            for gi, names in enumerate(groups):
                g = _Group(sorted(names, key=lambda n: arena.offsets[arena.index[n]]), gi)
                g.segments = [[arena.extent(arena.index[n]) for n in g.names]]
                self._groups.append(g)
        self._group_of_param = {}
        for g in self._groups:
            for n in g.names:
                self._group_of_param[self._named_parameters[n]] = g
            g.extents = [arena.extent(arena.index[n]) for n in g.names]
            g.lo = min(o for o, _ in g.extents)
            g.hi = max(o + m for o, m in g.extents)
        self._layerwise_compressors = {':'.join(g.names): self._density for g in self._groups}
        if not self._active:
            return
        for g in self._groups:
            if self._sparse:
                comp = self._compressor
                use_res = comp not in ("topk2", "gaussian2", "randomk", "randomksame")
                if comp in ch.MODES:
                    kcomp = comp
                elif comp.startswith("randomk"):
                    kcomp = "threshold"        # random-k only uses K1 of the chain (gk_randomk_send selects)
                else:
                    # The fork's ordering selectors need a sort; they run with torch ops (library sort) on the
                    # tensors of the group and hand their (index, value) pairs to the same transports.  The
                    # chain object only provides tables, slots and the control block here.
                    if self._compression is None or not hasattr(self._compression, "compress"):
                        raise ValueError(f"compressor {comp!r} has no kernel and no compress() implementation")
                    kcomp = "threshold"
                    self._torch_select = True
                    logger.warning("compressor %r runs on the torch-op path (library sort, eager, NCCL all_gather); "
                                   "the fused kernels cover %s", comp, sorted(_KERNEL_COMPRESSORS))
                g.chain = ch.CompressChain(arena.grad, self._resid, g.segments, compressor=kcomp,
                                           density=min(self.get_current_density(), 0.999), use_residual=use_res,
                                           zero_grad=True, chunk=self._chunk)
                g.transport = make_transport(g.chain.cap, self.device,
                                             "allgather" if self._torch_select and size() > 1 else self._transport_kind)
                if rank() == 0:
                    logger.info('group %d: %d tensors, %d elements, cap %d pairs, transport %s', g.idx,
                                len(g.names), g.chain.n_total, g.chain.cap, g.transport.name)
            elif self._compressor == "bucket":
                g.chain = ch.CompressChain(arena.grad, self._resid, g.segments, compressor="topk",
                                           density=0.5, use_residual=False, zero_grad=False, chunk=self._chunk)
                g.means = torch.zeros(2 * g.chain.nseg, dtype=torch.float32, device=self.device)
                g.gmeans = torch.zeros_like(g.means)

    def _register_hooks(self):
        for p in self._arena.params:
            self._hooks.append(p.register_post_accumulate_grad_hook(self._make_hook(p)))
        for g in self._groups:
            g.pending = len(g.names)
        if self._steal:
            # Gradients are taken over from autograd instead of being accumulated into the arena: with
            # p.grad unset, AccumulateGrad hands the freshly computed tensor to p.grad without a kernel and
            # K1 reads it in place (GkChain.gsrc).  Saves one add kernel per parameter tensor and step
            # (161 for ResNet-50) and the zeroing of the gradient arena.
            for p in self._arena.params:
                p.grad = None

    def _collect_grad_sources(self, g: _Group) -> None:
        """Point K1 at the tensors autograd produced for group ``g`` (arena layout required: same strides
        as the parameter, which AccumulateGrad guarantees for dense parameters; anything else is copied into
        the arena first)."""
        arena = self._arena
        ptrs, dirty = [], []
        for n in g.names:
            i = arena.index[n]
            p = arena.params[i]
            view = arena.grad_views[i]
            gr = p.grad
            if gr is None:
                ptrs.append(view.data_ptr())                      # no gradient this step: arena zeros
            elif gr.data_ptr() == view.data_ptr():
                ptrs.append(view.data_ptr())
                dirty.append(view)
            elif (gr.dtype == torch.float32 and gr.device == p.device and gr.data_ptr() % 16 == 0
                  and all(a == b or n == 1 for a, b, n in zip(gr.stride(), p.stride(), p.shape))):
                # same element order in memory (strides of size-1 dimensions are arbitrary: the gradient of a
                # 1x1 convolution weight comes back with other ones than the channels_last parameter has)
                ptrs.append(gr.data_ptr())
            else:
                view.copy_(gr)
                ptrs.append(view.data_ptr())
                dirty.append(view)
        g.chain.set_grad_sources(ptrs)
        g._dirty_views = dirty

    def _release_grad_sources(self, g: _Group, keep_alive: bool = False) -> None:
        """After the chain is enqueued: the arena views become the visible gradients again (they will hold
        the averaged sparse gradient); arena regions that served as a dense source are cleared.  When the chain
        runs on the side stream (``keep_alive``) the tensors autograd produced stay referenced until
        ``synchronize()`` has joined the streams: they were allocated on the compute stream, and dropping the
        last reference here would let the caching allocator hand their memory to backward kernels that run
        while K1 is still reading them."""
        arena = self._arena
        for v in getattr(g, "_dirty_views", ()):
            v.zero_()
        g._dirty_views = []
        if keep_alive:
            g._stolen = [arena.params[arena.index[n]].grad for n in g.names]
        for n in g.names:
            i = arena.index[n]
            arena.params[i].grad = arena.grad_views[i]

    def _launch_group_torch(self, g: _Group) -> None:
        """Send side for the selectors that have no kernel (the fork's ordering selectors,
        compression.py:60-177): the registry object's own ``compress()`` — torch ops, library sort — selects on
        the dense group vector and keeps its residual under the group's name, exactly as it would under the
        reference's optimizer (distributed_optimizer.py:416-427); the pairs are packed into the send slot in
        the wire format of the kernels and travel through the all-gather transport."""
        arena, chain = self._arena, g.chain
        if not hasattr(g, "_dense_map"):
            g._dense_map = arena.dense_index_map(g.names).to(self.device)
            g._torch_seq = 0
        acc = arena.grad.index_select(0, g._dense_map)
        arena.grad[g.lo:g.hi].zero_()
        _t, idx, vals = self._compression.compress(acc, "gk_group%d" % g.idx, ratio=min(self.get_current_density(), 0.999))
        idx = idx.reshape(-1)[:chain.cap].to(torch.int64)
        vals = vals.reshape(-1)[:chain.cap].to(torch.float32).contiguous()
        aidx = g._dense_map[idx]
        g._torch_seq += 1
        words = aidx | ((vals.view(torch.int32).to(torch.int64) & 0xFFFFFFFF) << 32)
        send = g.transport.send_tensor()
        send[0] = int(idx.numel()) | (g._torch_seq << 32)
        send[ch.N.GK_SLOT_HDR: ch.N.GK_SLOT_HDR + idx.numel()] = words
        chain.ctrl.view(torch.int32)[10] = g._torch_seq          # GkCtrl.seq: what the receive kernel keys on
        chain.ctrl.view(torch.int32)[8] = int(idx.numel())       # GkCtrl.total
        self._selected_num_gradients.append(int(idx.numel()))

    def _make_hook(self, p):
        group = self._group_of_param[p]

        def hook(_param):
            self._grad_state = "dense"
            if self.local:
                return
            group.pending -= 1
            if group.pending == 0 and self._overlap and not group.launched:
                self._launch_group(group, from_hook=True)
        return hook

    # ------------------------------------------------------------------ density schedule
    def increase_one_epoch(self):
        """distributed_optimizer.py:136-144: advance the density schedule and log the selection stats."""
        self.train_epoch += 1
        if rank() == 0 and self._sparse and self._active:
            density = self.get_current_density()
            total = int(np.sum(self._sizes))
            k = max(int(total * density), 1)
            sel = self.selected_num_gradients()
            if sel:
                logger.info('Average number of selected gradients: %f, exact k: %d', float(np.mean(sel)), k)
                logger.info('The number of selected gradients: %s', sel)
        self._selected_num_gradients = []
        self._apply_density()

    def get_current_density(self, name=None):
        """Effective density = max(user density, schedule[epoch]) (distributed_optimizer.py:146-160)."""
        density = self._density
        if self._dynamic_densities:
            sched = self._dynamic_densities[min(self.train_epoch, len(self._dynamic_densities) - 1)]
            density = max(self._density, sched)
        return density

    def _apply_density(self):
        """Push the current density into the chains.  The receive arenas follow it: their slots hold 8x k pairs
        of the CURRENT density, so after the two warm-up epochs of the reference's schedule (:60) the
        peer-mapped arena of ResNet-50 shrinks from 409 MB to 26 MB per rank.  Re-creating a transport is a
        collective (every rank changes density at the same step) and invalidates captured CUDA graphs
        (``graph_version``)."""
        if not (self._active and self._sparse):
            return
        d = min(self.get_current_density(), 0.999)
        for g in self._groups:
            if g.chain is None:
                continue
            g.chain.set_density(d)
            if g.chain.auto_cap and g.transport is not None and g.chain.cap_for(d) != g.chain.cap:
                if self.device.type == "cuda":
                    torch.cuda.synchronize(self.device)
                g.transport.close()
                g.transport = None
                g.chain.set_cap(g.chain.cap_for(d))
                g.transport = make_transport(g.chain.cap, self.device,
                                             "allgather" if self._torch_select and size() > 1 else self._transport_kind)
                self.graph_version += 1
                if rank() == 0:
                    logger.info('group %d: density %g -> slot capacity %d pairs (%.1f MB receive arena)', g.idx, d,
                                g.chain.cap, 16e-6 * size() * (g.chain.cap + 2))

    def selected_num_gradients(self) -> List[int]:
        """Pairs selected by the most recent sends (<= 64, oldest first), read from the device-side ring
        (the reference appends ``ctx.numel()`` on the host every iteration, :420)."""
        out: List[int] = []
        for g in self._groups:
            if g.chain is None or not self._sparse:
                continue
            c = g.chain.read_ctrl()
            seq = c["seq"]
            n = min(seq, 64)
            out.extend(int(c["ring"][(seq - n + 1 + i) & 63]) for i in range(n))
        return out

    # ------------------------------------------------------------------ send side
    def _launch_group(self, g: _Group, from_hook: bool = False):
        if g.launched:
            return
        g.launched = True
        stream_ctx = None
        if self._overlap and self.device.type == "cuda":
            if self._comm_stream is None:
                self._comm_stream = torch.cuda.Stream(self.device)
            self._comm_stream.wait_stream(torch.cuda.current_stream(self.device))
            stream_ctx = torch.cuda.stream(self._comm_stream)
            stream_ctx.__enter__()
        try:
            if self._sparse:
                comp = self._compressor
                dump = (settings.LOGGING_GRADIENTS and self._gradient_path and rank() == 0
                        and self.train_iter % max(settings.GRADIENT_DUMP_EVERY, 1) == 0)
                seed = self._random_seed if comp.startswith("randomksame") else self._random_seed * 1000003 + rank() + 1
                if self._torch_select:
                    self._launch_group_torch(g)
                elif comp.startswith("randomk"):
                    g.chain.randomk_send(g.transport.dst(), seed)
                else:
                    if self._steal:
                        self._collect_grad_sources(g)
                    if dump:
                        # opt-in gradient-distribution study (the reference does this every iteration by
                        # default, distributed_optimizer.py:422-424)
                        self._dump_gradients(g)
                    if self._verify_every and self.train_iter % self._verify_every == 0:
                        self._verify_snapshot(g)
                    with self._phase("compress"):
                        g.chain.compress(g.transport.dst(), seed)     # the seed only matters to dgcsampling
                    if self._steal:
                        self._release_grad_sources(g, keep_alive=stream_ctx is not None)
                g.transport.exchange()
            elif self._compressor == "bucket":
                g.chain.bucket_means(g.means)
                g.gmeans.copy_(g.means)
                allreduce_(g.gmeans, average=True)
            else:
                buf = self._arena.grad[g.lo:g.hi]
                g.work = allreduce_(buf, average=True, async_op=True)
        finally:
            if stream_ctx is not None:
                stream_ctx.__exit__(None, None, None)

    def _dump_gradients(self, g: _Group) -> None:
        """``<gradient_path>/r<rank>_gradients_iter_<i>.npy`` = acc (gradient + residual) of the group in the
        reference's dense order, for logs/plot_cdf-style distribution plots.  Host copy: debugging only."""
        if not hasattr(g, "_dense_map"):
            g._dense_map = self._arena.dense_index_map(g.names).to(self.device)
        if self._steal:
            arena = self._arena
            flat = torch.zeros_like(arena.grad)
            for n in g.names:
                i = arena.index[n]
                if arena.params[i].grad is not None:
                    arena.view(flat, i).copy_(arena.params[i].grad)
            acc = flat.index_select(0, g._dense_map)
        else:
            acc = self._arena.grad.index_select(0, g._dense_map)
        if g.chain.use_residual:
            acc = acc + self._resid.index_select(0, g._dense_map)
        acc = acc.cpu().numpy()
        os.makedirs(self._gradient_path, exist_ok=True)
        np.save('%s/r%d_gradients_iter_%d' % (self._gradient_path, rank(), self.train_iter), acc)

    def gradient_histogram(self, lo: float, hi: float, nbins: int = 256) -> torch.Tensor:
        """On-device histogram of the residual arena (acc right after a compress) — the cheap alternative
        to dumping the whole tensor; returns an int32 tensor of ``nbins`` counts."""
        import ctypes as C
        assert self._resid is not None and self.device.type == "cuda"
        out = torch.zeros(nbins, dtype=torch.int32, device=self.device)
        ch.N.check(ch.N.cuda_lib().gk_histogram(self._resid.data_ptr(), self._resid.numel(), C.c_float(lo),
                                                C.c_float(hi), nbins, out.data_ptr(),
                                                torch.cuda.current_stream(self.device).cuda_stream), "gk_histogram")
        return out

    # ------------------------------------------------------------------ phase timers / NVTX
    def _phase(self, name: str):
        """Context manager around one phase of the step: an NVTX range when ``GK_NVTX=1`` and, when
        ``GK_PROFILE_PHASES=1``, a CUDA-event pair on the current stream (the reference wraps the same three
        phases in host timers, distributed_optimizer.py:428-429, :465-466, :484-485)."""
        return _Phase(self, name)

    def _print_profiling(self):
        """Every 40 samples: mean device time of the phases, in the reference's line format (:522-540)."""
        if not self._profiling or len(self._phase_events) < 40 * 2:
            return
        if self.device.type == "cuda":
            torch.cuda.synchronize(self.device)
        acc: Dict[str, List[float]] = {}
        for name, e0, e1 in self._phase_events:
            acc.setdefault(name, []).append(e0.elapsed_time(e1) * 1e-3 if e0 is not None else e1)
        self._phase_events = []
        for k, v in acc.items():
            self._timers.setdefault(k, []).extend(v)
        tcp = float(np.mean(acc.get("compress", [0.0])))
        tar = float(np.mean(acc.get("comm", [0.0])))
        tup = float(np.mean(acc.get("update", [0.0])))
        tcu = float(np.mean(acc.get("comm+update", [0.0])))
        if rank() == 0:
            logger.info('[%d]: Total compress: %f, allreduce: %f, update: %f, total: %f', rank(), tcp, tar + tcu, tup,
                        tcp + tar + tup + tcu)
        self.last_phase_means = {"compress": tcp, "comm": tar, "update": tup, "comm+update": tcu}

    # ------------------------------------------------------------------ oracle cross-check (GK_VERIFY_EVERY)
    def _verify_snapshot(self, g: _Group) -> None:
        """Before the chain of a verify step runs: keep a copy of acc = gradient (+ residual) in arena
        layout.  compute-sanitizer is not available on every pool; this is the in-situ check that the fused
        path still agrees with the oracle."""
        arena = self._arena
        flat = torch.zeros_like(arena.grad)
        for n in g.names:
            i = arena.index[n]
            gr = arena.params[i].grad
            if gr is not None:
                arena.view(flat, i).copy_(gr)
        if g.chain.use_residual:
            flat += self._resid
        self._verify = (g, flat, self.train_iter)

    def _verify_check(self) -> None:
        """After the materialised receive of a verify step: (1) this rank's pairs are exactly the oracle's
        selection on the snapshot, (2) the dense gradient is exactly the rank-ordered mean of all ranks' pairs
        (distributed_optimizer.py:468-482 with the exact mean), (3) the replicas hold identical parameters."""
        from .ops import oracle
        g, acc, it = self._verify
        self._verify = None
        if self.device.type == "cuda":
            torch.cuda.synchronize(self.device)
        chain, P, me = g.chain, size(), rank()
        ctrl = chain.read_ctrl()
        if ctrl["err"]:
            raise RuntimeError(f"verify step {it}: chain error flag {ctrl['err']} on rank {me}")
        src = g.transport.src()
        nslots, stride = src.nsrc, int(src.slot_stride)
        parity = (ctrl["seq"] & 1) if src.parity_stride else 0
        words = self._recv_words(g)
        lists = []
        for r in range(nslots):
            base = (parity * nslots + r) * stride
            hdr = int(words[base])
            cnt = min(hdr & 0xFFFFFFFF, chain.cap)
            w = words[base + ch.N.GK_SLOT_HDR: base + ch.N.GK_SLOT_HDR + cnt]
            lists.append(((w & 0xFFFFFFFF).to(torch.int64), (w >> 32).to(torch.int32).view(torch.float32)))
        mine = lists[me if nslots > 1 else 0]
        comp = self._compressor
        acc_c = acc.cpu()
        if comp in ("gaussian", "gaussiank", "gaussian2", "topk", "topk2", "mink", "redsync", "redsynctrim"):
            stats = chain.read_stats()
            exp = []
            for sidx, extents in enumerate(g.segments):
                amap = torch.cat([torch.arange(o, o + m) for o, m in extents])
                seg = acc_c[amap]
                if comp.startswith("gaussian"):
                    oi, _ = oracle.gaussian_select(seg, chain.density, chain.nloops, mean=float(stats[sidx]["mean"]),
                                                   std=float(stats[sidx]["std"]))
                else:
                    oi = oracle.select(comp, seg, chain.density)
                exp.append(amap[oi])
            exp = torch.cat(exp)[:chain.cap]
            if not torch.equal(mine[0].cpu(), exp):
                raise RuntimeError(f"verify step {it} rank {me}: selection differs from the oracle "
                                   f"({mine[0].numel()} vs {exp.numel()} indices)")
            if not torch.equal(mine[1].cpu(), acc_c[exp]):
                raise RuntimeError(f"verify step {it} rank {me}: sent values differ from acc[idx]")
        dense = torch.zeros_like(acc_c)
        scale = torch.tensor(1.0 / (P * self.loss_scale), dtype=torch.float32)
        for idx, val in lists:
            dense[idx.cpu()] += val.cpu() * scale
        if not torch.equal(dense, self._arena.grad.cpu()):
            bad = int((dense != self._arena.grad.cpu()).sum())
            raise RuntimeError(f"verify step {it} rank {me}: received gradient differs from the rank-ordered mean "
                               f"of the pairs at {bad} coordinates")
        if P > 1:
            cs = self.param_checksum()
            allcs = [None] * P
            dist.all_gather_object(allcs, cs)
            if any(c != allcs[0] for c in allcs):
                raise RuntimeError(f"verify step {it}: replicas differ: {allcs}")
        self.verified_steps = getattr(self, "verified_steps", 0) + 1
        if me == 0:
            logger.info('verify step %d ok: %d pairs sent, %d received, replicas identical', it, mine[0].numel(),
                        sum(l[0].numel() for l in lists))

    def _recv_words(self, g: _Group) -> torch.Tensor:
        """The local receive arena of group ``g`` as an int64 CPU tensor (debug / verification only)."""
        t = g.transport
        for attr in ("recv", "buf"):
            if hasattr(t, attr):
                return getattr(t, attr).t.cpu()
        return t.local_words().cpu()

    def param_checksum(self):
        """(fp64 sum, xor of the bit patterns) of the parameter arena: equal on every rank iff the replicas are
        bit-identical."""
        p = self._arena.param
        bits = p.view(torch.int32)
        x = bits[: bits.numel() // 2 * 2].view(torch.int64)
        xor = 0
        # xor-reduce on the device: fold halves until small, finish on the host
        while x.numel() > 1024:
            half = x.numel() // 2
            rest = x[2 * half:]
            x = torch.bitwise_xor(x[:half], x[half:2 * half])
            if rest.numel():
                x = torch.cat([x, rest])
        for v in x.cpu().tolist():
            xor ^= v & 0xFFFFFFFFFFFFFFFF
        return (float(p.double().sum().item()), xor)

    def status_tensor(self) -> Optional[torch.Tensor]:
        """Device tensor [total, sent, seq, err, overflow] of the first chain's control block as float32 (one
        tiny conversion kernel, no host sync): the trainer reads it back together with the loss and fails
        within one step of a receive time-out instead of at the end of training."""
        for g in self._groups:
            if g.chain is not None and self._sparse and self._active:
                return g.chain.ctrl.view(torch.int32)[8:13].to(torch.float32)
        return None

    # ------------------------------------------------------------------ receive side
    def synchronize(self):
        """Finish the exchange: afterwards every ``p.grad`` holds the averaged (sparse) gradient
        (distributed_optimizer.py:449-519)."""
        if not self._active:
            return
        P = size()
        any_launched = False
        for g in self._groups:
            if not g.launched and self._grad_state == "dense":
                self._launch_group(g)
            if not g.launched:
                continue
            any_launched = True
            if self._overlap and self._comm_stream is not None:
                torch.cuda.current_stream(self.device).wait_stream(self._comm_stream)
            g._stolen = None               # the side stream has been joined: autograd's tensors may go
            if self._sparse:
                with self._phase("comm"):
                    ch.scatter(g.transport.src(), g.chain.ctrl, self._arena.grad, 1.0 / (P * self.loss_scale))
                if self._verify is not None and self._verify[0] is g:
                    self._verify_check()
            elif self._compressor == "bucket":
                g.chain.bucket_reconstruct(g.means, g.gmeans)
            else:
                if g.work is not None:
                    g.work.wait()
                    g.work = None
                if self.loss_scale != 1.0:
                    self._arena.grad[g.lo:g.hi].mul_(1.0 / self.loss_scale)
            g.launched = False
            g.pending = len(g.names)
        if any_launched:
            self._random_seed += 1
            self.train_iter += 1
            # "synced": the dense / bucket exchange has run; a second synchronize() (the reference-style
            # synchronize(); clip; step() sequence calls it again from step()) must not communicate again
            self._grad_state = "sparse" if self._sparse else "synced"
            if self._norm_clip is not None:
                # honour norm_clip (the reference's factory discards it); same scaling
                # as its dense path (:498-505)
                self.clip_grad_norm(float(np.sqrt(1.0 / P) * self._norm_clip))

    def clip_grad_norm(self, max_norm: float) -> torch.Tensor:
        """Fused global-norm clipping of the arena (dist_trainer.py:80-85 does it with
        ``clip_grad_norm_`` after ``synchronize()``).  The coefficient stays on the device and is folded
        into the next fused step; with a non-fused optimizer the gradient is scaled in place."""
        out = ch.grad_norm(self._arena.grad, self._tables, max_norm)
        if self._fused_kind == "sgd":
            self._clip_coef = out[1:2]
        else:
            ch.scale_(self._arena.grad, self._tables, out[1:2])
        return out

    @property
    def graph_safe(self) -> bool:
        """True when a training step contains no host-issued collective, i.e. it can be captured in a
        CUDA graph: the P2P / local transports communicate from inside the kernels."""
        if not self._active:
            return True
        nccl_ok = (settings.NCCL_GRAPH and comm.is_distributed() and dist.get_backend() == "nccl"
                   and self.device.type == "cuda")      # replays of captured NCCL collectives work on this stack
        if not self._sparse:
            return size() == 1 or (nccl_ok and self._compressor == "none")
        if self._torch_select or self._verify_every:
            return False
        if nccl_ok and all(g.transport is not None and g.transport.name == "allgather" for g in self._groups):
            return True
        return all(g.transport is not None and g.transport.name in ("p2p", "local") for g in self._groups)

    def read_overflow(self) -> int:
        """Pairs withheld so far because a send exceeded the slot capacity (they stay in the residual)."""
        return sum(int(g.chain.read_ctrl()["overflow"]) for g in self._groups if g.chain is not None and self._sparse)

    def check_errors(self) -> None:
        """Host-side check of the device error flags (receive timeouts).  Synchronises."""
        for g in self._groups:
            if g.chain is not None and self._sparse:
                c = g.chain.read_ctrl()
                if c["err"]:
                    raise RuntimeError(f"sparse exchange of group {g.idx} timed out on rank {rank()} (seq {c['seq']})")

    # ------------------------------------------------------------------ update
    def _refresh_hyper(self):
        first = self._steps_done == 0
        key = tuple((g['lr'], g.get('momentum', 0.0), g.get('dampening', 0.0), g.get('weight_decay', 0.0),
                     bool(g.get('nesterov', False)), g.get('eeta', 0.0), g.get('epsilon', 0.0))
                    for g in self.param_groups) + (first,)
        if key == self._hyper_key:
            return
        self._hyper_key = key
        host = ch.make_hyper(self.param_groups, "cpu", first)
        if self.device.type == "cuda":
            if self._hyper_host is None:
                self._hyper_host = torch.empty_like(host).pin_memory()
                self._hyper_dev = torch.empty_like(host, device=self.device)
            self._hyper_host.copy_(host)
            self._hyper_dev.copy_(self._hyper_host, non_blocking=True)
        else:
            self._hyper_dev = host

    def fused_update(self):
        """The fused optimizer kernel alone (graph-capturable: reads lr & co from device memory)."""
        a = self._arena
        with self._phase("update"):
            if self._fused_kind == "sgd":
                ch.sgd_step(a.param, a.grad, self._mom, self._tables, self._hyper_dev, self._clip_coef)
            else:
                ch.lars_step(a.param, a.grad, self._mom, self._tables, self._hyper_dev)
        self._clip_coef = None

    def set_loss_scale(self, scale: float) -> None:
        """Static loss scale of fp16 training (dl_trainer.py:41-45 uses 128): the gradients, the residuals and the
        pairs on the wire live in the scaled domain (selection is scale-invariant), the receive side divides by
        it.  Supported on the sparse path and the dense all-reduce path with more than one worker (or
        ``force_compress``); clipping thresholds refer to the unscaled gradient, so LSTM recipes use bf16."""
        if scale != 1.0 and (not self._active or self._compressor == "bucket"):
            raise ValueError("fp16 loss scaling needs the exchange path (several workers or force_compress)")
        self.loss_scale = float(scale)
        self.graph_version += 1

    def can_fold(self) -> bool:
        """True when the pending exchange can be finished by the folded receive+update kernel: sparse path,
        fused SGD, ONE group whose send side is (or is about to be) launched, index-sorted pair lists (every
        kernel compressor except random-k), no clipping pending.  Otherwise ``step()`` materialises the
        averaged gradient first (``synchronize()``), as it does when the user calls ``synchronize()`` to look
        at or clip the gradients (LSTM models, dist_trainer.py:80-85)."""
        if not (settings.FOLDED_UPDATE and self._active and self._sparse and self._fused_kind == "sgd"):
            return False
        if len(self._groups) != 1 or self._norm_clip is not None or self._clip_coef is not None:
            return False
        if self._compressor.startswith("randomk") or self._torch_select or not self._tables.foldable:
            return False
        if size() > ch.N.GK_MAXP:
            return False
        if self._verify_every and (self._verify is not None or self.train_iter % self._verify_every == 0):
            return False                   # verify steps materialise the gradient so that it can be checked
        g = self._groups[0]
        return g.launched or self._grad_state == "dense"

    def folded_update(self):
        """Finish the exchange and update in one kernel (ops.chain.sgd_recv_step).  ``p.grad`` is NOT filled
        with the averaged gradient on this path (the arena stays zero) — call ``synchronize()`` before
        ``step()`` if you need to see it."""
        g = self._groups[0]
        if not g.launched:
            self._launch_group(g)
        if self._overlap and self._comm_stream is not None:
            torch.cuda.current_stream(self.device).wait_stream(self._comm_stream)
        a = self._arena
        skip = all(pg.get("momentum", 0.0) == 0 and pg.get("weight_decay", 0.0) == 0 for pg in self.param_groups)
        with self._phase("comm+update"):
            ch.sgd_recv_step(a.param, a.grad.view(torch.int32), self._mom, self._tables, self._hyper_dev,
                             g.transport.src(), g.chain.ctrl, 1.0 / (size() * self.loss_scale), skip_dense=skip)
        g.launched = False
        g.pending = len(g.names)
        g._stolen = None
        self._random_seed += 1
        self.train_iter += 1
        self._grad_state = "zero"          # the kernel leaves the gradient arena all zero

    def step(self, closure=None):
        """Synchronise the exchange, then update (distributed_optimizer.py:543-546)."""
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        if self._fused_kind is not None:
            self._refresh_hyper()
            if not self.local and self.can_fold():
                self.folded_update()
            else:
                if not self.local:
                    self.synchronize()
                self.fused_update()
            self._steps_done += 1
            self._print_profiling()
            return loss
        if not self.local:
            self.synchronize()
        return super(self.__class__, self).step()

    def zero_grad(self, set_to_none: bool = False):
        """Keep the arena views alive (torch's default ``set_to_none=True`` would drop them) and clear
        only what was written: after a sparse step just the received coordinates."""
        a = self._arena
        if self._grad_state == "sparse" and self._active and self._sparse and settings.SPARSE_ZERO:
            for g in self._groups:
                ch.unscatter(g.transport.src(), g.chain.ctrl, a.grad)
        elif self._grad_state != "zero" or not self._active:
            a.grad.zero_()
        self._grad_state = "zero"
        if self._steal:
            for p in a.params:
                p.grad = None          # the next backward hands its gradient tensors over (no accumulate kernels)
        else:
            a.rebind_grads()

    # ------------------------------------------------------------------ state
    def residual_state(self) -> Optional[torch.Tensor]:
        return self._resid

    def extra_state_dict(self) -> dict:
        """Compressor residuals + schedule position (not saved by the reference)."""
        return {"residual": None if self._resid is None else self._resid.detach().cpu().clone(),
                "train_epoch": self.train_epoch, "train_iter": self.train_iter,
                "steps_done": self._steps_done, "random_seed": self._random_seed}

    def load_extra_state_dict(self, sd: dict) -> None:
        if sd.get("residual") is not None and self._resid is not None and sd["residual"].numel() == self._resid.numel():
            self._resid.copy_(sd["residual"])
        self.train_epoch = int(sd.get("train_epoch", 0))
        self.train_iter = int(sd.get("train_iter", 0))
        self._steps_done = int(sd.get("steps_done", 0))
        self._random_seed = int(sd.get("random_seed", 0))
        self._hyper_key = None
        self._apply_density()

    def sync_resume_state(self, root_rank: int = 0) -> None:
        """After a resume: only ``root_rank`` has read the checkpoint (dist_trainer.py:26-27 loads it on rank 0
        only), so its momentum (or LARS acceleration) arena and its step / schedule counters are broadcast —
        otherwise the other ranks would run their next step as a *first* step (momentum buffer := gradient)
        and the replicas would drift apart for good.  Residuals are per-rank error feedback: ``root_rank``
        keeps the one it saved, the others start from zero.  (The reference restores none of this,
        SURVEY.md Q12, and stays consistent only because every rank then has zero momentum.)"""
        if not comm.is_distributed():
            return
        meta = None
        if rank() == root_rank:
            meta = {"steps_done": self._steps_done, "random_seed": self._random_seed, "train_iter": self.train_iter,
                    "train_epoch": self.train_epoch}
        meta = comm.bcast_object(meta, root_rank)
        self._steps_done = int(meta["steps_done"])
        self._random_seed = int(meta["random_seed"])
        self.train_iter = int(meta["train_iter"])
        self.train_epoch = int(meta["train_epoch"])
        if self._mom is not None:
            comm.broadcast_(self._mom, root_rank)
        elif self._fused_kind is None:
            broadcast_optimizer_state(self, root_rank)
        self._hyper_key = None
        self._apply_density()

    def close(self):
        for h in self._hooks:
            h.remove()
        self._hooks = []
        for g in self._groups:
            if g.transport is not None:
                g.transport.close()


def DistributedOptimizer(optimizer, named_parameters=None, compression=None, is_sparse=False, density=0.001,
                         seq_layernames=None, layerwise_times=None, norm_clip=None, threshold=0, writer=None,
                         gradient_path=None, **kwargs):
    """Wrap ``optimizer`` (factory of distributed_optimizer.py:550-585).  Like the reference, the result
    is an instance of a dynamically created subclass of the wrapped optimizer's class, so it *is* an
    ``SGD`` / ``LARS`` whose ``step()`` first finishes the gradient exchange.

    Extra keyword arguments (all optional): ``fused`` (default True: arena kernel instead of the per-tensor
    torch update), ``transport`` (auto|p2p|allgather), ``dynamic_densities`` (list or None; default the
    reference's warm-up schedule [0.015625, 0.004, 0.001]), ``force_compress`` (run the chain even with
    one worker), ``chunk``, ``overlap``.  ``norm_clip`` is honoured (the reference drops it, Q4)."""
    cls = type(optimizer.__class__.__name__, (optimizer.__class__,), dict(_DistributedOptimizer.__dict__))
    return cls(optimizer.param_groups, named_parameters, compression, is_sparse, density,
               seq_layernames=seq_layernames, layerwise_times=layerwise_times, norm_clip=norm_clip,
               threshold=threshold, writer=writer, gradient_path=gradient_path, **kwargs)


def broadcast_parameters(params, root_rank=0):
    """Make every rank start from ``root_rank``'s values (distributed_optimizer.py:588-617).  The
    reference broadcasts tensor by tensor; here tensors are packed per dtype into one flat buffer."""
    if isinstance(params, dict):
        items = sorted(params.items())
    elif isinstance(params, list):
        items = [p if isinstance(p, tuple) else (None, p) for p in params]
    else:
        raise ValueError('invalid params of type: %s' % type(params))
    if not comm.is_distributed():
        return
    by_dtype: Dict[torch.dtype, List[torch.Tensor]] = {}
    for _, t in items:
        if torch.is_tensor(t):
            by_dtype.setdefault(t.dtype, []).append(t)
    for dtype, tensors in by_dtype.items():
        flat = torch.cat([t.detach().reshape(-1) for t in tensors]) if len(tensors) > 1 else tensors[0].detach().reshape(-1).clone()
        comm.broadcast_(flat, root_rank)
        off = 0
        with torch.no_grad():
            for t in tensors:
                n = t.numel()
                t.copy_(flat[off:off + n].view(t.shape))
                off += n


def broadcast_optimizer_state(optimizer, root_rank=0):
    """Broadcast the optimizer state and hyper-parameters (distributed_optimizer.py:620-736)."""
    if not comm.is_distributed():
        return
    sd = optimizer.state_dict() if rank() == root_rank else None
    meta = comm.bcast_object({"param_groups": sd["param_groups"]} if sd is not None else None, root_rank)
    for g, m in zip(optimizer.param_groups, meta["param_groups"]):
        for k, v in m.items():
            if k != "params":
                g[k] = v
    tensors = []
    for g in optimizer.param_groups:
        for p in g["params"]:
            for k, v in sorted(optimizer.state.get(p, {}).items()):
                if torch.is_tensor(v):
                    tensors.append(v)
    if tensors:
        broadcast_parameters([(None, t) for t in tensors], root_rank)
    # the compressor residuals are deliberately NOT broadcast: they are each rank's own error-feedback memory
