"""Fused training-mode BatchNorm2d (+ residual add) (+ ReLU) on channels_last fp32 activations.

The reference's CNNs take ``bn -> (+ identity) -> relu`` from torch.nn as separate modules
(torchvision ResNet-50 through dl_trainer.py:90, models/resnet.py, models/vgg.py ...).  On a B200 those
HBM-bound passes are 40 % of the ResNet-50 step (profiles/step_breakdown_resnet50_v1.txt), so here they
are one cooperative kernel per direction (csrc/gk_bn.cu: ``gk_bn_fwd_fused`` / ``gk_bn_bwd_fused``; two
kernels each, ``gk_bn_fwd`` / ``gk_bn_bwd``, for shapes the cooperative tiling does not cover), plus the
variants for the ResNet stem (``bn_relu_maxpool``) and for the downsample junctions (``bn2_act``).

``bn_act(bn, x, residual=None, relu=True)`` is the entry point: ``bn`` stays an ordinary
``nn.BatchNorm2d`` (same parameters, buffers, state_dict keys); anything the kernels do not cover
(CPU tensors, eval mode, non-fp32, NCHW-contiguous input, C % 4 != 0, momentum=None) takes the plain
PyTorch path, so the numerics contract is "same as F.batch_norm + add + relu".
"""
from __future__ import annotations

import torch
import torch.nn.functional as F
from torch.autograd.function import once_differentiable

from .. import settings
from . import native
from . import chain as _chain


def _ptr(t):
    return None if t is None else t.data_ptr()


def _stream():
    return torch.cuda.current_stream().cuda_stream


def _scratch(bn, device, C):
    """Per-layer accumulators (2*C doubles) and ticket; zeroed once, the kernels leave them zero."""
    sc = bn.__dict__.get("_gk_bn_scratch")
    if sc is None or sc[0].device != device or sc[0].numel() != 2 * C:
        sc = (torch.zeros(2 * C, dtype=torch.float64, device=device),
              torch.zeros(1, dtype=torch.int32, device=device))
        bn.__dict__["_gk_bn_scratch"] = sc
    return sc


ERR_WORD, STAMP_WORD = 127, 96
_COOP_OFF = {"off": False}
_COOP_LAUNCH_ERRORS = (720, 701)   # cudaErrorCooperativeLaunchTooLarge, cudaErrorLaunchOutOfResources


def _scratch_coop(bn, device, C):
    """Scratch of the single-kernel path, one per layer: 4*C doubles and 128 barrier words, zero at
    creation and never reset (include/gk_b200.h)."""
    table = bn.__dict__.setdefault("_gk_bn_coop", {})
    key = (str(device), C)
    sc = table.get(key)
    if sc is None:
        sc = (torch.zeros(4 * C, dtype=torch.float64, device=device),
              torch.zeros(128, dtype=torch.int64, device=device))
        table[key] = sc
    return sc


def phase_stamps(bn):
    """[forward, backward] phase boundaries in ns relative to kernel start (needs GK_BN_STAMPS=1):
    start, loads done, sums posted, barrier passed, constants ready, end."""
    out = []
    for _, bar in bn.__dict__.get("_gk_bn_coop", {}).values():
        w = bar[STAMP_WORD:STAMP_WORD + 16].tolist()
        out = [[v - w[0] for v in w[0:6]], [v - w[8] for v in w[8:14]]]
    return out


def coop_errors(model) -> int:
    """Number of layers whose barrier wait ever expired (0 on a healthy run)."""
    bad = 0
    for m in model.modules():
        for _, bar in m.__dict__.get("_gk_bn_coop", {}).values():
            bad += int(bar[ERR_WORD].item() != 0)
    return bad


def _rows_view_ok(t):
    return t.dim() == 4 and t.is_contiguous(memory_format=torch.channels_last) and t.data_ptr() % 16 == 0


def eligible(bn, x, residual=None) -> bool:
    if not (x.is_cuda and x.dtype == torch.float32 and bn.training and bn.affine and x.dim() == 4):
        return False
    if bn.momentum is None or x.shape[1] % 4 or x.numel() == 0 or x.numel() // x.shape[1] < 2:
        return False
    if bn.weight.dtype != torch.float32 or not _rows_view_ok(x):
        return False
    if residual is not None and (residual.shape != x.shape or residual.dtype != x.dtype):
        return False
    return True


class _BnAct(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, residual, weight, bias, bn, relu, twice=False):
        N, C, H, W = x.shape
        M = N * H * W
        if residual is not None and not _rows_view_ok(residual):
            residual = residual.contiguous(memory_format=torch.channels_last)
        y = torch.empty_like(x, memory_format=torch.channels_last)
        stats = torch.empty(2, C, dtype=torch.float32, device=x.device)
        track = bn.track_running_stats and bn.running_mean is not None
        lib = native.cuda_lib()
        ctx.coop = bool(settings.FUSED_BN_COOP and not _COOP_OFF["off"] and lib.gk_bn_fused_supported(M, C))
        head = (_ptr(x), _ptr(residual), _ptr(y), _ptr(weight), _ptr(bias),
                _ptr(bn.running_mean) if track else None, _ptr(bn.running_var) if track else None,
                _ptr(bn.num_batches_tracked) if track else None, stats[0].data_ptr(), stats[1].data_ptr())
        tail = (M, C, float(bn.eps), float(bn.momentum), int(relu) | (2 if settings.BN_STAMPS else 0), _stream())
        mask = None
        if ctx.coop:
            accum, bar = ctx.scratch = _scratch_coop(bn, x.device, C)
            if relu and residual is not None and settings.FUSED_BN_MASK:
                # the backward needs to know where the ReLU was active: one byte per channel quad instead of y
                mask = torch.empty(M * (C // 4), dtype=torch.uint8, device=x.device)
            rc = lib.gk_bn_fwd_fused_m(*head, accum.data_ptr(), bar.data_ptr(), *tail[:-1], _ptr(mask), tail[-1])
            if rc in _COOP_LAUNCH_ERRORS:
                # the device cannot keep one block per SM resident (SM share limited by MPS, ...): use the
                # two-kernel variant from now on
                _COOP_OFF["off"] = True
                ctx.coop, mask = False, None
                settings.logger.warning("cooperative launch refused (cudaError %d): fused BN falls back to two kernels", rc)
            else:
                native.check(rc, "gk_bn_fwd_fused")
                _chain.LAUNCHES["n"] += 1
        if not ctx.coop:
            accum, ticket = ctx.scratch = _scratch(bn, x.device, C)
            native.check(lib.gk_bn_fwd(*head, accum.data_ptr(), ticket.data_ptr(), *tail[:4], int(relu), tail[5]),
                         "gk_bn_fwd")
            _chain.LAUNCHES["n"] += 2
        ctx.relu = bool(relu)
        ctx.has_res = residual is not None
        # the ReLU mask is recomputed from x unless a residual went into the activation
        ctx.has_mask = mask is not None
        ctx.save_for_backward(x, weight, bias, stats,
                              mask if mask is not None else (y if (relu and residual is not None) else None))
        if twice:
            # the same activation handed out twice (convolution input + identity shortcut of the next block):
            # autograd then delivers the two gradients separately and the backward kernel adds them on the
            # fly instead of autograd running an add kernel over the whole activation first
            ctx.set_materialize_grads(False)
            return y, y.view_as(y)
        return y

    @staticmethod
    @once_differentiable
    def backward(ctx, dy, dy2=None):
        x, weight, bias, stats, y = ctx.saved_tensors
        N, C, H, W = x.shape
        M = N * H * W
        if dy is None:
            dy, dy2 = dy2, None
        if dy is None:
            return None, None, None, None, None, None, None
        if not _rows_view_ok(dy):
            dy = dy.contiguous(memory_format=torch.channels_last)
        if dy2 is not None and not _rows_view_ok(dy2):
            dy2 = dy2.contiguous(memory_format=torch.channels_last)
        if dy2 is not None and not ctx.coop:
            dy, dy2 = dy + dy2, None
        dx = torch.empty_like(x, memory_format=torch.channels_last)
        dres = torch.empty_like(x, memory_format=torch.channels_last) if (ctx.has_res and ctx.relu) else None
        # separate tensors: autograd hands a fresh dense tensor to .grad without a copy, a view it would clone
        dgamma = torch.empty(C, dtype=torch.float32, device=x.device)
        dbeta = torch.empty(C, dtype=torch.float32, device=x.device)
        lib = native.cuda_lib()
        head = (_ptr(dy), _ptr(x), _ptr(y), _ptr(weight), _ptr(bias), stats[0].data_ptr(), stats[1].data_ptr(),
                _ptr(dx), _ptr(dres), dgamma.data_ptr(), dbeta.data_ptr())
        if ctx.coop:
            accum, bar = ctx.scratch
            native.check(lib.gk_bn_bwd_fused2(head[0], _ptr(dy2), *head[1:], accum.data_ptr(), bar.data_ptr(), M, C,
                                              int(ctx.relu) | (2 if settings.BN_STAMPS else 0)
                                              | (4 if ctx.has_mask else 0), _stream()),
                         "gk_bn_bwd_fused2")
            _chain.LAUNCHES["n"] += 1
        else:
            accum, ticket = ctx.scratch
            coef = torch.empty(2, C, dtype=torch.float32, device=x.device)
            native.check(lib.gk_bn_bwd(*head, coef.data_ptr(), accum.data_ptr(), ticket.data_ptr(), M, C,
                                       int(ctx.relu), _stream()), "gk_bn_bwd")
            _chain.LAUNCHES["n"] += 2
        if ctx.has_res and dres is None:      # no ReLU: the residual branch receives dy unchanged
            dres = dy if dy2 is None else dy + dy2
        return dx, (dres if ctx.has_res else None), dgamma, dbeta, None, None, None


class _BnReluPool(torch.autograd.Function):
    """relu(bn(x)) followed by max_pool2d(3, stride 2, padding 1): the full-resolution activation is never
    written; the backward gathers the pooled gradient through the saved window positions."""

    @staticmethod
    def forward(ctx, x, weight, bias, bn, presums=None):
        N, C, H, W = x.shape
        OH, OW = (H - 1) // 2 + 1, (W - 1) // 2 + 1
        yp = torch.empty((N, C, OH, OW), dtype=x.dtype, device=x.device, memory_format=torch.channels_last)
        pidx = torch.empty((N, OH, OW, C), dtype=torch.uint8, device=x.device)
        stats = torch.empty(2, C, dtype=torch.float32, device=x.device)
        track = bn.track_running_stats and bn.running_mean is not None
        accum, bar = ctx.scratch = _scratch_coop(bn, x.device, C)
        lib = native.cuda_lib()
        if presums is not None:       # per-CTA sums / sums of squares from the producer of x (ops/stem.py): no statistics pass
            native.check(lib.gk_bn_pool_fwd_sums(_ptr(x), _ptr(yp), _ptr(pidx), _ptr(weight), _ptr(bias),
                                                 _ptr(bn.running_mean) if track else None,
                                                 _ptr(bn.running_var) if track else None,
                                                 _ptr(bn.num_batches_tracked) if track else None,
                                                 stats[0].data_ptr(), stats[1].data_ptr(), presums.data_ptr(),
                                                 presums.shape[0], N, H, W, C, float(bn.eps), float(bn.momentum),
                                                 _stream()), "gk_bn_pool_fwd_sums")
            _chain.LAUNCHES["n"] += 2
            ctx.save_for_backward(x, weight, bias, stats, pidx)
            return yp
        native.check(lib.gk_bn_pool_fwd(_ptr(x), _ptr(yp), _ptr(pidx), _ptr(weight), _ptr(bias),
                                        _ptr(bn.running_mean) if track else None,
                                        _ptr(bn.running_var) if track else None,
                                        _ptr(bn.num_batches_tracked) if track else None,
                                        stats[0].data_ptr(), stats[1].data_ptr(), accum.data_ptr(), None,
                                        bar.data_ptr(), N, H, W, C, float(bn.eps), float(bn.momentum), _stream()),
                     "gk_bn_pool_fwd")
        _chain.LAUNCHES["n"] += 2
        ctx.save_for_backward(x, weight, bias, stats, pidx)
        return yp

    @staticmethod
    @once_differentiable
    def backward(ctx, dyp):
        x, weight, bias, stats, pidx = ctx.saved_tensors
        N, C, H, W = x.shape
        if not _rows_view_ok(dyp):
            dyp = dyp.contiguous(memory_format=torch.channels_last)
        dx = torch.empty_like(x, memory_format=torch.channels_last)
        dgamma = torch.empty(C, dtype=torch.float32, device=x.device)
        dbeta = torch.empty(C, dtype=torch.float32, device=x.device)
        accum, bar = ctx.scratch
        native.check(native.cuda_lib().gk_bn_pool_bwd(
            _ptr(dyp), _ptr(pidx), _ptr(x), _ptr(weight), _ptr(bias), stats[0].data_ptr(), stats[1].data_ptr(),
            _ptr(dx), _ptr(dx), dgamma.data_ptr(), dbeta.data_ptr(), accum.data_ptr(), bar.data_ptr(), N, H, W, C,
            2 if settings.BN_STAMPS else 0, _stream()), "gk_bn_pool_bwd")     # dx doubles as the gather buffer
        _chain.LAUNCHES["n"] += 2
        return dx, dgamma, dbeta, None, None


def _ptr_table(ptrs):
    arr = (native.C.c_void_p * len(ptrs))(*[p if p else None for p in ptrs])
    return arr


class _Bn2Act(torch.autograd.Function):
    """relu(bn_a(xa) + bn_b(xb)): last BN of a residual block plus the BN of its downsample shortcut."""

    @staticmethod
    def forward(ctx, xa, xb, wa, ba, wb, bb, bna, bnb, relu, twice):
        N, C, H, W = xa.shape
        M = N * H * W
        y = torch.empty_like(xa, memory_format=torch.channels_last)
        stats = torch.empty(4, C, dtype=torch.float32, device=xa.device)
        table = bna.__dict__.setdefault("_gk_bn_coop", {})
        key = (str(xa.device), C, "pair")
        if key not in table:
            table[key] = (torch.zeros(8 * C, dtype=torch.float64, device=xa.device),
                          torch.zeros(128, dtype=torch.int64, device=xa.device))
        accum, bar = ctx.scratch = table[key]

        def side(x, w, b, bn, row):
            track = bn.track_running_stats and bn.running_mean is not None
            return _ptr_table([_ptr(x), _ptr(w), _ptr(b), stats[row].data_ptr(), stats[row + 1].data_ptr(),
                               _ptr(bn.running_mean) if track else None, _ptr(bn.running_var) if track else None,
                               _ptr(bn.num_batches_tracked) if track else None])
        ta, tb = side(xa, wa, ba, bna, 0), side(xb, wb, bb, bnb, 2)
        mask = (torch.empty(M * (C // 4), dtype=torch.uint8, device=xa.device)
                if (relu and settings.FUSED_BN_MASK) else None)
        native.check(native.cuda_lib().gk_bn2_fwd_fused(ta, tb, float(bna.eps), float(bna.momentum), float(bnb.eps),
                                                        float(bnb.momentum), _ptr(y), _ptr(mask), accum.data_ptr(),
                                                        bar.data_ptr(), M, C, int(relu), _stream()), "gk_bn2_fwd_fused")
        _chain.LAUNCHES["n"] += 1
        ctx.relu = bool(relu)
        ctx.has_mask = mask is not None
        ctx.save_for_backward(xa, xb, wa, ba, wb, bb, stats, mask if mask is not None else y)
        if twice:
            ctx.set_materialize_grads(False)
            return y, y.view_as(y)
        return y

    @staticmethod
    @once_differentiable
    def backward(ctx, dy, dy2=None):
        xa, xb, wa, ba, wb, bb, stats, y = ctx.saved_tensors
        N, C, H, W = xa.shape
        M = N * H * W
        if dy is None:
            dy, dy2 = dy2, None
        if dy is None:
            return (None,) * 10
        if not _rows_view_ok(dy):
            dy = dy.contiguous(memory_format=torch.channels_last)
        if dy2 is not None and not _rows_view_ok(dy2):
            dy2 = dy2.contiguous(memory_format=torch.channels_last)
        dxa = torch.empty_like(xa, memory_format=torch.channels_last)
        dxb = torch.empty_like(xb, memory_format=torch.channels_last)
        dga, dba, dgb, dbb = (torch.empty(C, dtype=torch.float32, device=xa.device) for _ in range(4))
        accum, bar = ctx.scratch
        ta = _ptr_table([_ptr(xa), _ptr(wa), _ptr(ba), stats[0].data_ptr(), stats[1].data_ptr(), _ptr(dxa), _ptr(dga),
                         _ptr(dba)])
        tb = _ptr_table([_ptr(xb), _ptr(wb), _ptr(bb), stats[2].data_ptr(), stats[3].data_ptr(), _ptr(dxb), _ptr(dgb),
                         _ptr(dbb)])
        native.check(native.cuda_lib().gk_bn2_bwd_fused(_ptr(dy), _ptr(dy2), _ptr(y), ta, tb, accum.data_ptr(),
                                                        bar.data_ptr(), M, C,
                                                        int(ctx.relu) | (4 if ctx.has_mask else 0), _stream()),
                     "gk_bn2_bwd_fused")
        _chain.LAUNCHES["n"] += 1
        return dxa, dxb, dga, dba, dgb, dbb, None, None, None, None


def bn2_act(bna, xa, bnb, xb, relu=True, twice=False):
    """``relu(bna(xa) + bnb(xb))`` — e.g. ``bn3(conv3(.)) + downsample_bn(downsample_conv(.))`` of a ResNet block."""
    if (settings.FUSED_BN and settings.FUSED_BN_PAIR and settings.FUSED_BN_COOP and not _COOP_OFF["off"]
            and xa.shape == xb.shape
            and eligible(bna, xa) and eligible(bnb, xb)
            and native.cuda_lib().gk_bn_fused_supported(xa.numel() // xa.shape[1], xa.shape[1])):
        return _Bn2Act.apply(xa, xb, bna.weight, bna.bias, bnb.weight, bnb.bias, bna, bnb, relu, twice)
    return bn_act(bna, xa, residual=bn_act(bnb, xb, relu=False), relu=relu, twice=twice)


def bn_act(bn, x, residual=None, relu=True, twice=False):
    """``relu(bn(x) + residual)`` with ``residual`` and ``relu`` optional; ``bn`` is an nn.BatchNorm2d.
    ``twice=True`` returns the result as two tensors sharing one storage (see _BnAct.forward)."""
    if settings.FUSED_BN and eligible(bn, x, residual):
        return _BnAct.apply(x, residual, bn.weight, bn.bias, bn, relu, twice)
    y = _bn_act_plain(bn, x, residual, relu)
    return (y, y) if twice else y


def bn_relu_maxpool(bn, x, pool=None, presums=None):
    """``max_pool2d(relu(bn(x)), 3, 2, 1)`` (the ResNet stem) through the fused kernels when possible.  ``presums``
    ([parts, 2C] fp32: per-channel sums and sums of squares of x, in partial rows) replaces the statistics pass."""
    N, C, H, W = x.shape if x.dim() == 4 else (0, 0, 0, 0)
    if (settings.FUSED_BN and settings.FUSED_BN_POOL and settings.FUSED_BN_COOP and not _COOP_OFF["off"]
            and eligible(bn, x) and H >= 2 and W >= 2
            and N * H * W < 2 ** 31 and native.cuda_lib().gk_bn_fused_supported(N * H * W, C)):
        if presums is not None and not (bn.training and presums.shape[1] == 2 * C and presums.is_cuda):
            presums = None
        return _BnReluPool.apply(x, bn.weight, bn.bias, bn, presums)
    y = bn_act(bn, x, relu=True)
    return pool(y) if pool is not None else F.max_pool2d(y, 3, 2, 1)


def _bn_act_plain(bn, x, residual, relu):
    """The PyTorch sequence; BatchNorm2d.forward directly because ``bn`` may be a FusedBNReLU (models/fuse.py)."""
    y = torch.nn.BatchNorm2d.forward(bn, x) if isinstance(bn, torch.nn.BatchNorm2d) else bn(x)
    if residual is not None:
        y = y + residual
    return F.relu(y, inplace=True) if relu else y
