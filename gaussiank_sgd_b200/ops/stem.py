"""ResNet stem convolution on the sm_100a tensor cores (csrc/gk_stem.cu: TMA + tcgen05.mma kind::tf32 + TMEM).

The reference trains torchvision's resnet50 (dl_trainer.py:187-196); its first layer, the 7x7 / stride 2 convolution
on 3 channels, is the one convolution cuDNN has no Blackwell kernel for (it runs a pre-Hopper kernel plus a layout
conversion).  ``stem_conv_tc`` computes it as a 4x4 convolution on the padded 2x2 space-to-depth image: one kernel
writes that image, one packs the weight into operand order, the implicit-GEMM kernel does the rest.  The weight
gradient is a second tcgen05 kernel (reduction over the pixels, both operands MN-major) plus a fold of the per-CTA
partial sums; shapes it does not cover go to cuDNN (``aten.convolution_backward`` on the same 4x4 form).  The input
is the image and needs no gradient.
"""
from __future__ import annotations

import torch

from .. import settings
from . import chain as _chain
from . import native

_err = {}
calls = {"fwd": 0, "wgrad": 0}                                   # launches of the tensor-core kernel (tests assert on it)


def _err_word(device):
    t = _err.get(device)
    if t is None:
        t = _err[device] = torch.zeros(1, dtype=torch.int32, device=device)
    return t


def error_codes():
    """Wait-timeout codes recorded by the kernel, per device (all zero in a healthy run)."""
    return {str(d): int(t.item()) for d, t in _err.items()}


_driver = {}


def _driver_ok(device) -> bool:
    """Once per device: can this driver encode the tensor maps of csrc/gk_stem.cu?  (If not, cuDNN keeps the stem.)"""
    ok = _driver.get(device)
    if ok is None:
        ok = _driver[device] = bool(native.cuda_lib().gk_stem_driver_ok(_err_word(device).data_ptr()))
        if not ok:
            settings.logger.warning("stem tensor-core kernels disabled: the driver cannot encode their tensor maps")
    return ok


def supported(x: torch.Tensor, weight: torch.Tensor) -> bool:
    """x: the image batch [N, 3, H, W], channels_last or plain NCHW memory (the space-to-depth kernel reads either, so
    a batch straight from the loader needs no layout conversion); weight: the stem's [64, 3, 7, 7] parameter."""
    return (x.is_cuda and x.dtype == torch.float32 and weight.dtype == torch.float32 and x.dim() == 4
            and x.shape[1] == 3 and tuple(weight.shape) == (64, 3, 7, 7)
            and (x.is_contiguous(memory_format=torch.channels_last) or x.is_contiguous()) and not x.requires_grad
            and not torch.is_autocast_enabled()
            and torch.backends.cudnn.allow_tf32        # the kernel is tf32: honour the switch that forbids it
            and x.shape[0] * (x.shape[2] // 2 + 3) * (x.shape[3] // 2 + 3) < 2 ** 31 - 2048
            and x.shape[0] * (x.shape[2] // 2) * (x.shape[3] // 2) * 256 < 2 ** 32 - 1   # 32-bit byte offsets of y
            and bool(native.cuda_lib().gk_stem_supported(int(x.shape[2]), int(x.shape[3])))
            and _driver_ok(x.device))


class _StemConvTC(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, want_stats=False):
        lib = native.cuda_lib()
        st = torch.cuda.current_stream().cuda_stream
        N, _, H, W = x.shape
        OH, OW = H // 2, W // 2
        X = torch.empty((N, OH + 3, OW + 3, 12), dtype=torch.float32, device=x.device)   # padded space-to-depth image
        nchw = 0 if x.is_contiguous(memory_format=torch.channels_last) else 1
        native.check(lib.gk_stem_s2d_layout(x.data_ptr(), X.data_ptr(), N, H, W, nchw, st), "gk_stem_s2d")
        wpk = torch.empty(24 * 512, dtype=torch.float32, device=x.device)
        wc = weight.detach().contiguous()                                       # channels_last models stride it
        native.check(lib.gk_stem_pack_w(wc.data_ptr(), wpk.data_ptr(), st), "gk_stem_pack_w")
        y = torch.empty((N, OH, OW, 64), dtype=torch.float32, device=x.device)
        sums = (torch.empty((lib.gk_stem_stats_rows(), 128), dtype=torch.float32, device=x.device)
                if want_stats else None)
        native.check(lib.gk_stem_fwd_stats(X.data_ptr(), wpk.data_ptr(), y.data_ptr(),
                                           sums.data_ptr() if want_stats else None, N, OH, OW,
                                           _err_word(x.device).data_ptr(), st), "gk_stem_fwd")
        calls["fwd"] += 1
        _chain.LAUNCHES["n"] += 3                                                # s2d, weight pack, implicit GEMM
        ctx.save_for_backward(X)
        ctx.want_stats = want_stats
        y = y.permute(0, 3, 1, 2)                                               # NCHW over NHWC memory
        if want_stats:
            ctx.mark_non_differentiable(sums)
            return y, sums
        return y

    @staticmethod
    def backward(ctx, gy, *unused):
        if not ctx.needs_input_grad[1]:
            return None, None, None
        X, = ctx.saved_tensors
        gy = gy.contiguous(memory_format=torch.channels_last)
        N, OH, OW = gy.shape[0], gy.shape[2], gy.shape[3]
        lib = native.cuda_lib()
        if settings.STEM_TC_WGRAD and gy.dtype == torch.float32 and lib.gk_stem_wgrad_supported(OH, OW):
            work = torch.empty(lib.gk_stem_wgrad_workspace(), dtype=torch.float32, device=gy.device)
            gw = torch.empty((64, 3, 7, 7), dtype=torch.float32, device=gy.device)
            native.check(lib.gk_stem_wgrad(gy.data_ptr(), X.data_ptr(), gw.data_ptr(), work.data_ptr(), N, OH, OW,
                                           _err_word(gy.device).data_ptr(), torch.cuda.current_stream().cuda_stream),
                         "gk_stem_wgrad")
            calls["wgrad"] += 1
            _chain.LAUNCHES["n"] += 2                                            # pixel-reduction GEMM, fold
            return None, gw, None
        # weight gradient of the 4x4 form from cuDNN (only the weight's shape matters), folded back to 7x7:
        # gw[o, ch, 2a+dy-1, 2b+dx-1] = gw4[o, (dy*2+dx)*3 + ch, a, b]
        w4 = torch.empty((64, 12, 4, 4), dtype=gy.dtype, device=gy.device).contiguous(memory_format=torch.channels_last)
        _, gw4, _ = torch.ops.aten.convolution_backward(
            gy, X.permute(0, 3, 1, 2), w4, None, [1, 1], [0, 0], [1, 1], False, [0, 0], 1, [False, True, False])
        gw = gw4.reshape(64, 2, 2, 3, 4, 4).permute(0, 3, 4, 1, 5, 2).reshape(64, 3, 8, 8)[:, :, 1:, 1:]
        return None, gw.contiguous(), None


def stem_conv_tc(x: torch.Tensor, weight: torch.Tensor, want_stats: bool = False):
    """``conv2d(x, weight, stride=2, padding=3)`` for a 3-channel channels_last image batch and a [64, 3, 7, 7]
    weight; returns [N, 64, H/2, W/2] in channels_last memory.  With ``want_stats`` also the per-channel sums and
    sums of squares of the output ([parts, 128] fp32 partial rows, accumulated in the kernel's epilogue) for
    ``ops.bn.bn_relu_maxpool(..., presums=...)``."""
    return _StemConvTC.apply(x, weight, want_stats)


# ------------------------------------------------------------------------------------------------------------------
# The same decomposition in plain torch (any device, any float dtype): what the kernels of csrc/gk_stem.cu compute,
# step by step.  tests/test_components_cpu.py holds it against F.conv2d, so the index conventions of the
# space-to-depth image, the packed weight and the fold of the weight gradient are pinned on machines without a GPU.
def space_to_depth_reference(x: torch.Tensor) -> torch.Tensor:
    """[N, 3, H, W] -> padded space-to-depth image [N, H/2+3, W/2+3, 12] (k_stem_s2d):
    X[n, i, j, (dy*2+dx)*3 + c] = x[n, c, 2(i-2)+dy, 2(j-2)+dx], zero in the two leading / one trailing rows and columns."""
    N, C, H, W = x.shape
    xs = x.permute(0, 2, 3, 1).reshape(N, H // 2, 2, W // 2, 2, C).permute(0, 1, 3, 2, 4, 5).reshape(N, H // 2, W // 2, 4 * C)
    X = x.new_zeros((N, H // 2 + 3, W // 2 + 3, 4 * C))
    X[:, 2:-1, 2:-1] = xs
    return X


def pack_weight_reference(w: torch.Tensor) -> torch.Tensor:
    """[64, 3, 7, 7] -> [24, 2, 64, 4] (k_stem_pack_w; the kernel stores the 64 rows as 8 groups of 8): step
    m = (a*3 + t)*2 + bp, half kc, row o, float e  <-  W4[o, 4t + e, a, 2bp + kc] with
    W4[o, (dy*2+dx)*3 + ch, a, b] = w[o, ch, 2a+dy-1, 2b+dx-1] (zero where an index is -1)."""
    O = w.shape[0]
    w4 = torch.nn.functional.pad(w, (1, 0, 1, 0)).reshape(O, 3, 4, 2, 4, 2).permute(0, 3, 5, 1, 2, 4).reshape(O, 12, 4, 4)
    # [o, (t, e), a, (bp, kc)] -> [a, t, bp, kc, o, e]
    return w4.reshape(O, 3, 4, 4, 2, 2).permute(3, 1, 4, 5, 0, 2).reshape(24, 2, O, 4)


def forward_reference(x: torch.Tensor, w: torch.Tensor) -> torch.Tensor:
    """conv2d(x, w, stride 2, padding 3) as the 24 K = 8 steps of k_stem_fwd_tc over the flat padded positions."""
    N, _, H, W = x.shape
    OH, OW, Hp, Wp = H // 2, W // 2, H // 2 + 3, W // 2 + 3
    Xf = space_to_depth_reference(x).reshape(N * Hp * Wp, 12)
    Xf = torch.cat([Xf, Xf.new_zeros((3 * Wp + 4, 12))])                 # the tensor map's zero fill past the end
    wpk = pack_weight_reference(w)
    P = N * Hp * Wp
    y = x.new_zeros((P, w.shape[0]))
    for a in range(4):
        for t in range(3):
            for bp in range(2):
                m = (a * 3 + t) * 2 + bp
                for kc in range(2):
                    off = a * Wp + 2 * bp + kc
                    y += Xf[off:off + P, 4 * t:4 * t + 4] @ wpk[m, kc].t()
    return y.reshape(N, Hp, Wp, -1)[:, :OH, :OW].permute(0, 3, 1, 2)


def wgrad_reference(gy: torch.Tensor, x: torch.Tensor) -> torch.Tensor:
    """Weight gradient [64, 3, 7, 7] as k_stem_wgrad_tc + k_stem_wgrad_fold compute it: per kernel row a the 64 x 48
    product of gY (pixels x channels) with the 48-float windows of X, then the fold of the 4x4 form back to 7x7."""
    N, O, OH, OW = gy.shape
    X = space_to_depth_reference(x)
    g = gy.permute(0, 2, 3, 1)                                          # NHWC
    part = gy.new_zeros((O, 192))
    for a in range(4):
        win = torch.cat([X[:, a:a + OH, b:b + OW] for b in range(4)], dim=3)       # [N, OH, OW, 48]: b-major, then c
        part[:, a * 48:(a + 1) * 48] = torch.einsum("nhwo,nhwk->ok", g, win)
    gw = gy.new_zeros((O, 3, 7, 7))
    for u in range(7):
        for v in range(7):
            dy, a, dx, b = (u + 1) & 1, (u + 1) >> 1, (v + 1) & 1, (v + 1) >> 1
            for ch in range(3):
                gw[:, ch, u, v] = part[:, a * 48 + b * 12 + (dy * 2 + dx) * 3 + ch]
    return gw
