"""ResNet stem convolution on the sm_100a tensor cores (csrc/gk_stem.cu: TMA + tcgen05.mma kind::tf32 + TMEM).

The reference trains torchvision's resnet50 (dl_trainer.py:187-196); its first layer, the 7x7 / stride 2 convolution
on 3 channels, is the one convolution cuDNN has no Blackwell kernel for (it runs a pre-Hopper kernel plus a layout
conversion).  ``stem_conv_tc`` computes it as a 4x4 convolution on the padded 2x2 space-to-depth image with the
implicit-GEMM kernel; the weight gradient still comes from cuDNN (``aten.convolution_backward`` on the same 4x4 form),
the input is the image and needs no gradient.
"""
from __future__ import annotations

import torch

from . import native

_err = {}
calls = {"fwd": 0}                                   # launches of the tensor-core kernel (tests assert on it)


def _err_word(device):
    t = _err.get(device)
    if t is None:
        t = _err[device] = torch.zeros(1, dtype=torch.int32, device=device)
    return t


def error_codes():
    """Wait-timeout codes recorded by the kernel, per device (all zero in a healthy run)."""
    return {str(d): int(t.item()) for d, t in _err.items()}


def supported(X: torch.Tensor, w4: torch.Tensor) -> bool:
    return (X.is_cuda and X.dtype == torch.float32 and w4.dtype == torch.float32 and X.dim() == 4 and X.shape[3] == 12
            and X.is_contiguous() and tuple(w4.shape) == (64, 12, 4, 4) and not torch.is_autocast_enabled()
            and X.shape[0] * X.shape[1] * X.shape[2] < 2 ** 31 - 128)


class _StemConvTC(torch.autograd.Function):
    @staticmethod
    def forward(ctx, X, w4):
        # X [N, Hp, Wp, 12] padded space-to-depth image (NHWC), w4 [64, 12, 4, 4]
        N, Hp, Wp, _ = X.shape
        OH, OW = Hp - 3, Wp - 3
        wk = w4.detach().permute(0, 2, 3, 1).reshape(64, 192).contiguous()      # [o][a*48 + b*12 + c]
        y = torch.empty((N, OH, OW, 64), dtype=torch.float32, device=X.device)
        rc = native.cuda_lib().gk_stem_fwd(X.data_ptr(), wk.data_ptr(), y.data_ptr(), N, OH, OW,
                                           _err_word(X.device).data_ptr(), torch.cuda.current_stream().cuda_stream)
        native.check(rc, "gk_stem_fwd")
        calls["fwd"] += 1
        ctx.save_for_backward(X, w4)
        return y.permute(0, 3, 1, 2)                                            # NCHW over NHWC memory

    @staticmethod
    def backward(ctx, gy):
        X, w4 = ctx.saved_tensors
        if not ctx.needs_input_grad[1]:
            return None, None
        gy = gy.contiguous(memory_format=torch.channels_last)
        _, gw, _ = torch.ops.aten.convolution_backward(
            gy, X.permute(0, 3, 1, 2), w4.contiguous(memory_format=torch.channels_last), None,
            [1, 1], [0, 0], [1, 1], False, [0, 0], 1, [False, True, False])
        return None, gw


def stem_conv_tc(X: torch.Tensor, w4: torch.Tensor) -> torch.Tensor:
    """``conv2d(X.permute(0, 3, 1, 2), w4)`` for the padded space-to-depth image X [N, Hp, Wp, 12] and the 4x4 form
    w4 [64, 12, 4, 4] of the stem weight; returns [N, 64, Hp-3, Wp-3] in channels_last memory."""
    return _StemConvTC.apply(X, w4)
