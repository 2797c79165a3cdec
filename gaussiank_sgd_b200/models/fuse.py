"""Route the BN -> (+identity) -> ReLU sites of a model through the fused kernels (ops/bn.py).

The models of this package call ``bn_act`` directly.  Models that come from torchvision (the reference
trains ``torchvision.models.resnet50``, dl_trainer.py:90) and ``nn.Sequential`` stacks (VGG-BN, the
Caffe CIFAR nets) are rewired here without touching their parameters, buffers or state_dict keys:

* torchvision ``Bottleneck`` / ``BasicBlock`` / ``ResNet``: the instance gets a forward that computes
  the same expression with ``bn_act``;
* ``[..., BatchNorm2d, ReLU, ...]`` inside a Sequential: the BatchNorm2d instance becomes a
  ``FusedBNReLU`` (class swap, same tensors) and the ReLU slot becomes ``nn.Identity``.
"""
from __future__ import annotations

import types

import torch
import torch.nn as nn
import torch.nn.functional as F

from .. import settings
from ..ops import stem
from ..ops.bn import bn2_act, bn_act, bn_relu_maxpool


class FusedBNReLU(nn.BatchNorm2d):
    """BatchNorm2d whose forward also applies the ReLU that followed it."""

    def forward(self, x):
        return bn_act(self, x, relu=True)


def _split(x):
    """A block input is either one tensor or (convolution input, identity shortcut): the same activation
    handed out twice by the previous block so that its two gradients are not pre-added by autograd."""
    return x if isinstance(x, tuple) else (x, x)


def _plain_shortcut(mod):
    ds = mod.downsample
    return isinstance(ds, nn.Sequential) and len(ds) == 2 and type(ds[1]) is nn.BatchNorm2d


def _bottleneck_forward(self, x):
    xc, xi = _split(x)
    y = bn_act(self.bn1, self.conv1(xc))
    y = bn_act(self.bn2, self.conv2(y))
    twice = getattr(self, "_gk_twice", False)
    if self.downsample is not None and _plain_shortcut(self):      # both BNs of the junction in one kernel
        return bn2_act(self.bn3, self.conv3(y), self.downsample[1], self.downsample[0](xi), twice=twice)
    s = xi if self.downsample is None else self.downsample(xi)
    return bn_act(self.bn3, self.conv3(y), residual=s, twice=twice)


def _basic_forward(self, x):
    xc, xi = _split(x)
    y = bn_act(self.bn1, self.conv1(xc))
    twice = getattr(self, "_gk_twice", False)
    if self.downsample is not None and _plain_shortcut(self):
        return bn2_act(self.bn2, self.conv2(y), self.downsample[1], self.downsample[0](xi), twice=twice)
    s = xi if self.downsample is None else self.downsample(xi)
    return bn_act(self.bn2, self.conv2(y), residual=s, twice=twice)


def _is_stem_pool(m):
    def two(v):
        return tuple(v) if isinstance(v, (tuple, list)) else (v, v)
    return (type(m) is nn.MaxPool2d and two(m.kernel_size) == (3, 3) and two(m.stride) == (2, 2)
            and two(m.padding) == (1, 1) and two(m.dilation) == (1, 1) and not m.ceil_mode and not m.return_indices)


def _stem_conv_s2d(conv, x):
    """7x7 / stride 2 / padding 3 on 3 channels == 4x4 / stride 1 on the 2x2 space-to-depth image (12 channels)
    with one zero tap in front of the kernel:  y[o] = sum_k w[k] x[2o+k-3] = sum_{a<4,b<2} w'[2a+b] X[o+a-2][b],
    w'[0] = 0, w'[k+1] = w[k], X[i][b] = x[2i+b]  (padding 2 before / 1 after).  Same arithmetic, but a shape cuDNN
    has aligned tensor-core kernels for (the 3-channel form runs pre-Hopper indexed kernels: 162 + 176 us at
    batch 32 on a B200).  The parameter keeps its [C_out, 3, 7, 7] shape; autograd carries the gradient back
    through the rearrangement."""
    N, _, H, W = x.shape
    xs = x.permute(0, 2, 3, 1).reshape(N, H // 2, 2, W // 2, 2, 3).permute(0, 1, 3, 2, 4, 5).reshape(N, H // 2, W // 2, 12)
    X = F.pad(xs, (0, 0, 2, 1, 2, 1)).permute(0, 3, 1, 2)          # logical NCHW over NHWC memory
    O = conv.out_channels
    w4 = F.pad(conv.weight, (1, 0, 1, 0)).reshape(O, 3, 4, 2, 4, 2).permute(0, 3, 5, 1, 2, 4).reshape(O, 12, 4, 4)
    return F.conv2d(X, w4.contiguous(memory_format=torch.channels_last), None, 1, 0)


def _stem_conv(conv, x, want_stats=False):
    """The stem convolution in a shape cuDNN has aligned kernels for: the space-to-depth rewrite when the
    convolution is the 7x7 / stride 2 / padding 3 one on an even-sized image (_stem_conv_s2d), else the 3 input
    channels padded to 4 (zeros in the image and in the weight: same result), else the module as it is.  With
    C_in = 3 the NHWC rows are 12 bytes and cuDNN falls back to its pre-Hopper indexed kernels.  The parameter
    keeps its [C_out, 3, k, k] shape in every case.  The tensor-core path (ops/stem.py) also reads a plain NCHW batch;
    the cuDNN paths want channels_last, so an NCHW batch that does not take it is converted here."""
    s2d_shape = (settings.STEM_S2D and x.is_cuda and x.dim() == 4 and x.shape[1] == 3 and conv.in_channels == 3
                 and conv.groups == 1 and conv.bias is None and conv.padding_mode == 'zeros'
                 and tuple(conv.kernel_size) == (7, 7) and tuple(conv.stride) == (2, 2) and tuple(conv.padding) == (3, 3)
                 and tuple(conv.dilation) == (1, 1) and x.shape[2] % 2 == 0 and x.shape[3] % 2 == 0)
    if s2d_shape and settings.STEM_TC and conv.out_channels == 64 and stem.supported(x, conv.weight):
        # TMA + tcgen05 implicit GEMM (csrc/gk_stem.cu); with want_stats it returns (y, sums for the BN)
        return stem.stem_conv_tc(x, conv.weight, want_stats and settings.STEM_BN_STATS)
    if (settings.CHANNELS_LAST and x.is_cuda and x.dim() == 4
            and not x.is_contiguous(memory_format=torch.channels_last)):
        x = x.contiguous(memory_format=torch.channels_last)      # channels_last model, NCHW batch, no tensor-core stem
    if (settings.STEM_S2D and x.is_cuda and x.dim() == 4 and x.shape[1] == 3 and conv.in_channels == 3
            and conv.groups == 1 and conv.bias is None and conv.padding_mode == 'zeros'
            and tuple(conv.kernel_size) == (7, 7) and tuple(conv.stride) == (2, 2) and tuple(conv.padding) == (3, 3)
            and tuple(conv.dilation) == (1, 1) and x.shape[2] % 2 == 0 and x.shape[3] % 2 == 0
            and x.is_contiguous(memory_format=torch.channels_last)):
        return _stem_conv_s2d(conv, x)
    if not (settings.STEM_PAD4 and x.is_cuda and x.dim() == 4 and x.shape[1] == 3 and conv.in_channels == 3
            and conv.groups == 1 and conv.bias is None and conv.padding_mode == 'zeros'
            and x.is_contiguous(memory_format=torch.channels_last)):
        return conv(x)
    N, _, H, W = x.shape
    x4 = torch.zeros((N, 4, H, W), dtype=x.dtype, device=x.device).contiguous(memory_format=torch.channels_last)
    x4[:, :3] = x
    w4 = F.pad(conv.weight, (0, 0, 0, 0, 0, 1)).contiguous(memory_format=torch.channels_last)
    return F.conv2d(x4, w4, None, conv.stride, conv.padding, conv.dilation, 1)


def _resnet_forward(self, x):
    if _is_stem_pool(self.maxpool):
        y = _stem_conv(self.conv1, x, want_stats=self.bn1.training)
        y, sums = y if isinstance(y, tuple) else (y, None)          # batch statistics from the convolution's epilogue
        x = bn_relu_maxpool(self.bn1, y, self.maxpool, presums=sums)
    else:
        x = self.maxpool(bn_act(self.bn1, _stem_conv(self.conv1, x)))
    x = self.layer4(self.layer3(self.layer2(self.layer1(x))))
    if isinstance(x, tuple):
        x = x[0]
    return self.fc(torch.flatten(self.avgpool(x), 1))


def fuse_bn_act(model: nn.Module) -> int:
    """Rewire ``model`` in place; returns the number of BatchNorm sites now on the fused path."""
    sites = 0
    try:
        from torchvision.models import resnet as tv
    except Exception:                       # torchvision is optional
        tv = None
    for m in model.modules():
        if tv is not None and type(m) is tv.Bottleneck and all(type(b) is nn.BatchNorm2d for b in (m.bn1, m.bn2, m.bn3)):
            m.forward = types.MethodType(_bottleneck_forward, m)
            sites += 3 + (m.downsample is not None)
        elif tv is not None and type(m) is tv.BasicBlock and all(type(b) is nn.BatchNorm2d for b in (m.bn1, m.bn2)):
            m.forward = types.MethodType(_basic_forward, m)
            sites += 2 + (m.downsample is not None)
        elif tv is not None and type(m) is tv.ResNet and type(m.bn1) is nn.BatchNorm2d:
            m.forward = types.MethodType(_resnet_forward, m)
            m._gk_stem_reads_nchw = True            # the stem takes the loader's NCHW batch (see _stem_conv)
            sites += 1
            # the last block of a stage feeds the next stage's first convolution and its downsample convolution
            stages = [m.layer1, m.layer2, m.layer3, m.layer4]
            for la, lb in zip(stages, stages[1:]):
                if (settings.FUSED_BN_TWICE and len(la) and len(lb)
                        and type(la[-1]) in (tv.Bottleneck, tv.BasicBlock) and type(lb[0]) in (tv.Bottleneck, tv.BasicBlock)):
                    la[-1]._gk_stage_twice = True
        elif isinstance(m, nn.Sequential):
            blocks = list(m)
            if tv is not None and blocks and all(type(b) in (tv.Bottleneck, tv.BasicBlock) for b in blocks):
                for b, nxt in zip(blocks, blocks[1:] + [None]):      # a ResNet stage
                    b._gk_twice = bool(settings.FUSED_BN_TWICE and ((nxt is not None and nxt.downsample is None)
                                                                      or getattr(b, "_gk_stage_twice", False)))
            for i in range(len(m) - 1):
                if type(m[i]) is nn.BatchNorm2d and type(m[i + 1]) is nn.ReLU:
                    m[i].__class__ = FusedBNReLU
                    m[i + 1] = nn.Identity()
                    sites += 1
    return sites
