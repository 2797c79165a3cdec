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

from .. import settings
from ..ops.bn import bn_act, bn_relu_maxpool


class FusedBNReLU(nn.BatchNorm2d):
    """BatchNorm2d whose forward also applies the ReLU that followed it."""

    def forward(self, x):
        return bn_act(self, x, relu=True)


def _downsample(mod, x):
    ds = mod.downsample
    if ds is None:
        return x
    if isinstance(ds, nn.Sequential) and len(ds) == 2 and type(ds[1]) is nn.BatchNorm2d:
        return bn_act(ds[1], ds[0](x), relu=False)
    return ds(x)


def _split(x):
    """A block input is either one tensor or (convolution input, identity shortcut): the same activation
    handed out twice by the previous block so that its two gradients are not pre-added by autograd."""
    return x if isinstance(x, tuple) else (x, x)


def _bottleneck_forward(self, x):
    xc, xi = _split(x)
    s = xi if self.downsample is None else _downsample(self, xc)
    y = bn_act(self.bn1, self.conv1(xc))
    y = bn_act(self.bn2, self.conv2(y))
    return bn_act(self.bn3, self.conv3(y), residual=s, twice=getattr(self, "_gk_twice", False))


def _basic_forward(self, x):
    xc, xi = _split(x)
    s = xi if self.downsample is None else _downsample(self, xc)
    y = bn_act(self.bn1, self.conv1(xc))
    return bn_act(self.bn2, self.conv2(y), residual=s, twice=getattr(self, "_gk_twice", False))


def _is_stem_pool(m):
    def two(v):
        return tuple(v) if isinstance(v, (tuple, list)) else (v, v)
    return (type(m) is nn.MaxPool2d and two(m.kernel_size) == (3, 3) and two(m.stride) == (2, 2)
            and two(m.padding) == (1, 1) and two(m.dilation) == (1, 1) and not m.ceil_mode and not m.return_indices)


def _resnet_forward(self, x):
    if _is_stem_pool(self.maxpool):
        x = bn_relu_maxpool(self.bn1, self.conv1(x), self.maxpool)
    else:
        x = self.maxpool(bn_act(self.bn1, self.conv1(x)))
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
            sites += 1
        elif isinstance(m, nn.Sequential):
            blocks = list(m)
            if tv is not None and blocks and all(type(b) in (tv.Bottleneck, tv.BasicBlock) for b in blocks):
                for b, nxt in zip(blocks, blocks[1:] + [None]):      # a ResNet stage
                    b._gk_twice = bool(settings.FUSED_BN_TWICE and nxt is not None and nxt.downsample is None)
            for i in range(len(m) - 1):
                if type(m[i]) is nn.BatchNorm2d and type(m[i + 1]) is nn.ReLU:
                    m[i].__class__ = FusedBNReLU
                    m[i + 1] = nn.Identity()
                    sites += 1
    return sites
