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

from ..ops.bn import bn_act


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


def _bottleneck_forward(self, x):
    s = _downsample(self, x)
    y = bn_act(self.bn1, self.conv1(x))
    y = bn_act(self.bn2, self.conv2(y))
    return bn_act(self.bn3, self.conv3(y), residual=s)


def _basic_forward(self, x):
    s = _downsample(self, x)
    y = bn_act(self.bn1, self.conv1(x))
    return bn_act(self.bn2, self.conv2(y), residual=s)


def _resnet_forward(self, x):
    x = self.maxpool(bn_act(self.bn1, self.conv1(x)))
    x = self.layer4(self.layer3(self.layer2(self.layer1(x))))
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
            for i in range(len(m) - 1):
                if type(m[i]) is nn.BatchNorm2d and type(m[i + 1]) is nn.ReLU:
                    m[i].__class__ = FusedBNReLU
                    m[i + 1] = nn.Identity()
                    sites += 1
    return sites
