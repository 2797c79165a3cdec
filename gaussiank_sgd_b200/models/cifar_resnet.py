"""CIFAR ResNets (He et al. 2015, section 4.2): 6n+2 layers, 16/32/64 channels, parameter-free
"option A" shortcuts.  Same architecture and parameter counts as the reference's
models/resnet.py:40-147 + models/res_utils.py:4-13 (ResNet-20: 269 722 parameters in 59 tensors).
Also the pre-activation variant (models/preresnet.py) and the feature-returning variant
(models/resnet_mod.py) since they only differ in block wiring."""
from __future__ import annotations

import math

import torch
import torch.nn as nn
import torch.nn.functional as F
from ..ops.bn import bn_act


class PadShortcut(nn.Module):
    """Option-A shortcut: spatial subsampling by striding + zero channels appended (no parameters)."""

    def __init__(self, stride: int, extra_channels: int):
        super().__init__()
        self.stride, self.extra = stride, extra_channels

    def forward(self, x):
        x = x[:, :, ::self.stride, ::self.stride]
        return F.pad(x, (0, 0, 0, 0, 0, self.extra))


class BasicBlock(nn.Module):
    expansion = 1

    def __init__(self, cin, cout, stride=1, shortcut=None):
        super().__init__()
        self.conv_a = nn.Conv2d(cin, cout, 3, stride, 1, bias=False)
        self.bn_a = nn.BatchNorm2d(cout)
        self.conv_b = nn.Conv2d(cout, cout, 3, 1, 1, bias=False)
        self.bn_b = nn.BatchNorm2d(cout)
        self.downsample = shortcut

    def forward(self, x):
        y = bn_act(self.bn_a, self.conv_a(x))
        s = x if self.downsample is None else self.downsample(x)
        return bn_act(self.bn_b, self.conv_b(y), residual=s)


class PreActBlock(nn.Module):
    expansion = 1

    def __init__(self, cin, cout, stride=1, shortcut=None):
        super().__init__()
        self.bn_a = nn.BatchNorm2d(cin)
        self.conv_a = nn.Conv2d(cin, cout, 3, stride, 1, bias=False)
        self.bn_b = nn.BatchNorm2d(cout)
        self.conv_b = nn.Conv2d(cout, cout, 3, 1, 1, bias=False)
        self.downsample = shortcut

    def forward(self, x):
        a = bn_act(self.bn_a, x)
        s = x if self.downsample is None else self.downsample(a)
        y = self.conv_a(a)
        y = self.conv_b(bn_act(self.bn_b, y))
        return s + y


class CifarResNet(nn.Module):
    def __init__(self, depth: int, num_classes: int = 10, preact: bool = False, return_features: bool = False):
        super().__init__()
        if (depth - 2) % 6:
            raise ValueError('depth should be one of 20, 32, 44, 56, 110')
        n = (depth - 2) // 6
        self.name = ('preresnet%d' if preact else 'resnet%d') % depth
        self.num_classes = num_classes
        self.return_features = return_features
        block = PreActBlock if preact else BasicBlock
        self.conv_1_3x3 = nn.Conv2d(3, 16, 3, 1, 1, bias=False)
        self.bn_1 = nn.BatchNorm2d(16)
        widths, cin, stages = (16, 32, 64), 16, []
        for si, w in enumerate(widths):
            blocks = []
            for bi in range(n):
                stride = 2 if (si > 0 and bi == 0) else 1
                sc = PadShortcut(stride, w - cin) if (stride != 1 or cin != w) else None
                blocks.append(block(cin, w, stride, sc))
                cin = w
            stages.append(nn.Sequential(*blocks))
        self.stage_1, self.stage_2, self.stage_3 = stages
        self.final_bn = nn.BatchNorm2d(64) if preact else None
        self.avgpool = nn.AvgPool2d(8)
        self.classifier = nn.Linear(64, num_classes)
        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                fan = m.kernel_size[0] * m.kernel_size[1] * m.out_channels
                m.weight.data.normal_(0, math.sqrt(2.0 / fan))
            elif isinstance(m, nn.BatchNorm2d):
                m.weight.data.fill_(1)
                m.bias.data.zero_()
            elif isinstance(m, nn.Linear):
                nn.init.kaiming_normal_(m.weight)
                m.bias.data.zero_()

    def forward(self, x):
        x = self.conv_1_3x3(x)
        if self.final_bn is None:
            x = bn_act(self.bn_1, x)
        f1 = self.stage_1(x)
        f2 = self.stage_2(f1)
        f3 = self.stage_3(f2)
        y = f3 if self.final_bn is None else bn_act(self.final_bn, f3)
        y = self.avgpool(y).flatten(1)
        out = self.classifier(y)
        if self.return_features:
            return out, [f1, f2, f3]
        return out


def _mk(depth, **kw):
    def build(num_classes=10):
        return CifarResNet(depth, num_classes, **kw)
    return build


resnet20, resnet32, resnet44, resnet56, resnet110 = (_mk(d) for d in (20, 32, 44, 56, 110))
preresnet20, preresnet32, preresnet44, preresnet56, preresnet110 = (_mk(d, preact=True) for d in (20, 32, 44, 56, 110))
resnet_mod20, resnet_mod32, resnet_mod44, resnet_mod56, resnet_mod110 = (
    _mk(d, return_features=True) for d in (20, 32, 44, 56, 110))
