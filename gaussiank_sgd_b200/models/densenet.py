"""DenseNet-BC-100 (k=12) for CIFAR (Huang et al. 2016; models/densenet.py in the reference)."""
import math

import torch
import torch.nn as nn
import torch.nn.functional as F
from ..ops.bn import bn_act


class DenseLayer(nn.Module):
    def __init__(self, cin, growth, bottleneck=True):
        super().__init__()
        inter = 4 * growth
        self.bn1 = nn.BatchNorm2d(cin)
        self.conv1 = nn.Conv2d(cin, inter if bottleneck else growth, 1 if bottleneck else 3,
                               padding=0 if bottleneck else 1, bias=False)
        self.bn2 = nn.BatchNorm2d(inter) if bottleneck else None
        self.conv2 = nn.Conv2d(inter, growth, 3, padding=1, bias=False) if bottleneck else None

    def forward(self, x):
        y = self.conv1(bn_act(self.bn1, x))
        if self.conv2 is not None:
            y = self.conv2(bn_act(self.bn2, y))
        return torch.cat((x, y), 1)


class Transition(nn.Module):
    def __init__(self, cin, cout):
        super().__init__()
        self.bn1 = nn.BatchNorm2d(cin)
        self.conv1 = nn.Conv2d(cin, cout, 1, bias=False)

    def forward(self, x):
        return F.avg_pool2d(self.conv1(bn_act(self.bn1, x)), 2)


class DenseNet(nn.Module):
    def __init__(self, growth_rate=12, depth=100, reduction=0.5, num_classes=10, bottleneck=True):
        super().__init__()
        self.name = 'densenet%d_%d' % (depth, growth_rate)
        n = (depth - 4) // 3
        if bottleneck:
            n //= 2
        c = 2 * growth_rate
        self.conv1 = nn.Conv2d(3, c, 3, padding=1, bias=False)
        stages = []
        for si in range(3):
            for _ in range(n):
                stages.append(DenseLayer(c, growth_rate, bottleneck))
                c += growth_rate
            if si < 2:
                cout = int(math.floor(c * reduction))
                stages.append(Transition(c, cout))
                c = cout
        self.dense = nn.Sequential(*stages)
        self.bn1 = nn.BatchNorm2d(c)
        self.fc = nn.Linear(c, num_classes)
        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                m.weight.data.normal_(0, math.sqrt(2.0 / (m.kernel_size[0] * m.kernel_size[1] * m.out_channels)))
            elif isinstance(m, nn.Linear):
                m.bias.data.zero_()

    def forward(self, x):
        x = bn_act(self.bn1, self.dense(self.conv1(x)))
        return self.fc(F.adaptive_avg_pool2d(x, 1).flatten(1))


def densenet100_12(num_classes=10):
    return DenseNet(12, 100, 0.5, num_classes, True)
