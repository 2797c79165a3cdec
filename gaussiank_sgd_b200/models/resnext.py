"""ResNeXt-29 for CIFAR (Xie et al. 2016; models/resnext.py in the reference): 3 stages x 3 bottlenecks with
grouped 3x3 convolutions, cardinality 8 or 16, base width 64."""
import torch.nn as nn
import torch.nn.functional as F
from ..ops.bn import bn_act


class ResNeXtBottleneck(nn.Module):
    expansion = 4

    def __init__(self, cin, planes, cardinality, base_width, stride=1):
        super().__init__()
        d = int(planes * (base_width / 64.0)) * cardinality
        cout = planes * self.expansion
        self.conv_reduce = nn.Conv2d(cin, d, 1, bias=False)
        self.bn_reduce = nn.BatchNorm2d(d)
        self.conv_conv = nn.Conv2d(d, d, 3, stride, 1, groups=cardinality, bias=False)
        self.bn = nn.BatchNorm2d(d)
        self.conv_expand = nn.Conv2d(d, cout, 1, bias=False)
        self.bn_expand = nn.BatchNorm2d(cout)
        self.downsample = None
        if stride != 1 or cin != cout:
            self.downsample = nn.Sequential(nn.Conv2d(cin, cout, 1, stride, bias=False), nn.BatchNorm2d(cout))

    def forward(self, x):
        y = bn_act(self.bn_reduce, self.conv_reduce(x))
        y = bn_act(self.bn, self.conv_conv(y))
        s = x if self.downsample is None else self.downsample(x)
        return bn_act(self.bn_expand, self.conv_expand(y), residual=s)


class CifarResNeXt(nn.Module):
    def __init__(self, cardinality, depth, num_classes, base_width=64):
        super().__init__()
        assert (depth - 2) % 9 == 0
        n = (depth - 2) // 9
        self.name = 'resnext%d_%d_%d' % (depth, cardinality, base_width)
        self.conv_1_3x3 = nn.Conv2d(3, 64, 3, 1, 1, bias=False)
        self.bn_1 = nn.BatchNorm2d(64)
        blocks, cin = [], 64
        for si, planes in enumerate((64, 128, 256)):
            for bi in range(n):
                blocks.append(ResNeXtBottleneck(cin, planes, cardinality, base_width, 2 if (si > 0 and bi == 0) else 1))
                cin = planes * ResNeXtBottleneck.expansion
        self.stages = nn.Sequential(*blocks)
        self.classifier = nn.Linear(cin, num_classes)
        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                nn.init.kaiming_normal_(m.weight, mode='fan_out')
            elif isinstance(m, nn.Linear):
                nn.init.kaiming_normal_(m.weight)
                m.bias.data.zero_()

    def forward(self, x):
        x = bn_act(self.bn_1, self.conv_1_3x3(x))
        return self.classifier(F.adaptive_avg_pool2d(self.stages(x), 1).flatten(1))


def resnext29_8_64(num_classes=10):
    return CifarResNeXt(8, 29, num_classes, 64)


def resnext29_16_64(num_classes=10):
    return CifarResNeXt(16, 29, num_classes, 64)
