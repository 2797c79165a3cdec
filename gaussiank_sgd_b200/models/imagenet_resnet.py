"""ImageNet ResNet-18/34/50/101/152 (He et al. 2015) — the local copy the reference keeps next to the
torchvision model it actually trains (models/imagenet_resnet.py:82-192).  Plan-driven: a depth is a block
type plus the number of blocks per stage."""
import torch.nn as nn
import torch.nn.functional as F
from ..ops.bn import bn_act

PLANS = {18: ('basic', (2, 2, 2, 2)), 34: ('basic', (3, 4, 6, 3)), 50: ('bottleneck', (3, 4, 6, 3)),
         101: ('bottleneck', (3, 4, 23, 3)), 152: ('bottleneck', (3, 8, 36, 3))}


class Block(nn.Module):
    def __init__(self, kind, cin, width, stride):
        super().__init__()
        self.expansion = 1 if kind == 'basic' else 4
        cout = width * self.expansion
        if kind == 'basic':
            specs = [(cin, width, 3, stride, 1), (width, width, 3, 1, 1)]
        else:
            specs = [(cin, width, 1, 1, 0), (width, width, 3, stride, 1), (width, cout, 1, 1, 0)]
        self.convs = nn.ModuleList(nn.Conv2d(a, b, k, s, p, bias=False) for a, b, k, s, p in specs)
        self.bns = nn.ModuleList(nn.BatchNorm2d(b) for _, b, *_ in specs)
        self.downsample = None
        if stride != 1 or cin != cout:
            self.downsample = nn.Sequential(nn.Conv2d(cin, cout, 1, stride, bias=False), nn.BatchNorm2d(cout))

    def forward(self, x):
        s = x if self.downsample is None else bn_act(self.downsample[1], self.downsample[0](x), relu=False)
        y, last = x, len(self.convs) - 1
        for i, (conv, bn) in enumerate(zip(self.convs, self.bns)):
            y = bn_act(bn, conv(y), residual=s if i == last else None)
        return y


class ResNet(nn.Module):
    def __init__(self, depth=50, num_classes=1000):
        super().__init__()
        kind, counts = PLANS[depth]
        self.name = 'resnet%d' % depth
        self.conv1 = nn.Conv2d(3, 64, 7, 2, 3, bias=False)
        self.bn1 = nn.BatchNorm2d(64)
        layers, cin = [], 64
        for si, (width, n) in enumerate(zip((64, 128, 256, 512), counts)):
            for bi in range(n):
                blk = Block(kind, cin, width, 2 if (bi == 0 and si > 0) else 1)
                layers.append(blk)
                cin = width * blk.expansion
        self.layers = nn.Sequential(*layers)
        self.fc = nn.Linear(cin, num_classes)
        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                nn.init.kaiming_normal_(m.weight, mode='fan_out', nonlinearity='relu')

    def forward(self, x):
        x = F.max_pool2d(bn_act(self.bn1, self.conv1(x)), 3, 2, 1)
        x = F.adaptive_avg_pool2d(self.layers(x), 1).flatten(1)
        return self.fc(x)


def _mk(depth):
    def build(pretrained=False, num_classes=1000, **kw):
        if pretrained:
            raise RuntimeError('no network access: pretrained weights cannot be downloaded')
        return ResNet(depth, num_classes)
    return build


resnet18, resnet34, resnet50, resnet101, resnet152 = (_mk(d) for d in (18, 34, 50, 101, 152))
