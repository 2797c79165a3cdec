"""Inception-v4 (Szegedy et al. 2016) — the architecture the reference takes from the Cadene port
(models/inceptionv4.py:264-337): stem, 4x Inception-A, Reduction-A, 7x Inception-B, Reduction-B,
3x Inception-C, 1536 -> classes.  42 679 816 parameters in 449 tensors.  Works at 299x299 and, through the
adaptive final pooling, at 224x224 (:300-306).

Every block is described as a list of branches; a branch is a list of layer specs
``(cout, kernel, stride, padding)`` or the strings 'max' / 'avg'."""
import torch
import torch.nn as nn
import torch.nn.functional as F
from ..ops.bn import bn_act


class ConvBN(nn.Module):
    def __init__(self, cin, cout, kernel, stride=1, padding=0):
        super().__init__()
        self.conv = nn.Conv2d(cin, cout, kernel, stride, padding, bias=False)
        self.bn = nn.BatchNorm2d(cout, eps=0.001, momentum=0.1, affine=True)

    def forward(self, x):
        return bn_act(self.bn, self.conv(x))


def _chain(cin, specs):
    layers = []
    for spec in specs:
        if spec == 'max':
            layers.append(nn.MaxPool2d(3, stride=2))
        elif spec == 'avg':
            layers.append(nn.AvgPool2d(3, stride=1, padding=1, count_include_pad=False))
        else:
            cout, k, s, p = spec
            layers.append(ConvBN(cin, cout, k, s, p))
            cin = cout
    return nn.Sequential(*layers), cin


class Branches(nn.Module):
    """Parallel branches concatenated along channels."""

    def __init__(self, cin, branches):
        super().__init__()
        mods, self.cout = [], 0
        for specs in branches:
            m, c = _chain(cin, specs)
            mods.append(m)
            self.cout += c
        self.branches = nn.ModuleList(mods)

    def forward(self, x):
        return torch.cat([b(x) for b in self.branches], 1)


class InceptionC(nn.Module):
    """The only block whose branches fork again (1x3 / 3x1 pairs)."""

    def __init__(self, cin=1536):
        super().__init__()
        self.branch0 = ConvBN(cin, 256, 1)
        self.branch1_0 = ConvBN(cin, 384, 1)
        self.branch1_1a = ConvBN(384, 256, (1, 3), 1, (0, 1))
        self.branch1_1b = ConvBN(384, 256, (3, 1), 1, (1, 0))
        self.branch2_0 = ConvBN(cin, 384, 1)
        self.branch2_1 = ConvBN(384, 448, (3, 1), 1, (1, 0))
        self.branch2_2 = ConvBN(448, 512, (1, 3), 1, (0, 1))
        self.branch2_3a = ConvBN(512, 256, (1, 3), 1, (0, 1))
        self.branch2_3b = ConvBN(512, 256, (3, 1), 1, (1, 0))
        self.branch3 = nn.Sequential(nn.AvgPool2d(3, stride=1, padding=1, count_include_pad=False), ConvBN(cin, 256, 1))

    def forward(self, x):
        b1 = self.branch1_0(x)
        b2 = self.branch2_2(self.branch2_1(self.branch2_0(x)))
        return torch.cat([self.branch0(x), self.branch1_1a(b1), self.branch1_1b(b1),
                          self.branch2_3a(b2), self.branch2_3b(b2), self.branch3(x)], 1)


A = [[(96, 1, 1, 0)], [(64, 1, 1, 0), (96, 3, 1, 1)], [(64, 1, 1, 0), (96, 3, 1, 1), (96, 3, 1, 1)],
     ['avg', (96, 1, 1, 0)]]
RED_A = [[(384, 3, 2, 0)], [(192, 1, 1, 0), (224, 3, 1, 1), (256, 3, 2, 0)], ['max']]
B = [[(384, 1, 1, 0)],
     [(192, 1, 1, 0), (224, (1, 7), 1, (0, 3)), (256, (7, 1), 1, (3, 0))],
     [(192, 1, 1, 0), (192, (7, 1), 1, (3, 0)), (224, (1, 7), 1, (0, 3)), (224, (7, 1), 1, (3, 0)),
      (256, (1, 7), 1, (0, 3))],
     ['avg', (128, 1, 1, 0)]]
RED_B = [[(192, 1, 1, 0), (192, 3, 2, 0)],
         [(256, 1, 1, 0), (256, (1, 7), 1, (0, 3)), (320, (7, 1), 1, (3, 0)), (320, 3, 2, 0)], ['max']]


class InceptionV4(nn.Module):
    def __init__(self, num_classes=1001):
        super().__init__()
        self.name = 'inceptionv4'
        self.input_size = (3, 299, 299)
        stem = [ConvBN(3, 32, 3, 2), ConvBN(32, 32, 3, 1), ConvBN(32, 64, 3, 1, 1),
                Branches(64, [['max'], [(96, 3, 2, 0)]]),                                       # Mixed_3a -> 160
                Branches(160, [[(64, 1, 1, 0), (96, 3, 1, 0)],
                               [(64, 1, 1, 0), (64, (1, 7), 1, (0, 3)), (64, (7, 1), 1, (3, 0)), (96, 3, 1, 0)]]),  # Mixed_4a -> 192
                Branches(192, [[(192, 3, 2, 0)], ['max']])]                                     # Mixed_5a -> 384
        body = [Branches(384, A) for _ in range(4)] + [Branches(384, RED_A)]
        body += [Branches(1024, B) for _ in range(7)] + [Branches(1024, RED_B)]
        body += [InceptionC(1536) for _ in range(3)]
        self.features = nn.Sequential(*stem, *body)
        self.last_linear = nn.Linear(1536, num_classes)

    def logits(self, features):
        x = F.adaptive_avg_pool2d(features, 1).flatten(1)
        return self.last_linear(x)

    def forward(self, x):
        return self.logits(self.features(x))


def inceptionv4(num_classes=1000, pretrained=None):
    if pretrained:
        raise RuntimeError('no network access: pretrained weights cannot be downloaded')
    return InceptionV4(num_classes=num_classes)
