"""GoogLeNet / Inception v1 as the reference trains it (models/googlenet.py:18-233): the torchvision
layout, but convolutions carry no batch norm (commented out at :223-233, so a block is conv(bias=False) +
ReLU) and the "5x5" branch is a 3x3 (:176-178).  6 610 344 parameters in 59 tensors without aux heads.
Table-driven: one row per inception block."""
import torch
import torch.nn as nn
import torch.nn.functional as F

# (in, 1x1, 3x3 reduce, 3x3, "5x5" reduce, "5x5", pool proj)
INCEPTION = {
    '3a': (192, 64, 96, 128, 16, 32, 32), '3b': (256, 128, 128, 192, 32, 96, 64),
    '4a': (480, 192, 96, 208, 16, 48, 64), '4b': (512, 160, 112, 224, 24, 64, 64),
    '4c': (512, 128, 128, 256, 24, 64, 64), '4d': (512, 112, 144, 288, 32, 64, 64),
    '4e': (528, 256, 160, 320, 32, 128, 128),
    '5a': (832, 256, 160, 320, 32, 128, 128), '5b': (832, 384, 192, 384, 48, 128, 128),
}


class ConvReLU(nn.Module):
    def __init__(self, cin, cout, **kw):
        super().__init__()
        self.conv = nn.Conv2d(cin, cout, bias=False, **kw)

    def forward(self, x):
        return F.relu(self.conv(x), inplace=True)


class Inception(nn.Module):
    def __init__(self, cin, c1, c3r, c3, c5r, c5, cp):
        super().__init__()
        self.branch1 = ConvReLU(cin, c1, kernel_size=1)
        self.branch2 = nn.Sequential(ConvReLU(cin, c3r, kernel_size=1), ConvReLU(c3r, c3, kernel_size=3, padding=1))
        self.branch3 = nn.Sequential(ConvReLU(cin, c5r, kernel_size=1), ConvReLU(c5r, c5, kernel_size=3, padding=1))
        self.branch4 = nn.Sequential(nn.MaxPool2d(3, stride=1, padding=1, ceil_mode=True),
                                     ConvReLU(cin, cp, kernel_size=1))

    def forward(self, x):
        return torch.cat([self.branch1(x), self.branch2(x), self.branch3(x), self.branch4(x)], 1)


class InceptionAux(nn.Module):
    def __init__(self, cin, num_classes):
        super().__init__()
        self.conv = ConvReLU(cin, 128, kernel_size=1)
        self.fc1 = nn.Linear(2048, 1024)
        self.fc2 = nn.Linear(1024, num_classes)

    def forward(self, x):
        x = self.conv(F.adaptive_avg_pool2d(x, (4, 4))).flatten(1)
        x = F.dropout(F.relu(self.fc1(x), inplace=True), 0.7, training=self.training)
        return self.fc2(x)


class GoogLeNet(nn.Module):
    def __init__(self, num_classes=1000, aux_logits=False, transform_input=False, init_weights=True):
        super().__init__()
        self.name = 'googlenet'
        self.aux_logits, self.transform_input = aux_logits, transform_input
        self.conv1 = ConvReLU(3, 64, kernel_size=7, stride=2, padding=3)
        self.conv2 = ConvReLU(64, 64, kernel_size=1)
        self.conv3 = ConvReLU(64, 192, kernel_size=3, padding=1)
        for key, cfg in INCEPTION.items():
            setattr(self, 'inception' + key, Inception(*cfg))
        if aux_logits:
            self.aux1, self.aux2 = InceptionAux(512, num_classes), InceptionAux(528, num_classes)
        self.dropout = nn.Dropout(0.2)
        self.fc = nn.Linear(1024, num_classes)
        if init_weights:
            for m in self.modules():
                if isinstance(m, (nn.Conv2d, nn.Linear)):
                    nn.init.trunc_normal_(m.weight, std=0.01, a=-0.02, b=0.02)   # scipy truncnorm in the reference (:91-99)

    def forward(self, x):
        if self.transform_input:
            scale = x.new_tensor([0.229, 0.224, 0.225]).view(1, 3, 1, 1) / 0.5
            shift = (x.new_tensor([0.485, 0.456, 0.406]).view(1, 3, 1, 1) - 0.5) / 0.5
            x = x * scale + shift
        pool = lambda t, k=3: F.max_pool2d(t, k, stride=2, ceil_mode=True)  # noqa: E731
        x = pool(self.conv1(x))
        x = pool(self.conv3(self.conv2(x)))
        x = pool(self.inception3b(self.inception3a(x)))
        x = self.inception4a(x)
        aux1 = self.aux1(x) if (self.aux_logits and self.training) else None
        x = self.inception4d(self.inception4c(self.inception4b(x)))
        aux2 = self.aux2(x) if (self.aux_logits and self.training) else None
        x = pool(self.inception4e(x), 2)
        x = self.inception5b(self.inception5a(x))
        x = self.fc(self.dropout(F.adaptive_avg_pool2d(x, 1).flatten(1)))
        if self.aux_logits and self.training:
            return x, aux2, aux1
        return x


def googlenet(pretrained=False, progress=True, **kwargs):
    if pretrained:
        raise RuntimeError('no network access: pretrained weights cannot be downloaded')
    return GoogLeNet(**kwargs)
