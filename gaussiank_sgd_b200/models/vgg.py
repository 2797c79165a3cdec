"""CIFAR VGG-11/13/16/19 with batch norm and a single 512->10 classifier
(models/vgg.py:14-38 in the reference; VGG-16 has 14 728 266 parameters in 54 tensors)."""
import torch.nn as nn

PLANS = {
    'VGG11': (1, 1, 2, 2, 2),
    'VGG13': (2, 2, 2, 2, 2),
    'VGG16': (2, 2, 3, 3, 3),
    'VGG19': (2, 2, 4, 4, 4),
}
STAGE_WIDTHS = (64, 128, 256, 512, 512)


class VGG(nn.Module):
    def __init__(self, vgg_name: str = 'VGG16', num_classes: int = 10):
        super().__init__()
        self.name = vgg_name.lower()
        layers, cin = [], 3
        for reps, width in zip(PLANS[vgg_name.upper()], STAGE_WIDTHS):
            for _ in range(reps):
                layers += [nn.Conv2d(cin, width, 3, padding=1), nn.BatchNorm2d(width), nn.ReLU(inplace=True)]
                cin = width
            layers.append(nn.MaxPool2d(2, 2))
        self.features = nn.Sequential(*layers)
        self.fc = nn.Linear(512, num_classes)

    def forward(self, x):
        return self.fc(self.features(x).flatten(1))
