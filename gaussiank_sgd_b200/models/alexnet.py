"""AlexNet with local response normalisation (the Krizhevsky 2012 layout the reference keeps in
models/alexnet.py:9-87; the CLI itself trains torchvision's LRN-free variant, dl_trainer.py:110)."""
import torch.nn as nn


class AlexNet(nn.Module):
    def __init__(self, num_classes=1000):
        super().__init__()
        self.name = 'alexnet'
        def lrn():
            return nn.LocalResponseNorm(size=5, alpha=1e-4, beta=0.75, k=2.0)
        self.features = nn.Sequential(
            nn.Conv2d(3, 96, 11, stride=4), nn.ReLU(inplace=True), lrn(), nn.MaxPool2d(3, 2),
            nn.Conv2d(96, 256, 5, padding=2, groups=2), nn.ReLU(inplace=True), lrn(), nn.MaxPool2d(3, 2),
            nn.Conv2d(256, 384, 3, padding=1), nn.ReLU(inplace=True),
            nn.Conv2d(384, 384, 3, padding=1, groups=2), nn.ReLU(inplace=True),
            nn.Conv2d(384, 256, 3, padding=1, groups=2), nn.ReLU(inplace=True), nn.MaxPool2d(3, 2))
        self.classifier = nn.Sequential(
            nn.Linear(256 * 6 * 6, 4096), nn.ReLU(inplace=True), nn.Dropout(),
            nn.Linear(4096, 4096), nn.ReLU(inplace=True), nn.Dropout(),
            nn.Linear(4096, num_classes))

    def forward(self, x):
        x = nn.functional.adaptive_avg_pool2d(self.features(x), 6)
        return self.classifier(x.flatten(1))
