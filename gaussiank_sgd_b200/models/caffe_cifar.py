"""The Caffe "CIFAR-10 quick" network (models/caffe_cifar.py in the reference): three 5x5 convolutions with
pooling and two fully connected layers."""
import torch.nn as nn
import torch.nn.functional as F


class CifarCaffeNet(nn.Module):
    def __init__(self, num_classes=10):
        super().__init__()
        self.name = 'caffe_cifar'
        def block(cin, cout):
            return nn.Sequential(nn.Conv2d(cin, cout, 3, padding=1), nn.ReLU(inplace=True), nn.BatchNorm2d(cout))
        self.block_1 = nn.Sequential(block(3, 32), nn.MaxPool2d(3, 2))
        self.block_2 = nn.Sequential(block(32, 64), block(64, 64), nn.AvgPool2d(3, 2))
        self.block_3 = nn.Sequential(block(64, 128), block(128, 128), nn.AvgPool2d(3, 2))
        self.classifier = nn.Linear(128 * 9, num_classes)
        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                nn.init.kaiming_normal_(m.weight, mode='fan_out')
            elif isinstance(m, nn.Linear):
                nn.init.kaiming_normal_(m.weight)
                m.bias.data.zero_()

    def forward(self, x):
        x = self.block_3(self.block_2(self.block_1(x)))
        return self.classifier(x.flatten(1))


def caffe_cifar(num_classes=10):
    return CifarCaffeNet(num_classes)
