"""MNIST-sized networks: MnistNet (dl_trainer.py:61-78), FCN5Net / LinearRegression
(models/fcn.py:9-35), LeNet (models/lenet.py:5-24)."""
import torch.nn as nn
import torch.nn.functional as F


class MnistNet(nn.Module):
    def __init__(self):
        super().__init__()
        self.conv1 = nn.Conv2d(1, 10, 5)
        self.conv2 = nn.Conv2d(10, 20, 5)
        self.conv2_drop = nn.Dropout2d()
        self.fc1 = nn.Linear(320, 50)
        self.fc2 = nn.Linear(50, 10)
        self.name = 'mnistnet'

    def forward(self, x):
        x = F.relu(F.max_pool2d(self.conv1(x), 2))
        x = F.relu(F.max_pool2d(self.conv2_drop(self.conv2(x)), 2))
        x = F.relu(self.fc1(x.flatten(1)))
        x = F.dropout(x, training=self.training)
        return self.fc2(x)


class FCN5Net(nn.Module):
    def __init__(self, hidden: int = 200):
        super().__init__()
        self.fc1 = nn.Linear(784, hidden)
        self.fc2 = nn.Linear(hidden, hidden)
        self.fc3 = nn.Linear(hidden, 10)
        self.name = 'fcn5net'

    def forward(self, x):
        x = F.relu(self.fc1(x.reshape(-1, 784)))
        return self.fc3(F.relu(self.fc2(x)))


class LinearRegression(nn.Module):
    def __init__(self):
        super().__init__()
        self.fc1 = nn.Linear(784, 10)
        self.name = 'lr'

    def forward(self, x):
        return self.fc1(x.reshape(-1, 784))


class LeNet(nn.Module):
    def __init__(self):
        super().__init__()
        self.conv1 = nn.Conv2d(1, 6, 5)
        self.conv2 = nn.Conv2d(6, 16, 5)
        self.fc1 = nn.Linear(400, 120)
        self.fc2 = nn.Linear(120, 84)
        self.fc3 = nn.Linear(84, 10)
        self.name = 'lenet'

    def forward(self, x):
        x = F.max_pool2d(F.relu(self.conv1(x)), 2)
        x = F.max_pool2d(F.relu(self.conv2(x)), 2)
        x = F.relu(self.fc2(F.relu(self.fc1(x.flatten(1)))))
        return self.fc3(x)
