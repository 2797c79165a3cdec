"""Model zoo (the reference's models/ package, models/__init__.py:16-30, plus MnistNet from
dl_trainer.py:61-78).  Builders that share their name with their module are imported after the module so
that the package attribute is the builder, like in the reference."""
from . import lstm  # noqa: F401  (used as a module: lstm.lstm, lstm.repackage_hidden)
from .alexnet import AlexNet  # noqa: F401
from .caffe_cifar import caffe_cifar  # noqa: F401
from .cifar_resnet import (CifarResNet, preresnet20, preresnet32, preresnet44, preresnet56, preresnet110,  # noqa: F401
                           resnet20, resnet32, resnet44, resnet56, resnet110, resnet_mod20, resnet_mod32,
                           resnet_mod44, resnet_mod56, resnet_mod110)
from .deepspeech import DeepSpeech  # noqa: F401
from .densenet import densenet100_12  # noqa: F401
from .googlenet import GoogLeNet, googlenet  # noqa: F401
from .imagenet_resnet import resnet18, resnet34, resnet50, resnet101, resnet152  # noqa: F401
from .inceptionv4 import InceptionV4, inceptionv4  # noqa: F401
from .lstman4 import create_net as LSTMAN4  # noqa: F401
from .resnext import resnext29_8_64, resnext29_16_64  # noqa: F401
from .small import FCN5Net, LeNet, LinearRegression, MnistNet  # noqa: F401
from .vgg import VGG  # noqa: F401
