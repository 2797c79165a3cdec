"""Model zoo (the reference's models/ package, models/__init__.py:16-30, plus MnistNet from
dl_trainer.py:61-78).  Heavy or rarely used families are imported on first use."""
from .cifar_resnet import (CifarResNet, preresnet20, preresnet32, preresnet44, preresnet56, preresnet110,  # noqa: F401
                           resnet20, resnet32, resnet44, resnet56, resnet110, resnet_mod20, resnet_mod32,
                           resnet_mod44, resnet_mod56, resnet_mod110)
from .small import FCN5Net, LeNet, LinearRegression, MnistNet  # noqa: F401
from .vgg import VGG  # noqa: F401

_LAZY = {
    'inceptionv4': ('inceptionv4', 'inceptionv4'),
    'googlenet': ('googlenet', 'googlenet'),
    'resnext29_8_64': ('resnext', 'resnext29_8_64'), 'resnext29_16_64': ('resnext', 'resnext29_16_64'),
    'caffe_cifar': ('caffe_cifar', 'caffe_cifar'),
    'densenet100_12': ('densenet', 'densenet100_12'),
    'resnet18': ('imagenet_resnet', 'resnet18'), 'resnet34': ('imagenet_resnet', 'resnet34'),
    'resnet50': ('imagenet_resnet', 'resnet50'), 'resnet101': ('imagenet_resnet', 'resnet101'),
    'resnet152': ('imagenet_resnet', 'resnet152'),
    'AlexNet': ('alexnet', 'AlexNet'),
    'LSTMAN4': ('lstman4', 'create_net'),
    'DeepSpeech': ('deepspeech', 'DeepSpeech'),
}


def __getattr__(name):
    if name in _LAZY:
        import importlib
        mod, attr = _LAZY[name]
        return getattr(importlib.import_module('.' + mod, __name__), attr)
    raise AttributeError(name)
