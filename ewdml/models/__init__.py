"""Model zoo + factory (the ``build_model`` of ``src/util.py:7-18``)."""
from .lenet import LeNet
from .mnist_net import MnistNet
from .resnet import ResNet, BasicBlock, Bottleneck, ResNet18, ResNet34, ResNet50, ResNet101, ResNet152
from .vgg import (VGG, vgg11, vgg11_bn, vgg13, vgg13_bn, vgg16, vgg16_bn, vgg19, vgg19_bn)

_FACTORY = {
    "LeNet": LeNet,
    "MnistNet": MnistNet,
    "ResNet18": ResNet18, "ResNet34": ResNet34, "ResNet50": ResNet50,
    "ResNet101": ResNet101, "ResNet152": ResNet152,
    # the reference maps "VGG11" to the BatchNorm variant (util.py:17-18)
    "VGG11": vgg11_bn, "VGG13": vgg13_bn, "VGG16": vgg16_bn, "VGG19": vgg19_bn,
}


def build_model(model_name: str, num_classes: int = 10):
    """Build a network by its ``--network`` name.  Accepts the README spellings too
    (``Resnet50``, ``Lenet``; reference README.md:9)."""
    key = {k.lower(): k for k in _FACTORY}.get(model_name.lower())
    if key is None:
        raise ValueError(f"unknown network {model_name!r}; choose from {sorted(_FACTORY)}")
    return _FACTORY[key](num_classes)


def input_shape(model_name: str):
    return (1, 28, 28) if model_name.lower() in ("lenet", "mnistnet") else (3, 32, 32)
