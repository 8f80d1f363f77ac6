"""CIFAR-sized VGG family (configurations A/B/D/E, with or without BatchNorm).

Parameter names/shapes follow ``src/model_ops/vgg.py:15-108``: ``features`` is a flat Sequential
of conv3x3(pad 1) [+BatchNorm] + ReLU groups and 2x2 max-pools, ``classifier`` is
Dropout-Linear(512,512)-ReLU-Dropout-Linear(512,512)-ReLU-Linear(512,C); conv weights are
He-normal with fan-out ``k*k*C_out`` and conv biases zero (``vgg.py:32-36``).  ``vgg11_bn`` has 38
parameter tensors / 9 756 426 parameters.

Execution differs from a plain Sequential: :class:`FusedFeatures` walks the same module list but
hands each (conv, bn, relu) group to the BatchNorm module's fused ``relu=True`` path, which on the
native backend is a single normalise+activate kernel after the conv's statistics epilogue.
"""
import math

import torch.nn as nn

from ..ops import layers as L

CFG = {
    "A": [64, "M", 128, "M", 256, 256, "M", 512, 512, "M", 512, 512, "M"],
    "B": [64, 64, "M", 128, 128, "M", 256, 256, "M", 512, 512, "M", 512, 512, "M"],
    "D": [64, 64, "M", 128, 128, "M", 256, 256, 256, "M", 512, 512, 512, "M", 512, 512, 512, "M"],
    "E": [64, 64, "M", 128, 128, "M", 256, 256, 256, 256, "M", 512, 512, 512, 512, "M",
          512, 512, 512, 512, "M"],
}


class FusedFeatures(nn.Sequential):
    """Sequential whose forward fuses conv -> [bn] -> relu -> [maxpool] groups (same children, same keys)."""

    def forward(self, x):
        mods = list(self)
        i = 0
        while i < len(mods):
            m = mods[i]
            if isinstance(m, nn.Conv2d):
                j = i + 1
                bn = mods[j] if j < len(mods) and isinstance(mods[j], nn.BatchNorm2d) else None
                j += bn is not None
                relu = j < len(mods) and isinstance(mods[j], nn.ReLU)
                j += relu
                pool = j < len(mods) and isinstance(mods[j], nn.MaxPool2d)
                j += pool
                x = L.conv_bn_act(x, m, bn, relu=relu, pool=pool)
                i = j
            else:
                x = m(x)
                i += 1
        return x


def make_layers(cfg, batch_norm: bool = False) -> nn.Sequential:
    mods, c_in = [], 3
    for v in cfg:
        if v == "M":
            mods.append(L.MaxPool2d(kernel_size=2, stride=2))
            continue
        mods.append(L.Conv2d(c_in, v, kernel_size=3, padding=1))
        if batch_norm:
            mods.append(L.BatchNorm2d(v))
        mods.append(L.ReLU(inplace=True))
        c_in = v
    return FusedFeatures(*mods)


class VGG(nn.Module):
    def __init__(self, features: nn.Sequential, num_classes: int = 10):
        super().__init__()
        self.features = features
        self.classifier = nn.Sequential(
            L.Dropout(), L.Linear(512, 512), L.ReLU(True),
            L.Dropout(), L.Linear(512, 512), L.ReLU(True),
            L.Linear(512, num_classes))
        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                fan_out = m.kernel_size[0] * m.kernel_size[1] * m.out_channels
                nn.init.normal_(m.weight, 0.0, math.sqrt(2.0 / fan_out))
                nn.init.zeros_(m.bias)

    def forward(self, x):
        x = self.features(x)
        x = x.reshape(x.shape[0], -1)
        c = self.classifier
        x = L.linear_act(c[0](x), c[1], relu=True)          # Dropout - Linear - ReLU
        x = L.linear_act(c[3](x), c[4], relu=True)          # Dropout - Linear - ReLU
        return c[6](x)


def _vgg(key, bn, num_classes):
    return VGG(make_layers(CFG[key], batch_norm=bn), num_classes=num_classes)


def vgg11(num_classes=10): return _vgg("A", False, num_classes)
def vgg11_bn(num_classes=10): return _vgg("A", True, num_classes)
def vgg13(num_classes=10): return _vgg("B", False, num_classes)
def vgg13_bn(num_classes=10): return _vgg("B", True, num_classes)
def vgg16(num_classes=10): return _vgg("D", False, num_classes)
def vgg16_bn(num_classes=10): return _vgg("D", True, num_classes)
def vgg19(num_classes=10): return _vgg("E", False, num_classes)
def vgg19_bn(num_classes=10): return _vgg("E", True, num_classes)
