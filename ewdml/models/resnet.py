"""CIFAR ResNets (3x3 stem, no stem max-pool, 4x4 average pool before the classifier).

Parameter names/shapes follow ``src/model_ops/resnet.py:14-112``: ``conv1/bn1``, ``layer1..4`` of
``BasicBlock`` (expansion 1) or ``Bottleneck`` (expansion 4) with a ``shortcut`` Sequential of
(1x1 conv, bn) when the shape changes, and ``linear``.  ResNet50 = 161 parameter tensors,
23 520 842 parameters, 320 state_dict keys.

Fixed relative to the reference: ``ResNet34`` takes ``num_classes`` (the
reference factory call is missing the argument, ``util.py:14``) and ``ResNet50/101/152`` honour
``num_classes`` instead of hard-coding 10 (``resnet.py:105-112``).

Every ``bn(conv(x))`` + ReLU (+ residual) is expressed through ``BatchNorm2d.forward(relu=...,
residual=...)`` so the native backend runs one fused normalise/add/activate kernel per block edge.
"""
import torch.nn as nn

from ..ops import layers as L


def _shortcut(in_planes, out_planes, stride):
    if stride != 1 or in_planes != out_planes:
        return nn.Sequential(L.Conv2d(in_planes, out_planes, kernel_size=1, stride=stride, bias=False),
                             L.BatchNorm2d(out_planes))
    return nn.Sequential()


def _run_shortcut(sc, x):
    if len(sc) == 0:
        return x
    return L.conv_bn_act(x, sc[0], sc[1])


class BasicBlock(nn.Module):
    expansion = 1

    def __init__(self, in_planes, planes, stride=1):
        super().__init__()
        self.conv1 = L.Conv2d(in_planes, planes, kernel_size=3, stride=stride, padding=1, bias=False)
        self.bn1 = L.BatchNorm2d(planes)
        self.conv2 = L.Conv2d(planes, planes, kernel_size=3, stride=1, padding=1, bias=False)
        self.bn2 = L.BatchNorm2d(planes)
        self.shortcut = _shortcut(in_planes, self.expansion * planes, stride)

    def forward(self, x):
        out = L.conv_bn_act(x, self.conv1, self.bn1, relu=True)
        return L.conv_bn_act(out, self.conv2, self.bn2, relu=True, residual=_run_shortcut(self.shortcut, x))


class Bottleneck(nn.Module):
    expansion = 4

    def __init__(self, in_planes, planes, stride=1):
        super().__init__()
        self.conv1 = L.Conv2d(in_planes, planes, kernel_size=1, bias=False)
        self.bn1 = L.BatchNorm2d(planes)
        self.conv2 = L.Conv2d(planes, planes, kernel_size=3, stride=stride, padding=1, bias=False)
        self.bn2 = L.BatchNorm2d(planes)
        self.conv3 = L.Conv2d(planes, self.expansion * planes, kernel_size=1, bias=False)
        self.bn3 = L.BatchNorm2d(self.expansion * planes)
        self.shortcut = _shortcut(in_planes, self.expansion * planes, stride)

    def forward(self, x):
        out = L.conv_bn_act(x, self.conv1, self.bn1, relu=True)
        out = L.conv_bn_act(out, self.conv2, self.bn2, relu=True)
        return L.conv_bn_act(out, self.conv3, self.bn3, relu=True, residual=_run_shortcut(self.shortcut, x))


class ResNet(nn.Module):
    def __init__(self, block, num_blocks, num_classes=10):
        super().__init__()
        self.in_planes = 64
        self.conv1 = L.Conv2d(3, 64, kernel_size=3, stride=1, padding=1, bias=False)
        self.bn1 = L.BatchNorm2d(64)
        self.layer1 = self._make_layer(block, 64, num_blocks[0], 1)
        self.layer2 = self._make_layer(block, 128, num_blocks[1], 2)
        self.layer3 = self._make_layer(block, 256, num_blocks[2], 2)
        self.layer4 = self._make_layer(block, 512, num_blocks[3], 2)
        self.linear = L.Linear(512 * block.expansion, num_classes)
        self.avgpool = L.AvgPool2d(4)

    def _make_layer(self, block, planes, n, stride):
        blocks = []
        for s in [stride] + [1] * (n - 1):
            blocks.append(block(self.in_planes, planes, s))
            self.in_planes = planes * block.expansion
        return nn.Sequential(*blocks)

    def forward(self, x):
        out = L.conv_bn_act(x, self.conv1, self.bn1, relu=True)
        out = self.layer4(self.layer3(self.layer2(self.layer1(out))))
        out = self.avgpool(out)
        return self.linear(out.reshape(out.shape[0], -1))


def ResNet18(num_classes=10): return ResNet(BasicBlock, [2, 2, 2, 2], num_classes)
def ResNet34(num_classes=10): return ResNet(BasicBlock, [3, 4, 6, 3], num_classes)
def ResNet50(num_classes=10): return ResNet(Bottleneck, [3, 4, 6, 3], num_classes)
def ResNet101(num_classes=10): return ResNet(Bottleneck, [3, 4, 23, 3], num_classes)
def ResNet152(num_classes=10): return ResNet(Bottleneck, [3, 8, 36, 3], num_classes)
