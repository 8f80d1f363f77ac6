"""LeNet for 28x28 single-channel input.

Same architecture, parameter names and shapes as ``src/model_ops/lenet.py:15-36`` (8 tensors,
431 080 parameters): conv 1->20 k5, maxpool 2, ReLU, conv 20->50 k5, maxpool 2, ReLU,
fc 800->500, fc 500->10 with no activation between the two fully connected layers.  The
``ceriation`` loss module is kept only so ``state_dict()._metadata`` has the same entries as the
reference checkpoint ``src/model_step_.zip``.  ``LeNetSplit`` (``lenet.py:38-259``, dead mpi4py
code) has no module here; its capability -- communication overlapped with a layer-wise backward
-- is provided for every model by :class:`ewdml.parallel.overlap.BucketScheduler`.
"""
import torch.nn as nn

from ..ops import layers as L


class LeNet(nn.Module):
    def __init__(self, num_classes: int = 10):
        super().__init__()
        self.conv1 = L.Conv2d(1, 20, 5, 1)
        self.conv2 = L.Conv2d(20, 50, 5, 1)
        self.fc1 = L.Linear(4 * 4 * 50, 500)
        self.fc2 = L.Linear(500, num_classes)
        self.ceriation = nn.CrossEntropyLoss()
        self.pool = L.MaxPool2d(2, 2)

    def forward(self, x):
        # relu(maxpool(conv(x))) as in lenet.py:24-29 (== maxpool(relu(.)): relu is monotonic)
        x = L.conv_bn_act(x, self.conv1, None, relu=True, pool=True)
        x = L.conv_bn_act(x, self.conv2, None, relu=True, pool=True)
        x = x.reshape(x.shape[0], -1)              # logical (C,H,W) order, as x.view(-1, 800)
        x = self.fc1(x)
        return self.fc2(x)

    def name(self):
        return "lenet"
