"""The small MNIST network of the Horovod trainer (``horvod_pytorch.py:43-59``).

conv 1->10 k5, maxpool+ReLU, conv 10->20 k5, Dropout2d, maxpool+ReLU, fc 320->50, ReLU, dropout,
fc 50->10, log-softmax output (the trainer pairs it with ``nll_loss``).
"""
import torch.nn as nn
import torch.nn.functional as F

from ..ops import layers as L


class MnistNet(nn.Module):
    def __init__(self, num_classes: int = 10):
        super().__init__()
        self.conv1 = L.Conv2d(1, 10, kernel_size=5)
        self.conv2 = L.Conv2d(10, 20, kernel_size=5)
        self.conv2_drop = nn.Dropout2d()
        self.fc1 = L.Linear(320, 50)
        self.fc2 = L.Linear(50, num_classes)
        self.pool = L.MaxPool2d(2, 2)
        self.log_softmax_output = True

    def forward(self, x):
        x = L.conv_bn_act(x, self.conv1, None, relu=True, pool=True)
        x = F.relu(F.max_pool2d(self.conv2_drop(self.conv2(x)), 2))
        x = x.reshape(x.shape[0], -1)
        x = F.relu(self.fc1(x))
        x = F.dropout(x, training=self.training)
        x = self.fc2(x)
        return F.log_softmax(x.float(), dim=1)
