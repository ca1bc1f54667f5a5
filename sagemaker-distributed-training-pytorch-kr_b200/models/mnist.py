"""MNIST CNN (reference: 1_training_mnist_ddp/model_def.py:22-45; 1,199,882 parameters).

conv(1->32,3) - ReLU - conv(32->64,3) - ReLU - maxpool2 - dropout(.25, channel-wise) - flatten -
fc(9216->128) - ReLU - dropout(.5) - fc(128->10) - log_softmax.  The reference applies Dropout2d to the
2-D fc activations (model_def.py:42), which torch treats as plain element dropout with a warning; that
behaviour is written out explicitly here.
"""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F


class Net(nn.Module):
    """`native_ops=True` (entry points: --native-ops): on a CUDA device the two conv + bias + ReLU stages, the max-pool and
    the log-softmax run the fp32 kernels of csrc/conv.cu instead of cuDNN / ATen (same parameters, same state_dict)."""

    def __init__(self, native_ops: bool = False):
        super().__init__()
        self.native_ops = native_ops
        self.conv1 = nn.Conv2d(1, 32, 3, 1)
        self.conv2 = nn.Conv2d(32, 64, 3, 1)
        self.dropout1 = nn.Dropout2d(0.25)
        self.dropout2 = nn.Dropout(0.5)
        self.fc1 = nn.Linear(9216, 128)
        self.fc2 = nn.Linear(128, 10)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if self.native_ops and x.dtype == torch.float32:
            from ..ops import functional as Fn
            x = Fn.conv2d_bias_relu(x, self.conv1.weight, self.conv1.bias)
            x = Fn.conv2d_bias_relu(x, self.conv2.weight, self.conv2.bias)
            x = Fn.max_pool2x2(x)
            x = self.dropout1(x)
            x = torch.flatten(x, 1)
            x = F.relu(self.fc1(x))
            x = self.dropout2(x)
            return Fn.log_softmax_small(self.fc2(x))
        x = F.relu(self.conv1(x))
        x = F.relu(self.conv2(x))
        x = F.max_pool2d(x, 2)
        x = self.dropout1(x)
        x = torch.flatten(x, 1)
        x = F.relu(self.fc1(x))
        x = self.dropout2(x)
        x = self.fc2(x)
        return F.log_softmax(x, dim=1)


def count_parameters(m: nn.Module) -> int:
    return sum(p.numel() for p in m.parameters())
