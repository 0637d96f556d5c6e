"""DenseNet 121/169/201/264/161 — module tree / state_dict keys of reference src/densenet.py:17-190:
pre-activation layers BN-ReLU-Dropout-conv1x1(4k) / BN-ReLU-Dropout-conv3x3(k) concatenated to
their input (:104-140), blocks (:143-168), transitions BN-ReLU-conv1x1(C/2)-AvgPool2/2 (:171-190),
final BN-ReLU-GAP (:80-82).  BatchNorm here is unconditional (not governed by setting.batch_norm)."""
from __future__ import annotations

import torch
import torch.nn as nn

from ._common import ZooModel

_CFG = {"121": (32, (6, 12, 24, 16), 64), "169": (32, (6, 12, 32, 32), 64),
        "201": (32, (6, 12, 48, 32), 64), "264": (32, (6, 12, 64, 48), 64),
        "161": (48, (6, 12, 36, 24), 96)}


def _preact(network, num_filters, **kwargs):
    return nn.Sequential(nn.BatchNorm2d(network.in_channels), nn.ReLU(inplace=True),
                         nn.Dropout(p=network.setting.dropout_rate),
                         network.conv2d(num_filters=num_filters, **kwargs))


class DenseLayer(nn.Module):
    def __init__(self, network, growth_rate, bottleneck_factor):
        super().__init__()
        self.bottleneck = _preact(network, bottleneck_factor * growth_rate, kernel_size=1)
        self.conv_block = _preact(network, growth_rate, kernel_size=3, padding=1)

    def forward(self, x):
        return torch.cat([x, self.conv_block(self.bottleneck(x))], dim=1)


class DenseBlock(nn.Module):
    def __init__(self, network, size, growth_rate, bottleneck_factor):
        super().__init__()
        layers = []
        for _ in range(size):
            width = network.in_channels
            layers.append(DenseLayer(network, growth_rate, bottleneck_factor))
            network.in_channels = width + growth_rate
        self.dense_block = nn.Sequential(*layers)

    def forward(self, x):
        return self.dense_block(x)


class TransitionBlock(nn.Sequential):
    def __init__(self, network, compression_factor):
        super().__init__()
        self.add_module("batch_norm", nn.BatchNorm2d(network.in_channels))
        self.add_module("relu", nn.ReLU(inplace=True))
        self.add_module("conv", network.conv2d(num_filters=network.in_channels // compression_factor,
                                               kernel_size=1))
        self.add_module("avg_pool", network.avgpool2d(kernel_size=2, stride=2))


class DenseNet(ZooModel):
    def make_feature_layers(self):
        growth, blocks, init = _CFG[str(self.setting.kind)]
        layers = self.stem(init)
        for i, size in enumerate(blocks):
            layers.append(DenseBlock(self, size, growth, bottleneck_factor=4))
            if i != len(blocks) - 1:
                layers.append(TransitionBlock(self, compression_factor=2))
        layers += [nn.BatchNorm2d(num_features=self.in_channels), nn.ReLU(inplace=True),
                   self.adapt_avgpool2d(output_size=1)]
        return nn.Sequential(*layers)
