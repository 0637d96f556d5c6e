"""MobileNet-v1 — module tree / state_dict keys of reference src/mobilenet_v1.py:17-125: a 3x3/2
Conv2dBlock stem (:48) and thirteen depthwise-separable blocks (:49-62) whose children are named
conv_dw, bn_dw, relu_dw, conv_pw, bn_pw, relu_pw (:90-125)."""
from __future__ import annotations

import torch.nn as nn

from ._common import ZooModel
from .base import Conv2dBlock

# (pointwise output channels, depthwise stride)
_BLOCKS = ((64, 1), (128, 2), (128, 1), (256, 2), (256, 1), (512, 2), (512, 1), (512, 1), (512, 1),
           (512, 1), (512, 1), (1024, 2), (1024, 1))


class Conv2dBlockDW(nn.Sequential):
    def __init__(self, network, num_filters, **kwargs):
        super().__init__()
        channels = network.in_channels
        bn = network.setting.batch_norm
        self.add_module("conv_dw", network.conv2d_depthwise(channels, channels, **kwargs))
        if bn:
            self.add_module("bn_dw", nn.BatchNorm2d(num_features=channels))
        self.add_module("relu_dw", nn.ReLU(inplace=True))
        self.add_module("conv_pw", network.conv2d(channels, num_filters, kernel_size=1))
        if bn:
            self.add_module("bn_pw", nn.BatchNorm2d(num_features=num_filters))
        self.add_module("relu_pw", nn.ReLU(inplace=True))


class MobileNetV1(ZooModel):
    def make_feature_layers(self):
        layers = [Conv2dBlock(self, num_filters=32, kernel_size=3, stride=2, padding=1)]
        for width, stride in _BLOCKS:
            layers.append(Conv2dBlockDW(self, width, kernel_size=3, stride=stride, padding=1))
        layers.append(self.adapt_avgpool2d(output_size=1))
        return nn.Sequential(*layers)
