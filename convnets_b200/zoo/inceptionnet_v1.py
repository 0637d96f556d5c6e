"""InceptionNet-v1 (GoogLeNet layout with BatchNorm, no auxiliary heads) — module tree /
state_dict keys of reference src/inceptionnet_v1.py:17-143: stem (:45-49), nine InceptionBlocks with
3x3/2 max-pools between the stages (:52-66), each block = cat(1x1, 1x1->3x3, 1x1->5x5,
maxpool3x3/1->1x1) (:104-142)."""
from __future__ import annotations

import torch
import torch.nn as nn

from ._common import ZooModel

# (1x1, 3x3 reduce, 3x3, 5x5 reduce, 5x5, pool projection); None = 3x3/2 max-pool
_BLOCKS = [(64, 96, 128, 16, 32, 32), (128, 128, 192, 32, 96, 64), None,
           (192, 96, 208, 16, 48, 64), (160, 112, 224, 24, 64, 64), (128, 128, 256, 24, 64, 64),
           (112, 144, 288, 32, 64, 64), (256, 160, 320, 32, 128, 128), None,
           (256, 160, 320, 32, 128, 128), (384, 192, 384, 48, 128, 128)]


class InceptionBlock(nn.Module):
    def __init__(self, network, c1, c3r, c3, c5r, c5, cp):
        super().__init__()
        cin = network.in_channels

        def cb(i, o, **kw):
            return network.conv2d_block(i, o, set_output=False, **kw)

        self.branch1 = cb(cin, c1, kernel_size=1)
        self.branch2 = nn.Sequential(cb(cin, c3r, kernel_size=1), cb(c3r, c3, kernel_size=3, padding=1))
        self.branch3 = nn.Sequential(cb(cin, c5r, kernel_size=1), cb(c5r, c5, kernel_size=5, padding=2))
        self.branch4 = nn.Sequential(network.maxpool2d(kernel_size=3, stride=1, padding=1),
                                     cb(cin, cp, kernel_size=1))
        network.in_channels = c1 + c3 + c5 + cp

    def forward(self, x):
        return torch.cat([self.branch1(x), self.branch2(x), self.branch3(x), self.branch4(x)], dim=1)


class InceptionNetV1(ZooModel):
    def make_feature_layers(self):
        layers = self.stem(64)
        layers += [self.conv2d_block(num_filters=64, kernel_size=1),
                   self.conv2d_block(num_filters=192, kernel_size=3, padding=1),
                   self.maxpool2d(kernel_size=3, stride=2, padding=1)]
        for cfg in _BLOCKS:
            layers.append(self.maxpool2d(kernel_size=3, stride=2, padding=1) if cfg is None
                          else InceptionBlock(self, *cfg))
        layers.append(self.adapt_avgpool2d(output_size=1))
        return nn.Sequential(*layers)
