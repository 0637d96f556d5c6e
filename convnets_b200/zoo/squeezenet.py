"""SqueezeNet 1.0 / 1.1 — module tree / state_dict keys of reference src/squeezenet.py:17-154:
un-padded stem conv and max-pools (:29-58), Fire = squeeze 1x1 -> cat(expand 1x1, expand 3x3)
(:127-153), classifier = Dropout -> 1x1 conv block -> GAP (:104-121)."""
from __future__ import annotations

import torch
import torch.nn as nn

from ._common import ZooModel

_F = {16: (16, 64, 64), 32: (32, 128, 128), 48: (48, 192, 192), 64: (64, 256, 256)}
_CFG = {
    "1.0": [("conv", 96, 7), "pool", 16, 16, 32, "pool", 32, 48, 48, 64, "pool", 64],
    "1.1": [("conv", 64, 3), "pool", 16, 16, "pool", 32, 32, "pool", 48, 48, 64, 64],
}


class Fire(nn.Module):
    def __init__(self, network, squeeze, expand1, expand3):
        super().__init__()
        self.squeeze = network.conv2d_block(num_filters=squeeze, kernel_size=1)
        self.expand_1x1 = network.conv2d_block(squeeze, expand1, set_output=False, kernel_size=1)
        self.expand_3x3 = network.conv2d_block(squeeze, expand3, kernel_size=3, padding=1)
        network.in_channels = expand1 + expand3

    def forward(self, x):
        x = self.squeeze(x)
        return torch.cat([self.expand_1x1(x), self.expand_3x3(x)], dim=1)


class SqueezeNet(ZooModel):
    def make_feature_layers(self):
        layers = []
        for e in _CFG[str(self.setting.kind)]:
            if e == "pool":
                layers.append(self.maxpool2d(kernel_size=3, stride=2))
            elif isinstance(e, tuple):
                layers.append(self.conv2d_block(num_filters=e[1], kernel_size=e[2], stride=2))
            else:
                layers.append(Fire(self, *_F[e]))
        return nn.Sequential(*layers)

    def make_classifier_layers(self):
        return nn.Sequential(nn.Dropout(p=self.setting.dropout_rate),
                             self.conv2d_block(num_filters=self.setting.num_classes, kernel_size=1),
                             self.adapt_avgpool2d(output_size=1))

    def forward(self, x):
        return torch.flatten(self.classifier(self.features(x)), start_dim=1)
