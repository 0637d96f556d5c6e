"""VGGNet 11/13/16/19 with BatchNorm, GAP and a three-Linear head — module tree / state_dict keys
of reference src/vggnet.py:17-101 (config :30-35, features :55-72, classifier :74-92)."""
from __future__ import annotations

import torch.nn as nn

from ._common import ZooModel

_CFG = {
    "11": (1, 1, 2, 2, 2), "13": (2, 2, 2, 2, 2), "16": (2, 2, 3, 3, 3), "19": (2, 2, 4, 4, 4),
}
_WIDTHS = (64, 128, 256, 512, 512)


class VGGNet(ZooModel):
    def make_feature_layers(self):
        layers = []
        for width, reps in zip(_WIDTHS, _CFG[str(self.setting.kind)]):
            layers += [self.conv2d_block(num_filters=width, kernel_size=3, padding=1)
                       for _ in range(reps)]
            layers.append(self.maxpool2d(kernel_size=2, stride=2))
        layers.append(self.adapt_avgpool2d(output_size=1))
        return nn.Sequential(*layers)

    def make_classifier_layers(self):
        p = self.setting.dropout_rate
        return nn.Sequential(
            nn.Dropout(p=p), nn.Linear(self.num_flat_features(), 4096), nn.ReLU(inplace=True),
            nn.Dropout(p=p), nn.Linear(4096, 4096), nn.ReLU(inplace=True),
            nn.Dropout(p=p), nn.Linear(4096, self.setting.num_classes))
