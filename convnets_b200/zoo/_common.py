"""Shared pieces of the zoo mirrors: the Dropout+Linear head, the `features -> flatten ->
classifier` forward, and the residual stage builder used by the ResNet-shaped families."""
from __future__ import annotations

import torch
import torch.nn as nn

from .base import MultiClassBaseModel


class ZooModel(MultiClassBaseModel):
    """features / classifier model skeleton common to every reference family
    (e.g. reference src/resnet.py:38-51,110-117)."""

    def __init__(self, setting):
        super().__init__(setting)
        self.features = self.make_feature_layers()
        self.classifier = self.make_classifier_layers()
        if setting.init_params:
            self.init_params()

    def make_classifier_layers(self):
        return nn.Sequential(nn.Dropout(p=self.setting.dropout_rate),
                             nn.Linear(self.num_flat_features(), self.setting.num_classes))

    def forward(self, x):
        x = self.features(x)
        x = torch.flatten(x, start_dim=1)
        return self.classifier(x)

    # 7x7/2 conv block + 3x3/2 max-pool opening of the ImageNet-style families
    def stem(self, num_filters=64):
        return [self.conv2d_block(num_filters=num_filters, kernel_size=7, stride=2, padding=3),
                self.maxpool2d(kernel_size=3, stride=2, padding=1)]

    def projection(self, out_channels, stride, track_shape=False):
        """1x1 (strided) conv + BN shortcut, or None when the shapes already agree.  With
        `track_shape` the shape tracker sees the projection but the channel count is restored
        (reference src/senet.py:93-101, src/sknet.py:92-100)."""
        if stride == 1 and self.in_channels == out_channels:
            return None
        keep = self.in_channels
        proj = self.conv2d_block(self.in_channels, out_channels, set_output=track_shape,
                                 activation=False, kernel_size=1, stride=stride)
        self.in_channels = keep
        return proj

    def stages(self, make_block, widths, depths, expansion, track_shape=False):
        """Four residual stages; the first block of stages 2-4 has stride 2."""
        layers = []
        for stage, (width, depth) in enumerate(zip(widths, depths)):
            for i in range(depth):
                stride = 2 if (i == 0 and stage > 0) else 1
                proj = self.projection(width * expansion, stride, track_shape)
                layers.append(make_block(width, expansion, proj, stride))
        return layers


class Residual(nn.Module):
    """body(x) += shortcut(x); relu — the forward every residual block of the zoo shares."""

    body_name = ""

    def forward(self, x):
        identity = x
        output = getattr(self, self.body_name)(x)
        if self.dim_synch is not None:
            identity = self.dim_synch(identity)
        output += identity
        return torch.relu(output)


DEPTHS = {"18": (2, 2, 2, 2), "34": (3, 4, 6, 3), "26": (2, 2, 2, 2), "50": (3, 4, 6, 3),
          "101": (3, 4, 23, 3), "152": (3, 8, 36, 3)}
