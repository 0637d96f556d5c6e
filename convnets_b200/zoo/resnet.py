"""ResNet 18/34 (basic) and 26/50/101/152 (bottleneck, stride on the 3x3) — module tree and
state_dict keys of reference src/resnet.py:17-196 (config :29-36, stem :66-67, residual_block
:85-98, ResBlock :119-156, ResBottleneck :158-196, GAP :81, Dropout+Linear :104-107)."""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F

from .base import MultiClassBaseModel

# kind -> (bottleneck?, [(width, blocks, first stride)])
_STAGES = {
    "18": (False, (2, 2, 2, 2)), "34": (False, (3, 4, 6, 3)),
    "26": (True, (2, 2, 2, 2)), "50": (True, (3, 4, 6, 3)),
    "101": (True, (3, 4, 23, 3)), "152": (True, (3, 8, 36, 3)),
}
_WIDTHS = (64, 128, 256, 512)


class _Residual(nn.Module):
    """Shared forward of both block types: body(x) += shortcut(x); relu."""

    body_name = ""

    def forward(self, x):
        identity = x
        output = getattr(self, self.body_name)(x)
        if self.dim_synch is not None:
            identity = self.dim_synch(identity)
        output += identity
        return F.relu(output)


class ResBlock(_Residual):
    body_name = "res_block"

    def __init__(self, network, num_filters, expansion, dim_synch=None, **kwargs):
        super().__init__()
        self.dim_synch = dim_synch
        self.res_block = nn.Sequential(
            network.conv2d_block(num_filters=num_filters, kernel_size=3, padding=1, **kwargs),
            network.conv2d_block(num_filters=num_filters * expansion, activation=False,
                                 kernel_size=3, padding=1))


class ResBottleneck(_Residual):
    body_name = "res_bottleneck"

    def __init__(self, network, num_filters, expansion, dim_synch=None, **kwargs):
        super().__init__()
        self.dim_synch = dim_synch
        self.res_bottleneck = nn.Sequential(
            network.conv2d_block(num_filters=num_filters, kernel_size=1),
            network.conv2d_block(num_filters=num_filters, kernel_size=3, padding=1, **kwargs),
            network.conv2d_block(num_filters=num_filters * expansion, activation=False,
                                 kernel_size=1))


class ResNet(MultiClassBaseModel):
    def __init__(self, setting):
        super().__init__(setting)
        self.features = self.make_feature_layers()
        self.classifier = nn.Sequential(
            nn.Dropout(p=setting.dropout_rate),
            nn.Linear(self.num_flat_features(), setting.num_classes))
        if setting.init_params:
            self.init_params()

    def make_feature_layers(self):
        bottleneck, depths = _STAGES[str(self.setting.kind)]
        block, expansion = (ResBottleneck, 4) if bottleneck else (ResBlock, 1)
        layers = [self.conv2d_block(num_filters=64, kernel_size=7, stride=2, padding=3),
                  self.maxpool2d(kernel_size=3, stride=2, padding=1)]
        for stage, (width, depth) in enumerate(zip(_WIDTHS, depths)):
            for i in range(depth):
                stride = 2 if (i == 0 and stage > 0) else 1
                layers.append(self.residual_block(block, width, expansion, stride))
        layers.append(self.adapt_avgpool2d(output_size=1))
        return nn.Sequential(*layers)

    def residual_block(self, block_type, num_filters, expansion, stride=1):
        """1x1 strided projection (conv + BN, no ReLU) when the shape changes (resnet.py:90-93);
        the stride keyword is always forwarded to the block's strided conv."""
        dim_synch = None
        if stride != 1 or self.in_channels != num_filters * expansion:
            dim_synch = self.conv2d_block(self.in_channels, num_filters * expansion,
                                          set_output=False, activation=False, kernel_size=1,
                                          stride=stride)
        return block_type(self, num_filters=num_filters, expansion=expansion,
                          dim_synch=dim_synch, stride=stride)

    def forward(self, x):
        x = self.features(x)
        x = torch.flatten(x, start_dim=1)
        return self.classifier(x)
