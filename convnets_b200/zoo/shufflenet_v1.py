"""ShuffleNet-v1 g1/g2/g3/g4/g8 — module tree / state_dict keys of reference
src/shufflenet_v1.py:17-190: 3x3/2 stem (24 ch) + max-pool (:52-53), three stages of ShuffleUnits
(:30-36,:56-62).  A unit (:96-171) is grouped 1x1 compress -> channel shuffle -> depthwise 3x3
(BN, no ReLU) -> grouped 1x1 expand (BN, no ReLU); downsampling units emit C_out - C_in channels and
concatenate the 3x3/2 average-pooled input in FRONT (:115-116,:148,:167); the first unit's compress
conv is ungrouped (:111) but its shuffle still uses `groups` (:162,:179)."""
from __future__ import annotations

import torch
import torch.nn as nn

from ._common import ZooModel
from .base import Conv2dBlock

_OUT = {"g1": 144, "g2": 200, "g3": 240, "g4": 272, "g8": 384}
_REPEATS = (3, 7, 3)


class ShuffleUnit(nn.Module):
    def __init__(self, network, num_output_channels, groups, stride, downsample, first_conv):
        super().__init__()
        self.groups = groups
        self.groups_first_conv1x1 = 1 if first_conv else groups
        self.downsample = downsample
        bottleneck = num_output_channels // 4
        identity_channels = network.in_channels
        branch_out = num_output_channels - identity_channels if downsample else num_output_channels
        self.conv1x1_group_compress = Conv2dBlock(network, num_filters=bottleneck, kernel_size=1,
                                                  groups=self.groups_first_conv1x1)
        self.conv3x3_depthwise = Conv2dBlock(network, num_filters=bottleneck, activation=False,
                                             kernel_size=3, stride=stride, padding=1,
                                             groups=bottleneck)
        self.conv1x1_group_expand = Conv2dBlock(network, num_filters=branch_out, activation=False,
                                                kernel_size=1, groups=groups)
        if downsample:
            network.in_channels += identity_channels
        self.avg_pool = network.avgpool2d(set_output=False, kernel_size=3, stride=2, padding=1)
        self.relu = nn.ReLU(inplace=True)

    def channel_shuffle(self, x):
        b, c, h, w = x.size()
        x = x.view(b, self.groups, c // self.groups, h, w)
        return torch.transpose(x, 1, 2).contiguous().view(b, -1, h, w)

    def forward(self, x):
        identity = self.avg_pool(x) if self.downsample else x
        output = self.conv1x1_group_compress(x)
        output = self.channel_shuffle(output)
        output = self.conv3x3_depthwise(output)
        output = self.conv1x1_group_expand(output)
        if self.downsample:
            output = torch.cat([identity, output], dim=1)
        else:
            output += identity
        return self.relu(output)


class ShuffleNetV1(ZooModel):
    def make_feature_layers(self):
        kind = str(self.setting.kind)
        groups = int(kind[1:])
        layers = [self.conv2d_block(num_filters=24, kernel_size=3, stride=2, padding=1),
                  self.maxpool2d(kernel_size=3, stride=2, padding=1)]
        for stage, repeat in enumerate(_REPEATS):
            width = _OUT[kind] << stage
            for j in range(1 + repeat):
                down = j == 0
                layers.append(ShuffleUnit(self, width, groups, 2 if down else 1, down,
                                          first_conv=(stage == 0 and j == 0)))
        layers.append(self.adapt_avgpool2d(output_size=1))
        return nn.Sequential(*layers)
