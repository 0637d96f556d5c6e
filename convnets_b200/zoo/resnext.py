"""ResNeXt 26/50/101/152 — module tree / state_dict keys of reference src/resnext.py:17-190:
widths 128..1024 (:29-36), expansion 2, cardinality 32 on the 3x3 (:84-95,:168) and NO global
average pool before the classifier (:66-80), so the Linear sees C*H*W features in NCHW order.
The basic-block kinds 18/34 of the reference are broken and are not
mirrored."""
from __future__ import annotations

import torch.nn as nn

from ._common import DEPTHS, Residual, ZooModel


class ResBottleneck(Residual):
    body_name = "res_bottleneck"

    def __init__(self, network, num_filters, expansion, dim_synch=None, **kwargs):
        super().__init__()
        self.dim_synch = dim_synch
        self.res_bottleneck = nn.Sequential(
            network.conv2d_block(num_filters=num_filters, kernel_size=1),
            network.conv2d_block(num_filters=num_filters, kernel_size=3, padding=1, **kwargs),
            network.conv2d_block(num_filters=num_filters * expansion, activation=False,
                                 kernel_size=1))


class ResNeXt(ZooModel):
    cardinality = 32

    def make_feature_layers(self):
        kind = str(self.setting.kind)
        if kind not in ("26", "50", "101", "152"):
            raise ValueError(f"ResNeXt kind {kind!r} is not supported (bottleneck kinds only)")

        def block(width, expansion, proj, stride):
            return ResBottleneck(self, width, expansion, proj, stride=stride,
                                 groups=self.cardinality)

        return nn.Sequential(*self.stem(), *self.stages(block, (128, 256, 512, 1024), DEPTHS[kind], 2))
