"""SE-ResNet 18/34/26/50/101/152 and SENet 26/50/101/152 — module trees / state_dict keys of
reference src/se_resnet.py:17-255 and src/senet.py:17-255.  Blocks are named-children Conv2dBlocks
followed by an SEUnit (GAP -> Linear(C, C/16, no bias) -> ReLU -> Linear(C/16, C, no bias) ->
Sigmoid -> scale, :216-255); SENet is the ResNeXt-shaped variant (widths 128..1024, expansion 2,
cardinality 32, body attribute `se_bottleneck`, projection seen by the shape tracker :93-101)."""
from __future__ import annotations

import torch.nn as nn

from ._common import DEPTHS, Residual, ZooModel
from .base import Conv2dBlock


class SEUnit(nn.Module):
    def __init__(self, network, reduction):
        super().__init__()
        c = network.in_channels
        self.squeeze = network.adapt_avgpool2d(set_output=False, output_size=1)
        self.excitation = nn.Sequential(nn.Linear(c, c // reduction, bias=False), nn.ReLU(inplace=True),
                                        nn.Linear(c // reduction, c, bias=False), nn.Sigmoid())

    def forward(self, x):
        b, c = x.size(0), x.size(1)
        s = self.excitation(self.squeeze(x).view(b, c)).view(b, c, 1, 1)
        return x * s.expand_as(x)


def _basic(network, width, expansion, reduction, **kwargs):
    return nn.Sequential(
        Conv2dBlock(network, num_filters=width, kernel_size=3, padding=1, **kwargs),
        Conv2dBlock(network, num_filters=width * expansion, activation=False, kernel_size=3, padding=1),
        SEUnit(network, reduction))


def _bottleneck(network, width, expansion, reduction, **kwargs):
    return nn.Sequential(
        Conv2dBlock(network, num_filters=width, kernel_size=1),
        Conv2dBlock(network, num_filters=width, kernel_size=3, padding=1, **kwargs),
        Conv2dBlock(network, num_filters=width * expansion, activation=False, kernel_size=1),
        SEUnit(network, reduction))


class SEResBlock(Residual):
    body_name = "res_block"

    def __init__(self, network, num_filters, expansion, reduction, dim_synch=None, **kwargs):
        super().__init__()
        self.dim_synch = dim_synch
        self.res_block = _basic(network, num_filters, expansion, reduction, **kwargs)


class SEResBottleneck(Residual):
    body_name = "res_bottleneck"

    def __init__(self, network, num_filters, expansion, reduction, dim_synch=None, **kwargs):
        super().__init__()
        self.dim_synch = dim_synch
        self.res_bottleneck = _bottleneck(network, num_filters, expansion, reduction, **kwargs)


class SEBottleneck(Residual):
    body_name = "se_bottleneck"

    def __init__(self, network, num_filters, expansion, reduction, dim_synch=None, **kwargs):
        super().__init__()
        self.dim_synch = dim_synch
        self.se_bottleneck = _bottleneck(network, num_filters, expansion, reduction, **kwargs)


class SEResNet(ZooModel):
    reduction = 16

    def make_feature_layers(self):
        kind = str(self.setting.kind)
        bottleneck = kind not in ("18", "34")
        cls, expansion = (SEResBottleneck, 4) if bottleneck else (SEResBlock, 1)

        def block(width, exp, proj, stride):
            return cls(self, width, exp, self.reduction, proj, stride=stride)

        return nn.Sequential(*self.stem(), *self.stages(block, (64, 128, 256, 512), DEPTHS[kind], expansion),
                             self.adapt_avgpool2d(output_size=1))


class SENet(ZooModel):
    reduction = 16
    cardinality = 32

    def make_feature_layers(self):
        kind = str(self.setting.kind)
        if kind not in ("26", "50", "101", "152"):
            raise ValueError(f"SENet kind {kind!r} is not supported")

        def block(width, exp, proj, stride):
            return SEBottleneck(self, width, exp, self.reduction, proj, stride=stride,
                                groups=self.cardinality)

        return nn.Sequential(*self.stem(),
                             *self.stages(block, (128, 256, 512, 1024), DEPTHS[kind], 2, track_shape=True),
                             self.adapt_avgpool2d(output_size=1))
