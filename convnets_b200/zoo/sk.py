"""SK-ResNet and SKNet 26/50/101/152 — module trees / state_dict keys of reference
src/sk_resnet.py:17-240 and src/sknet.py:17-242.  SKConv (:173-242): two grouped 3x3 Conv2dBlocks
(dilation 1 and 2, padding = dilation, 32 groups, shared stride) -> element-wise sum -> GAP ->
descriptor Conv2dBlock 1x1 (d = max(C/16, 32)) -> two 1x1 convs WITH bias -> softmax over the two
paths -> weighted sum."""
from __future__ import annotations

import torch
import torch.nn as nn

from ._common import DEPTHS, Residual, ZooModel
from .base import Conv2dBlock


class SKConv(nn.Module):
    def __init__(self, network, num_paths=2, groups=32, reduction=16, min_descriptor=32, **kwargs):
        super().__init__()
        c = network.in_channels
        self.num_paths = num_paths
        self.num_features = c
        d = max(c // reduction, min_descriptor)
        self.kernels = nn.ModuleList()
        self.attentions = nn.ModuleList()
        for i in range(num_paths):
            self.kernels.append(Conv2dBlock(network, num_filters=c, set_output=False, kernel_size=3,
                                            padding=1 + i, dilation=1 + i, groups=groups, **kwargs))
            self.attentions.append(nn.Sequential(nn.Conv2d(d, c, kernel_size=1)))
        self.global_avgpool = network.adapt_avgpool2d(set_output=False, output_size=1)
        self.descriptor = Conv2dBlock(network, num_filters=d, set_output=False, kernel_size=1)
        self.softmax = nn.Softmax(dim=1)

    def forward(self, x):
        fm = torch.cat([k(x) for k in self.kernels], dim=1)
        b, _, h, w = fm.size()
        fm = fm.view(b, self.num_paths, self.num_features, h, w)
        z = self.descriptor(self.global_avgpool(torch.sum(fm, dim=1)))
        a = torch.cat([att(z) for att in self.attentions], dim=1)
        a = self.softmax(a.view(b, self.num_paths, self.num_features, 1, 1))
        return torch.sum(fm * a, dim=1)


def _body(network, width, expansion, **kwargs):
    return nn.Sequential(
        Conv2dBlock(network, num_filters=width, kernel_size=1),
        SKConv(network, **kwargs),
        Conv2dBlock(network, num_filters=width * expansion, activation=False, kernel_size=1))


class SKResBottleneck(Residual):
    body_name = "res_bottleneck"

    def __init__(self, network, num_filters, expansion, dim_synch=None, **kwargs):
        super().__init__()
        self.dim_synch = dim_synch
        self.res_bottleneck = _body(network, num_filters, expansion, **kwargs)


class SKBottleneck(Residual):
    body_name = "sk_bottleneck"

    def __init__(self, network, num_filters, expansion, dim_synch=None, **kwargs):
        super().__init__()
        self.dim_synch = dim_synch
        self.sk_bottleneck = _body(network, num_filters, expansion, **kwargs)


class _SKFamily(ZooModel):
    block, widths, expansion, track_shape, extra = None, (), 1, False, {}

    def make_feature_layers(self):
        kind = str(self.setting.kind)
        if kind not in ("26", "50", "101", "152"):
            raise ValueError(f"{type(self).__name__} kind {kind!r} is not supported")

        def block(width, exp, proj, stride):
            return self.block(self, width, exp, proj, stride=stride, **self.extra)

        return nn.Sequential(*self.stem(),
                             *self.stages(block, self.widths, DEPTHS[kind], self.expansion,
                                          track_shape=self.track_shape),
                             self.adapt_avgpool2d(output_size=1))


class SKResNet(_SKFamily):
    block, widths, expansion, track_shape = SKResBottleneck, (64, 128, 256, 512), 4, False


class SKNet(_SKFamily):
    block, widths, expansion, track_shape = SKBottleneck, (128, 256, 512, 1024), 2, True
    extra = {"groups": 32}
