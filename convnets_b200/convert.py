"""convert(model): switch a ConvNets zoo model over to libconvnets_b200.

The reference builds every network from torch.nn layers through six factories
(reference src/basemodel.py:966-1067).  `convert` walks the finished module tree and
  * re-classes nn.Conv2d / BatchNorm2d / ReLU / MaxPool2d / AvgPool2d / AdaptiveAvgPool2d instances
    to the library-backed subclasses of convnets_b200.layers (same Parameter objects, same module
    names, same state_dict keys — `load_state_dict(strict=True)`, src/basemodel.py:940, still works
    and an optimizer created afterwards sees the same parameters);
  * fuses `BatchNorm2d -> ReLU(inplace)` pairs inside the conv2d_block Sequentials
    (src/basemodel.py:1004-1015) into one kernel (the ReLU slot becomes an Identity so the child
    indices and hence the state_dict keys do not move);
  * replaces the forward of the composite blocks whose tail is element-wise
    (`output += identity; F.relu(output)`, src/resnet.py:151-154,191-194 and the equivalents in the
    other families) by a version that calls the fused kernels.

It works on the reference's own classes (matched by class name; /root/reference is never edited)
and on the mirrors in convnets_b200.zoo.
"""
from __future__ import annotations

import types

import torch
import torch.nn as nn

from . import functional as F_
from . import layers as L

_CL = torch.channels_last

_SWAP = {
    nn.Conv2d: L.Conv2d,
    nn.BatchNorm2d: L.BatchNorm2d,
    nn.ReLU: L.ReLU,
    nn.MaxPool2d: L.MaxPool2d,
    nn.AvgPool2d: L.AvgPool2d,
    nn.AdaptiveAvgPool2d: L.AdaptiveAvgPool2d,
}


# ------------------------------------------------------------------------------------------------
# fused block forwards (bound to instances by class name)
def _tail(last: nn.Module):
    """Split a conv2d_block(activation=False) Sequential into (conv, bn-or-None)."""
    mods = [m for m in last.children() if not isinstance(m, (nn.Identity,))]
    conv = mods[0]
    bn = mods[1] if len(mods) > 1 and isinstance(mods[1], nn.BatchNorm2d) else None
    return conv, bn


def _residual_forward(body_name: str):
    def forward(self, x):
        """body(x) += shortcut(x); relu — the residual add and the ReLU run inside the last
        BatchNorm kernel of the body (reference src/resnet.py:139-156,176-196)."""
        blocks = list(getattr(self, body_name).children())
        out = x
        for blk in blocks[:-1]:
            out = blk(out)
        identity = x if self.dim_synch is None else self.dim_synch(x)
        conv, bn = _tail(blocks[-1])
        out = conv(out)
        if bn is not None:
            return bn(out, residual=identity, relu=True)
        return F_.add_relu(out, identity)
    return forward


_BLOCK_FORWARDS = {
    "ResBlock": _residual_forward("res_block"),
    "ResBottleneck": _residual_forward("res_bottleneck"),
}


def register_block_forward(class_name: str, fn) -> None:
    """Extension point used by convnets_b200.zoo for the other families' composite blocks."""
    _BLOCK_FORWARDS[class_name] = fn


# ------------------------------------------------------------------------------------------------
def _fuse_bn_relu(seq: nn.Sequential) -> None:
    names = list(seq._modules.keys())
    for a, b in zip(names, names[1:]):
        ma, mb = seq._modules[a], seq._modules[b]
        if isinstance(ma, L.BatchNorm2d) and type(mb) in (nn.ReLU, L.ReLU):
            ma.fuse_relu = True
            seq._modules[b] = L.Identity()


def convert(model: nn.Module, dtype: torch.dtype = torch.bfloat16, fuse: bool = True) -> nn.Module:
    """In-place conversion; returns `model`.  `dtype` is the arithmetic / storage type of the
    feature maps (bf16: tcgen05 tensor cores with fp32 accumulate; fp32: "fp32-IO" mode on CUDA
    cores).  Parameters stay fp32 masters."""
    if dtype not in (torch.bfloat16, torch.float32):
        raise TypeError("convert: dtype must be torch.bfloat16 or torch.float32")
    first_conv = None
    for m in model.modules():
        cls = _SWAP.get(type(m))
        if cls is not None:
            m.__class__ = cls
        if isinstance(m, L.Conv2d):
            m.compute_dtype = dtype
            if first_conv is None:
                first_conv = m
            w = m.weight
            if w.dim() == 4 and not w.data.is_contiguous(memory_format=_CL):
                # KRSC storage for the master weight: the bf16 operand copy is then a pure cast and
                # the fp32 weight gradient the library writes already has the parameter's layout
                w.data = w.data.contiguous(memory_format=_CL)
    if first_conv is not None:
        first_conv.cast_input = True
    if fuse:
        for m in model.modules():
            if isinstance(m, nn.Sequential):
                _fuse_bn_relu(m)
            fn = _BLOCK_FORWARDS.get(type(m).__name__)
            if fn is not None:
                m.forward = types.MethodType(fn, m)
    model._cnb_dtype = dtype
    return model
