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

import os
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


_BODY_NAMES = ("res_block", "res_bottleneck", "se_bottleneck", "sk_bottleneck")


# Identity-shortcut blocks: fold the skip gradient into the first conv's data-grad epilogue
# (functional._Conv2dSkipFn) instead of letting autograd add the two gradients of the block input
# (tests/test_gpu_fused_paths.py::test_dgrad_add_epilogue).  CNB_FUSED_SKIP_GRAD=0 = autograd's add.
_FUSED_SKIP = os.environ.get("CNB_FUSED_SKIP_GRAD", "1") == "1"


def _can_fold_skip(blocks, x) -> bool:
    first = blocks[0]
    if not (isinstance(first, nn.Sequential) and len(blocks) > 1 and torch.is_grad_enabled()
            and x.requires_grad and x.dtype == torch.bfloat16):
        return False
    conv = next(iter(first.children()), None)
    return (isinstance(conv, L.Conv2d) and not conv.cast_input and conv.stride in (1, (1, 1))
            and conv.padding_mode == "zeros" and not isinstance(conv.padding, str))


def _is_conv_block(m) -> bool:
    return isinstance(m, L.Conv2d) or (isinstance(m, nn.Sequential)
                                       and isinstance(next(iter(m.children()), None), L.Conv2d))


def _sole(t, consumer):
    """Mark `t` as read by `consumer` only — true for the tensors handed from one child of a block body
    to the next — when that consumer is a plain convolution (block): its data-grad may then form the
    backward sums of the BatchNorm that produced t (functional._dgrad_with_bn_sums)."""
    if _is_conv_block(consumer) and hasattr(t, "_cnb_bn_holder"):
        t._cnb_sole = True
    return t


def _residual_forward_folded(self, blocks, x):
    first = list(blocks[0].children())
    # x reaches this block only (the parent Sequential hands it over) and is read through the
    # two-output convolution below, whose data-grad forms the COMPLETE gradient of x
    out, identity = first[0](_sole(x, first[0]), with_skip=True)
    for m in first[1:]:
        out = m(out)
    last = blocks[-1]
    if isinstance(last, nn.Sequential) and isinstance(next(iter(last.children()), None), nn.Conv2d):
        for blk in blocks[1:-1]:
            out = blk(_sole(out, blk))
        conv, bn = _tail(last)
        out = conv(_sole(out, conv))
        if bn is not None:
            # `identity` is the alias _Conv2dSkipFn handed out and nothing else reads it: its
            # gradient can stay implicit (block-output gradient + the ReLU bit mask)
            return bn(out, residual=identity, relu=True, lazy_residual_grad=True)
        return F_.add_relu(out, identity)
    for blk in blocks[1:]:
        out = blk(out)
    return F_.add_relu(out, identity)


def _residual_forward(self, x):
    """body(x) += shortcut(x); relu.  When the body ends in a conv block, the residual add and the
    ReLU run inside that block's BatchNorm kernel; when it ends in an SE unit they are one
    add+ReLU kernel (reference src/resnet.py:139-156,176-196; src/se_resnet.py:199-211;
    src/sknet.py:160-171)."""
    body = next(getattr(self, n) for n in _BODY_NAMES if hasattr(self, n))
    blocks = list(body.children())
    if _FUSED_SKIP and self.dim_synch is None and _can_fold_skip(blocks, x):
        return _residual_forward_folded(self, blocks, x)
    identity = x if self.dim_synch is None else self.dim_synch(x)
    last = blocks[-1]
    if isinstance(last, nn.Sequential) and isinstance(next(iter(last.children()), None), nn.Conv2d):
        out = x
        for i, blk in enumerate(blocks[:-1]):
            out = blk(out if i == 0 else _sole(out, blk))   # x itself also feeds the shortcut
        conv, bn = _tail(last)
        out = conv(_sole(out, conv) if len(blocks) > 1 else out)
        if bn is not None:
            return bn(out, residual=identity, relu=True)
        return F_.add_relu(out, identity)
    return F_.add_relu(body(x), identity)


def _se_unit_forward(self, x):
    """GAP -> excitation MLP (fp32, two tiny bias-free Linears) -> x * s  (src/se_resnet.py:244-255)"""
    b, c = x.size(0), x.size(1)
    s = self.excitation(F_.global_avg_pool2d(x).view(b, c))
    return F_.se_scale(x, s.view(b, c, 1, 1))


def _sk_conv_forward(self, x):
    """Two grouped paths -> fused sum+GAP -> descriptor -> attention logits -> fused softmax-select
    (src/sknet.py:222-242).  The descriptor / attention convs see [B, C, 1, 1] fp32 vectors."""
    if self.num_paths != 2:
        raise NotImplementedError("convnets_b200 SKConv: two paths (the reference's configuration)")
    u0, u1 = self.kernels[0](x), self.kernels[1](x)
    z = self.descriptor(F_.sk_fuse_gap(u0, u1))
    logits = torch.cat([att(z) for att in self.attentions], dim=1)
    return F_.sk_select(u0, u1, logits)


def _shuffle_unit_forward(self, x):
    """src/shufflenet_v1.py:150-171 with the shuffle, concat and add+ReLU in library kernels."""
    out = self.conv1x1_group_compress(x)
    out = F_.channel_shuffle(out, self.groups)
    out = self.conv3x3_depthwise(out)
    out = self.conv1x1_group_expand(out)
    if self.downsample:
        return F_.relu(F_.cat_channels([self.avg_pool(x), out]))
    return F_.add_relu(out, x)


def _fire_forward(self, x):
    x = self.squeeze(x)
    return F_.cat_channels([self.expand_1x1(x), self.expand_3x3(x)])


def _inception_forward(self, x):
    return F_.cat_channels([self.branch1(x), self.branch2(x), self.branch3(x), self.branch4(x)])


def _dense_layer_forward(self, x):
    return F_.cat_channels([x, self.conv_block(self.bottleneck(x))])


_BLOCK_FORWARDS = {name: _residual_forward for name in (
    "ResBlock", "ResBottleneck", "SEResBlock", "SEResBottleneck", "SEBlock", "SEBottleneck",
    "SKResBottleneck", "SKBottleneck")}
def _dense_block_forward(self, x):
    from .dense import dense_block_forward
    return dense_block_forward(self, x)


_BLOCK_FORWARDS.update({
    "DenseBlock": _dense_block_forward,
    "SEUnit": _se_unit_forward, "SKConv": _sk_conv_forward, "ShuffleUnit": _shuffle_unit_forward,
    "Fire": _fire_forward, "InceptionBlock": _inception_forward, "DenseLayer": _dense_layer_forward,
})


def register_block_forward(class_name: str, fn) -> None:
    """Extension point used by convnets_b200.zoo for the other families' composite blocks."""
    _BLOCK_FORWARDS[class_name] = fn


# ------------------------------------------------------------------------------------------------
def _fuse_bn_relu(seq: nn.Sequential) -> None:
    names = list(seq._modules.keys())
    for a, b in zip(names, names[1:]):
        ma, mb = seq._modules[a], seq._modules[b]
        if isinstance(ma, L.Conv2d) and isinstance(mb, L.BatchNorm2d):
            ma.emit_bn_stats = True  # conv -> bn: statistics come out of the conv epilogue
            # inference: the conv folds this BatchNorm into its filter (not registered as a child:
            # the module tree and the state_dict keys must not change)
            ma.__dict__["_cnb_bn"] = mb
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
    if dtype == torch.bfloat16:
        dense = [m for m in model.modules() if isinstance(m, L.Conv2d) and m.groups == 1
                 and m.weight.shape[0] % 8 == 0 and m.weight.shape[1] % 8 == 0]
        if dense:
            group = L.PackGroup(dense)
            for m in dense:
                m.__dict__["_cnb_group"] = group
    if fuse:
        for m in model.modules():
            if isinstance(m, nn.Sequential):
                _fuse_bn_relu(m)
            fn = _BLOCK_FORWARDS.get(type(m).__name__)
            if fn is not None:
                m.forward = types.MethodType(fn, m)
    head = getattr(model, "classifier", None)
    if isinstance(head, nn.Module) and any(isinstance(m, nn.Linear) for m in head.modules()):
        # the Linear head keeps fp32 parameters (stays torch); models without
        # a global average pool (ResNeXt, src/resnext.py:66-80) hand it bf16 feature maps
        head.register_forward_pre_hook(lambda mod, args: tuple(
            a.float() if torch.is_tensor(a) and a.dtype == torch.bfloat16 else a for a in args))
    model._cnb_dtype = dtype
    opt = getattr(model, "optimizer", None)
    if hasattr(opt, "attach"):
        opt.attach(model)   # optim.FusedAdam built before convert(): let it maintain the operand copies
    return model
