"""Concat-aware DenseNet blocks.

The reference's DenseLayer ends in `torch.cat([x, output], dim=1)` (src/densenet.py:131-140): layer
l of a block re-copies the C_l channels it received next to the 32 it produced, its BatchNorm
re-reads the whole prefix for statistics it shares with every other layer, and autograd splits and
re-sums the prefix gradient layer by layer.  Here a block keeps ONE [N, C_total, H, W] buffer:

  forward   each 3x3 conv's 32 output channels are copied into their slot once (a 32-channel slab,
            not the prefix); the batch statistics of every channel are computed once — for the block
            input by layer 0's ordinary BatchNorm, for produced channels from the partial sums the
            conv epilogue already emits — and layer l's BatchNorm only APPLIES (its gamma / beta, its
            running statistics) to the row-strided prefix (`cnb_bn_fwd_train_prefix`);
  backward  one gradient buffer of the same shape; layer l's BatchNorm adds its input gradient to
            the prefix in place (`cnb_bn_bwd_prefix`), the 3x3 conv reads its slot.

The whole block is one torch.autograd.Function (an in-place buffer shared by 6-24 layers is not
something the tape can express); inside it the existing per-op functions run with a stand-in context
object, so every kernel call is the validated one.  bf16 training only: fp32-IO mode, inference and
no-grad forwards keep the per-layer path (functional.cat_channels).
"""
from __future__ import annotations

import ctypes as C
import os

import torch
import torch.nn as nn

from . import functional as F_
from . import layers as L
from . import lib as _L

_CL = torch.channels_last
_ENABLED = os.environ.get("CNB_DENSE_BUFFER", "1") == "1"


class _Ctx:
    """Stand-in for the autograd context when a Function's forward / backward is called directly."""

    def __init__(self, n_inputs=16):
        self.needs_input_grad = (True,) * n_inputs
        self.saved_tensors = ()

    def save_for_backward(self, *ts):
        self.saved_tensors = ts

    def mark_non_differentiable(self, *ts):
        pass

    def set_materialize_grads(self, value):
        pass


def _p(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _copy_channels(src, src_c, src_off, dst, dst_c, dst_off, m, ncopy):
    lib = _L.load()
    _L.check(lib.cnb_copy_channels(_p(src), src_c, src_off, _p(dst), dst_c, dst_off, m, ncopy,
                                   F_._dt(src), _stream()), "cnb_copy_channels")


def _split(seq: nn.Sequential):
    """(BatchNorm2d, dropout p, Conv2d) of a pre-activation Sequential BN-ReLU-Dropout-conv."""
    bn = conv = None
    p = 0.0
    for m in seq.children():
        if isinstance(m, L.BatchNorm2d):
            bn = m
        elif isinstance(m, nn.Dropout):
            p = float(m.p)
        elif isinstance(m, L.Conv2d):
            conv = m
    return bn, p, conv


def eligible(block, x) -> bool:
    if not (_ENABLED and x.is_cuda and x.dtype == torch.bfloat16 and block.training
            and torch.is_grad_enabled()):
        return False
    for layer in block.dense_block.children():
        for seq in (layer.bottleneck, layer.conv_block):
            bn, _, conv = _split(seq)
            if bn is None or conv is None or not bn.fuse_relu or not bn.track_running_stats \
                    or bn.momentum is None or bn.weight is None or conv.groups != 1:
                return False
            if conv.weight.shape[0] % 8 or conv.weight.shape[1] % 8 or conv.stride not in (1, (1, 1)):
                return False
    return True


def _block_params(block):
    ps = []
    for layer in block.dense_block.children():
        for seq in (layer.bottleneck, layer.conv_block):
            bn, _, conv = _split(seq)
            ps += [bn.weight, bn.bias, conv.weight]
            if conv.bias is not None:
                ps.append(conv.bias)
    return ps


def _dropout(y, p):
    """nn.Dropout(p) in training mode (reference src/densenet.py:119,126): torch's fused kernel — one
    read, one write, a byte mask for backward.  Stays torch, as in the per-layer path."""
    if p <= 0.0:
        return y, None
    out, keep = torch.native_dropout(y, p, True)
    return out, (keep, 1.0 / (1.0 - p))


def _dropout_bwd(dy, keep):
    if keep is None:
        return dy
    return torch.ops.aten.native_dropout_backward(dy, keep[0], keep[1])


class _DenseBlockFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x0, block, *params):
        lib = _L.load()
        x0 = F_._nhwc(x0)
        n, c0, h, w = x0.shape
        m = n * h * w
        layers = list(block.dense_block.children())
        growth = [_split(ly.conv_block)[2].weight.shape[0] for ly in layers]
        ct = c0 + sum(growth)
        dev = x0.device
        buf = torch.empty((n, ct, h, w), dtype=x0.dtype, device=dev, memory_format=_CL)
        _copy_channels(x0, c0, 0, buf, ct, 0, m, c0)
        mean_t = torch.empty(ct, dtype=torch.float32, device=dev)
        rstd_t = torch.empty(ct, dtype=torch.float32, device=dev)
        saved = []
        cl = c0
        for li, layer in enumerate(layers):
            bn1, p1, conv1 = _split(layer.bottleneck)
            bn2, p2, conv2 = _split(layer.conv_block)
            if bn1.num_features != cl:
                raise RuntimeError("dense block: BatchNorm width does not match the channel prefix")
            g1, b1 = bn1.weight.detach(), bn1.bias.detach()
            if li == 0:
                # the block input is a dense tensor: its BatchNorm is the ordinary one and leaves the
                # statistics of channels [0, C0) in the shared table
                c1 = _Ctx()
                y1 = F_._BatchNormFn.forward(c1, x0, None, g1, b1, bn1.running_mean, bn1.running_var, True,
                                             bn1.momentum, bn1.eps, True, None, bn1.num_batches_tracked)
                _, _, mask1, _, mean0, rstd0 = c1.saved_tensors
                mean_t[:c0].copy_(mean0)
                rstd_t[:c0].copy_(rstd0)
            else:
                mask1 = torch.empty(m * cl // 8, dtype=torch.uint8, device=dev)
                y1 = torch.empty((n, cl, h, w), dtype=x0.dtype, device=dev, memory_format=_CL)
                d = _L.BnDesc(m, cl, F_._dt(buf), _L.CNB_BN_RELU, float(bn1.eps), float(bn1.momentum),
                              bn1.num_batches_tracked.data_ptr())
                ws = F_._workspace(lib.cnb_bn_workspace_bytes(C.byref(d)), dev)
                _L.check(lib.cnb_bn_fwd_train_prefix(C.byref(d), _p(buf), ct, _p(mean_t), _p(rstd_t), _p(g1),
                                                     _p(b1), _p(bn1.running_mean), _p(bn1.running_var), _p(y1),
                                                     _p(mask1), _p(ws), ws.numel(), _stream()),
                         "cnb_bn_fwd_train_prefix")
            y1d, keep1 = _dropout(y1, p1 if block.training else 0.0)
            cc1 = _Ctx()
            wp, wpt = conv1._packed(y1d.dtype)
            t, part = F_._Conv2dFn.forward(cc1, y1d, conv1.weight, conv1.bias, wp, F_._pair(conv1.stride),
                                           F_._pair(conv1.padding), F_._pair(conv1.dilation), 1, conv1.algo,
                                           True, wpt, False)
            cb2 = _Ctx()
            y2 = F_._BatchNormFn.forward(cb2, t, None, bn2.weight.detach(), bn2.bias.detach(),
                                         bn2.running_mean, bn2.running_var, True, bn2.momentum, bn2.eps, True,
                                         part if part.shape[0] > 0 else None, bn2.num_batches_tracked)
            y2d, keep2 = _dropout(y2, p2 if block.training else 0.0)
            cc2 = _Ctx()
            wp, wpt = conv2._packed(y2d.dtype)
            o, part2 = F_._Conv2dFn.forward(cc2, y2d, conv2.weight, conv2.bias, wp, F_._pair(conv2.stride),
                                            F_._pair(conv2.padding), F_._pair(conv2.dilation), 1, conv2.algo,
                                            True, wpt, False)
            g = growth[li]
            if part2.shape[0] == 0:
                raise RuntimeError("dense block: the 3x3 convolution did not emit BatchNorm partials "
                                   "(not on the tensor-core path)")
            _copy_channels(o, g, 0, buf, ct, cl, m, g)
            d2 = _L.BnDesc(m, g, F_._dt(buf), 0, float(bn1.eps), 0.1, None)
            ws = F_._workspace(lib.cnb_bn_workspace_bytes(C.byref(d2)), dev)
            _L.check(lib.cnb_bn_stats_from_partials(C.byref(d2), _p(part2), part2.shape[0],
                                                    C.c_void_p(mean_t.data_ptr() + 4 * cl),
                                                    C.c_void_p(rstd_t.data_ptr() + 4 * cl), _p(ws), ws.numel(),
                                                    _stream()), "cnb_bn_stats_from_partials")
            saved.append((cl, g, mask1, keep1, cc1, cb2, keep2, cc2, bn1.eps))
            cl += g
        ctx.saved = saved
        ctx.buf, ctx.mean_t, ctx.rstd_t = buf, mean_t, rstd_t
        ctx.block = block
        ctx.dims = (n, c0, h, w, ct)
        return buf

    @staticmethod
    def backward(ctx, d_out):
        lib = _L.load()
        n, c0, h, w, ct = ctx.dims
        m = n * h * w
        buf, mean_t, rstd_t = ctx.buf, ctx.mean_t, ctx.rstd_t
        dev = buf.device
        dbuf = F_._nhwc(d_out)
        if dbuf.dtype != buf.dtype:
            dbuf = dbuf.to(buf.dtype)
        dbuf = dbuf.clone(memory_format=_CL)     # accumulated in place below; the incoming tensor is not ours
        layers = list(ctx.block.dense_block.children())
        grads = {}
        for li in range(len(layers) - 1, -1, -1):
            layer = layers[li]
            cl, g, mask1, keep1, cc1, cb2, keep2, cc2, eps1 = ctx.saved[li]
            bn1, _, conv1 = _split(layer.bottleneck)
            bn2, _, conv2 = _split(layer.conv_block)
            d_o = torch.empty((n, g, h, w), dtype=buf.dtype, device=dev, memory_format=_CL)
            _copy_channels(dbuf, ct, cl, d_o, g, 0, m, g)
            r = F_._Conv2dFn.backward(cc2, d_o)
            d_y2, grads[conv2.weight] = r[0], r[1]
            if conv2.bias is not None:
                grads[conv2.bias] = r[2]
            d_y2 = _dropout_bwd(d_y2, keep2)
            r = F_._BatchNormFn.backward(cb2, d_y2)
            d_t, grads[bn2.weight], grads[bn2.bias] = r[0], r[2], r[3]
            r = F_._Conv2dFn.backward(cc1, d_t)
            d_y1, grads[conv1.weight] = r[0], r[1]
            if conv1.bias is not None:
                grads[conv1.bias] = r[2]
            d_y1 = _dropout_bwd(d_y1, keep1)
            dgam = torch.empty(cl, dtype=torch.float32, device=dev)
            dbet = torch.empty(cl, dtype=torch.float32, device=dev)
            d = _L.BnDesc(m, cl, F_._dt(buf), _L.CNB_BN_RELU, float(eps1), 0.1, None)
            ws = F_._workspace(lib.cnb_bn_workspace_bytes(C.byref(d)), dev)
            _L.check(lib.cnb_bn_bwd_prefix(C.byref(d), _p(buf), ct, None, _p(mask1), _p(F_._nhwc(d_y1)),
                                           _p(mean_t), _p(rstd_t), _p(bn1.weight.detach()), _p(dbuf), ct,
                                           _p(dgam), _p(dbet), _p(ws), ws.numel(), _stream()),
                     "cnb_bn_bwd_prefix")
            grads[bn1.weight], grads[bn1.bias] = dgam, dbet
        dx0 = torch.empty((n, c0, h, w), dtype=buf.dtype, device=dev, memory_format=_CL)
        _copy_channels(dbuf, ct, 0, dx0, c0, 0, m, c0)
        ctx.buf = ctx.saved = None
        return (dx0, None) + tuple(grads.get(p) for p in _block_params(ctx.block))


def dense_block_forward(self, x):
    """DenseBlock.forward (reference src/densenet.py:143-168) on one shared buffer when eligible."""
    if eligible(self, x):
        return _DenseBlockFn.apply(x, self, *_block_params(self))
    return self.dense_block(x)
