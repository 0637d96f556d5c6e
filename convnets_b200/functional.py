"""torch.autograd.Function wrappers over the libconvnets_b200 C ABI.

Each function here is the drop-in for one library op the reference reaches through torch.nn
(SURVEY.md §2.1).  Tensors are logical NCHW / physical NHWC (torch.channels_last), dtype bf16
or fp32; parameters, statistics and weight gradients are fp32.  PyTorch only provides device
memory, streams and the autograd tape; all arithmetic on the feature maps happens in the CUDA
library.  There is no CPU path: a non-CUDA tensor raises.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence

import torch

from . import lib as _L

_CL = torch.channels_last


def _dt(t: torch.Tensor) -> int:
    if t.dtype == torch.float32:
        return _L.CNB_F32
    if t.dtype == torch.bfloat16:
        return _L.CNB_BF16
    raise TypeError(f"convnets_b200 supports float32 and bfloat16 activations, got {t.dtype}")


def _cuda(*ts: Optional[torch.Tensor]) -> None:
    for t in ts:
        if t is not None and not t.is_cuda:
            raise RuntimeError(
                "convnets_b200 ops run only on CUDA tensors (sm_100a); there is no CPU fallback. "
                "Use the oracle/ package for CPU reference results.")


def _p(t: Optional[torch.Tensor]):
    return None if t is None else C.c_void_p(t.data_ptr())


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _nhwc(t: torch.Tensor) -> torch.Tensor:
    if t.dim() != 4:
        raise ValueError(f"expected a 4-D NCHW tensor, got shape {tuple(t.shape)}")
    return t if t.is_contiguous(memory_format=_CL) else t.contiguous(memory_format=_CL)


def _empty_nhwc(n, c, h, w, dtype, device) -> torch.Tensor:
    return torch.empty((n, c, h, w), dtype=dtype, device=device, memory_format=_CL)


_workspaces: dict = {}


def _workspace(nbytes: int, device: torch.device) -> torch.Tensor:
    """Grow-only per-device scratch buffer.  Safe to share between ops because every op that
    uses it is enqueued on the same (current) stream, in order."""
    key = (device.index, torch.cuda.current_stream(device).cuda_stream)
    ws = _workspaces.get(key)
    if ws is None or ws.numel() < nbytes:
        ws = torch.empty(max(int(nbytes), 1 << 20), dtype=torch.uint8, device=device)
        _workspaces[key] = ws
    return ws


def _pair(v):
    return (int(v), int(v)) if isinstance(v, int) else (int(v[0]), int(v[1]))


# ----------------------------------------------------------------------------- layout / dtype
def to_nhwc(x: torch.Tensor, dtype: torch.dtype, c_dst: Optional[int] = None) -> torch.Tensor:
    """Network-input conversion: fp32 NCHW -> `dtype` NHWC (reference: the H2D'd batch enters
    the first conv as fp32 NCHW, src/basemodel.py:185-188).  With `c_dst` > C the result has
    c_dst channels, the extra ones zero (feeds the tensor-core stem)."""
    _cuda(x)
    if x.dim() != 4:
        raise ValueError("to_nhwc expects [N, C, H, W]")
    n, c, h, w = x.shape
    c_dst = c if c_dst is None else int(c_dst)
    if c_dst == c and x.dtype == dtype and x.is_contiguous(memory_format=_CL):
        return x
    if x.dtype != torch.float32 or not x.is_contiguous():
        x = x.float().contiguous()
    y = _empty_nhwc(n, c_dst, h, w, dtype, x.device)
    lib = _L.load()
    _L.check(lib.cnb_nchw_to_nhwc(_p(x), _p(y), n, c, h, w, c_dst, _dt(y), _stream()),
             "cnb_nchw_to_nhwc")
    return y


def normalize_u8(x: torch.Tensor, mean, std, nhwc: bool = False, crop: Optional[int] = None,
                 padding: int = 0, aug: Optional[torch.Tensor] = None,
                 out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """uint8 image batch on the device -> the fp32 NCHW tensor transforms.ToTensor + Normalize
    would have produced on the host (reference src/mngrdata.py:151-153), bit for bit.  `x` is
    [N, C, H, W] or, with nhwc=True, [N, H, W, C] (decoded-image layout).  `aug` (int32 [N, 3] on
    the device: top, left, flip) applies RandomCrop(crop, padding) / RandomHorizontalFlip
    (src/mngrdata.py:146-147) with the caller's random draws; without it the geometry is
    unchanged."""
    _cuda(x)
    if x.dtype != torch.uint8 or x.dim() != 4 or not x.is_contiguous():
        raise ValueError("normalize_u8 expects a contiguous 4-D uint8 tensor")
    if nhwc:
        n, h, w, c = x.shape
    else:
        n, c, h, w = x.shape
    ho = wo = int(crop) if crop else None
    if ho is None:
        ho, wo = h, w
    mean_a = (C.c_float * c)(*[float(v) for v in mean])
    std_a = (C.c_float * c)(*[float(v) for v in std])
    if aug is not None:
        _cuda(aug)
        if aug.dtype != torch.int32 or tuple(aug.shape) != (n, 3) or not aug.is_contiguous():
            raise ValueError("aug must be a contiguous int32 [N, 3] tensor (top, left, flip)")
    if out is None:
        out = torch.empty((n, c, ho, wo), dtype=torch.float32, device=x.device)
    elif out.dtype != torch.float32 or tuple(out.shape) != (n, c, ho, wo) or not out.is_contiguous():
        raise ValueError("out must be a contiguous fp32 [N, C, H_out, W_out] tensor")
    lib = _L.load()
    _L.check(lib.cnb_normalize_u8(_p(x), _p(out), n, c, h, w, 1 if nhwc else 0, ho, wo, int(padding),
                                  mean_a, std_a, _p(aug) if aug is not None else None, _stream()),
             "cnb_normalize_u8")
    return out


def cast(src: torch.Tensor, dtype: torch.dtype) -> torch.Tensor:
    """Element-wise precision conversion preserving strides (dense tensors only)."""
    _cuda(src)
    dst = torch.empty_like(src, dtype=dtype)
    lib = _L.load()
    _L.check(lib.cnb_cast(_p(src), _dt(src), _p(dst), _dt(dst), src.numel(), _stream()), "cnb_cast")
    return dst


def pack_weight(weight: torch.Tensor, dtype: torch.dtype, c_dst: Optional[int] = None) -> torch.Tensor:
    """fp32 master weight [K, C/g, R, S] -> KRSC operand copy in `dtype` (a channels_last tensor
    of the same logical shape, or with the channel dimension zero-padded to `c_dst`)."""
    w = weight.detach()
    if c_dst is not None and c_dst > w.shape[1]:
        wp = torch.empty((w.shape[0], c_dst, w.shape[2], w.shape[3]), dtype=w.dtype, device=w.device,
                         memory_format=_CL).zero_()
        wp[:, :w.shape[1]] = w
        w = wp
    if not w.is_contiguous(memory_format=_CL):
        w = w.contiguous(memory_format=_CL)
    if w.dtype == dtype:
        return w
    return cast(w, dtype)


def pack_filter(weight: torch.Tensor, transposed: bool = True):
    """fp32 master weight [K, C, R, S] of a dense conv (KRSC storage) -> (bf16 KRSC operand copy,
    bf16 [C, R*S, K] copy for data-grad or None) in one pass (cnb_pack_filter)."""
    w = weight.detach()
    _cuda(w)
    if w.dtype != torch.float32:
        raise TypeError("pack_filter expects the fp32 master weight")
    if not w.is_contiguous(memory_format=_CL):
        w = w.contiguous(memory_format=_CL)
    k, c, r, s = w.shape
    wb = torch.empty_like(w, dtype=torch.bfloat16)
    wt = torch.empty((c, r * s, k), dtype=torch.bfloat16, device=w.device) if transposed else None
    _L.check(_L.load().cnb_pack_filter(_p(w), _p(wb), _p(wt), k, r * s, c, _stream()),
             "cnb_pack_filter")
    return wb, wt


# ----------------------------------------------------------------------------- convolution
def _conv_desc(x_shape, w_shape, stride, padding, dilation, groups, dtype_code, flags, algo):
    n, c, h, w = x_shape
    k, cg, r, s = w_shape
    if c != cg * groups:
        raise RuntimeError(
            f"Given groups={groups}, weight of size {list(w_shape)}, expected input{list(x_shape)} "
            f"to have {cg * groups} channels, but got {c} channels instead")
    d = _L.ConvDesc(n, h, w, c, k, r, s, stride[0], stride[1], padding[0], padding[1],
                    dilation[0], dilation[1], groups, dtype_code, flags, algo)
    return d


# Autograd hands a Function one gradient per OUTPUT and, by default, materialises zeros for the ones
# nobody differentiated — here the partial-sum tables the conv functions return as non-differentiable
# second outputs: one `fill` launch per conv per step (~170 in a ResNet-50 step).  With
# CNB_SKIP_ZERO_GRADS=1 the conv functions take `None` instead (set_materialize_grads(False)) and
# return no gradient when their main output was not used at all, as torch's own convolution does.
_SKIP_ZERO_GRADS = __import__("os").environ.get("CNB_SKIP_ZERO_GRADS", "0") == "1"


class _Conv2dFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, wpack, stride, padding, dilation, groups, algo, want_stats,
                wpack_t=None, relu=False, bn_holder=None):
        if _SKIP_ZERO_GRADS:
            ctx.set_materialize_grads(False)
        ctx.bn_holder = bn_holder
        _cuda(x, wpack, bias)
        lib = _L.load()
        x = _nhwc(x)
        if x.dtype != wpack.dtype:
            raise TypeError(f"conv2d: activation dtype {x.dtype} != packed weight dtype {wpack.dtype}")
        flags = (_L.CNB_CONV_BIAS if bias is not None else 0) | (_L.CNB_CONV_RELU if relu else 0)
        d = _conv_desc(x.shape, wpack.shape, stride, padding, dilation, groups, _dt(x), flags, algo)
        P, Q = C.c_int32(), C.c_int32()
        _L.check(lib.cnb_conv2d_out_dims(C.byref(d), C.byref(P), C.byref(Q)), "cnb_conv2d_out_dims")
        y = _empty_nhwc(x.shape[0], wpack.shape[0], P.value, Q.value, x.dtype, x.device)
        nws = lib.cnb_conv2d_workspace_bytes(C.byref(d), _L.CNB_FWD)
        ws = _workspace(nws, x.device)
        b32 = None if bias is None else bias.detach().float().contiguous()
        rows = lib.cnb_conv2d_bn_partial_rows(C.byref(d)) if want_stats else 0
        part = torch.empty((max(rows, 0), 2, wpack.shape[0]), dtype=torch.float32, device=x.device)
        _L.check(lib.cnb_conv2d_fwd(C.byref(d), _p(x), _p(wpack), _p(b32), _p(y),
                                    _p(part) if rows > 0 else None, _p(ws), ws.numel(), _stream()),
                 "cnb_conv2d_fwd")
        ctx.save_for_backward(x, wpack, wpack_t)
        ctx.desc = d
        ctx.has_bias = bias is not None
        ctx.relu = bool(relu)
        ctx.wshape = tuple(weight.shape)
        ctx.mark_non_differentiable(part)
        return y, part

    @staticmethod
    def backward(ctx, dy, _dpart=None):
        if ctx.relu:
            raise NotImplementedError(
                "conv2d(relu=True) is the inference epilogue (BatchNorm folded into the filter, "
                "reference valid_model runs under no_grad, src/basemodel.py:304); no backward")
        if dy is None:       # only with CNB_SKIP_ZERO_GRADS: the output was not used
            return (None,) * 13
        lib = _L.load()
        x, wpack, wpack_t = ctx.saved_tensors
        d = ctx.desc
        dy = _nhwc(dy)
        if dy.dtype != x.dtype:
            dy = dy.to(x.dtype)
        dx = dw = db = None
        if ctx.needs_input_grad[0]:
            dx = torch.empty_like(x, memory_format=_CL)
            nws = lib.cnb_conv2d_workspace_bytes(C.byref(d), _L.CNB_DGRAD)
            ws = _workspace(nws, x.device)
            if not _dgrad_with_bn_sums(ctx, lib, d, dy, wpack, wpack_t, None, None, dx, ws):
                _L.check(lib.cnb_conv2d_dgrad(C.byref(d), _p(dy), _p(wpack), _p(wpack_t), _p(dx), _p(ws),
                                              ws.numel(), _stream()), "cnb_conv2d_dgrad")
        if ctx.needs_input_grad[1] or (ctx.has_bias and ctx.needs_input_grad[2]):
            dw = torch.empty(tuple(wpack.shape), dtype=torch.float32, device=x.device,
                             memory_format=_CL)
            if ctx.has_bias:
                db = torch.empty(ctx.wshape[0], dtype=torch.float32, device=x.device)
            nws = lib.cnb_conv2d_workspace_bytes(C.byref(d), _L.CNB_WGRAD)
            ws = _workspace(nws, x.device)
            _L.check(lib.cnb_conv2d_wgrad(C.byref(d), _p(x), _p(dy), _p(dw), _p(db), _p(ws),
                                          ws.numel(), _stream()), "cnb_conv2d_wgrad")
            if tuple(dw.shape) != ctx.wshape:  # zero-padded input channels (RGB stem)
                dw = dw[:, :ctx.wshape[1]].contiguous(memory_format=_CL)
        return dx, dw, db, None, None, None, None, None, None, None, None, None, None


# BatchNorm backward sums from the data-grad epilogue (cnb_conv2d_dgrad_bn): a conv whose input x is
# the output of a BatchNorm2d + ReLU, AND is that output's only consumer, computes dx — which is that
# BatchNorm's dy — and can form sum dy', sum dy' * (x_bn - mean) from the tile it just stored.  The
# BatchNorm layer hands a `holder` (its input, bit mask and batch mean) forward on its output tensor;
# the block forwards of convert.py mark tensors they know to be single-use (`_cnb_sole`); the conv
# leaves the partial table in the holder, keyed by dx's address, and BatchNorm backward skips its
# reduction pass when the gradient that arrives is that very tensor.  CNB_BN_BWD_SUMS=0 turns it off.
_BN_BWD_SUMS = __import__("os").environ.get("CNB_BN_BWD_SUMS", "1") == "1"


def _dgrad_with_bn_sums(ctx, lib, d, dy, wpack, wpack_t, addend, amask, dx, ws) -> bool:
    h = getattr(ctx, "bn_holder", None)
    if not (_BN_BWD_SUMS and h and "x" in h and d.stride_h == 1 and d.stride_w == 1
            and d.algo != _L.CNB_ALGO_SIMT and dx.dtype == torch.bfloat16):
        return False
    bx = h["x"]
    if tuple(bx.shape) != tuple(dx.shape) or bx.dtype != dx.dtype or not bx.is_contiguous(memory_format=_CL):
        return False
    rows = lib.cnb_conv2d_dgrad_partial_rows(C.byref(d))
    if rows <= 0:
        return False
    part = torch.empty((rows, 2, dx.shape[1]), dtype=torch.float32, device=dx.device)
    _L.check(lib.cnb_conv2d_dgrad_bn(C.byref(d), _p(dy), _p(wpack), _p(wpack_t), _p(addend), _p(amask),
                                     _p(bx), _p(h["mask"]), _p(h["mean"]), _p(part), _p(dx), _p(ws),
                                     ws.numel(), _stream()), "cnb_conv2d_dgrad_bn")
    h["bwd"] = (part, dx.data_ptr(), dx)
    return True


def conv2d(x, weight, bias=None, stride=1, padding=0, dilation=1, groups=1, wpack=None,
           algo=_L.CNB_ALGO_AUTO, bn_stats=False, wpack_t=None, relu=False, bn_holder=None):
    """Drop-in for F.conv2d as reached from nn.Conv2d (reference src/basemodel.py:974,992).
    `weight` is the fp32 master parameter; `wpack` an optional pre-packed operand copy.
    With `bn_stats` returns (y, partials): the per-tile BatchNorm partial sums emitted by the
    forward epilogue (None when this conv runs on a path that cannot produce them).  `relu`: fused
    ReLU epilogue, inference only (the BatchNorm-folded eval path, layers.Conv2d)."""
    if wpack is None:
        wpack = pack_weight(weight, x.dtype)
    y, part = _Conv2dFn.apply(x, weight, bias, wpack, _pair(stride), _pair(padding),
                              _pair(dilation), int(groups), int(algo), bool(bn_stats), wpack_t,
                              bool(relu), bn_holder)
    if bn_stats:
        return y, (part if part.shape[0] > 0 else None)
    return y


class _Conv2dSkipFn(torch.autograd.Function):
    """conv2d that also hands its input on as a second output (the identity shortcut of a residual
    block, reference src/resnet.py:139-156,176-196).  Autograd then calls ONE backward with both
    gradients, and the data-grad epilogue adds the skip gradient (cnb_conv2d_dgrad_add) instead of
    a separate element-wise add over the block input.  Used by convert._residual_forward for
    identity-shortcut blocks (CNB_FUSED_SKIP_GRAD=0 turns it off); everything but the data-grad is
    _Conv2dFn."""

    @staticmethod
    def forward(ctx, x, weight, bias, wpack, stride, padding, dilation, groups, algo, want_stats,
                wpack_t=None, bn_holder=None):
        y, part = _Conv2dFn.forward(ctx, x, weight, bias, wpack, stride, padding, dilation, groups,
                                    algo, want_stats, wpack_t, False, bn_holder)
        return y, part, x.view_as(x)

    @staticmethod
    def backward(ctx, dy, _dpart=None, dskip=None):
        # a lazy skip gradient (see _LAZY_DRES) is taken off the table whatever happens next: the
        # entry keeps dy alive, so no other tensor can sit at its address while it is registered
        lazy = _LAZY_DRES.pop(dskip.data_ptr(), None) if dskip is not None else None
        if dy is None and dskip is not None:
            # only with CNB_SKIP_ZERO_GRADS: the conv output was not used, the identity path was
            if not ctx.needs_input_grad[0]:
                return (None,) * 12
            dskip = _nhwc(dskip)
            if lazy is not None:
                dskip = mask_apply(dskip, lazy[0])
            return (dskip,) + (None,) * 11
        if dskip is None or not ctx.needs_input_grad[0]:
            return _Conv2dFn.backward(ctx, dy, _dpart)
        lib = _L.load()
        x, wpack, wpack_t = ctx.saved_tensors
        d = ctx.desc
        dy = _nhwc(dy)
        if dy.dtype != x.dtype:
            dy = dy.to(x.dtype)
        amask = None
        if lazy is not None:
            if lazy[1] is not dskip and (lazy[1].data_ptr() != dskip.data_ptr() or lazy[1].shape != dskip.shape):
                raise RuntimeError("lazy skip gradient: the registered tensor is not the one that arrived")
            amask = lazy[0]
        dskip = _nhwc(dskip)
        if dskip.dtype != x.dtype:
            dskip = dskip.to(x.dtype)
        dx = torch.empty_like(x, memory_format=_CL)
        nws = lib.cnb_conv2d_workspace_bytes(C.byref(d), _L.CNB_DGRAD)
        ws = _workspace(nws, x.device)
        fused = (d.stride_h == 1 and d.stride_w == 1 and d.algo != _L.CNB_ALGO_SIMT
                 and lib.cnb_conv2d_uses_tc(C.byref(d), _L.CNB_DGRAD) == 1)
        if amask is not None and not (fused and d.C % 32 == 0):
            # cannot mask in the epilogue: materialise the masked gradient after all
            dskip = mask_apply(dskip, amask)
            amask = None
        if fused and _dgrad_with_bn_sums(ctx, lib, d, dy, wpack, wpack_t, dskip, amask, dx, ws):
            pass
        elif fused:
            _L.check(lib.cnb_conv2d_dgrad_add(C.byref(d), _p(dy), _p(wpack), _p(wpack_t), _p(dskip),
                                              _p(amask), _p(dx), _p(ws), ws.numel(), _stream()),
                     "cnb_conv2d_dgrad_add")
        else:
            _L.check(lib.cnb_conv2d_dgrad(C.byref(d), _p(dy), _p(wpack), _p(wpack_t), _p(dx), _p(ws),
                                          ws.numel(), _stream()), "cnb_conv2d_dgrad")
            dx.add_(dskip)
        dw = db = None
        if ctx.needs_input_grad[1] or (ctx.has_bias and ctx.needs_input_grad[2]):
            dw = torch.empty(tuple(wpack.shape), dtype=torch.float32, device=x.device,
                             memory_format=_CL)
            if ctx.has_bias:
                db = torch.empty(ctx.wshape[0], dtype=torch.float32, device=x.device)
            nws = lib.cnb_conv2d_workspace_bytes(C.byref(d), _L.CNB_WGRAD)
            ws = _workspace(nws, x.device)
            _L.check(lib.cnb_conv2d_wgrad(C.byref(d), _p(x), _p(dy), _p(dw), _p(db), _p(ws),
                                          ws.numel(), _stream()), "cnb_conv2d_wgrad")
            if tuple(dw.shape) != ctx.wshape:
                dw = dw[:, :ctx.wshape[1]].contiguous(memory_format=_CL)
        return dx, dw, db, None, None, None, None, None, None, None, None, None


def mask_apply(t: torch.Tensor, mask: torch.Tensor) -> torch.Tensor:
    """t where bit (m*C + c) of the 1-bit ReLU mask is set, else 0 (torch ops; only the fallback of
    the lazy skip gradient uses it)."""
    n, c, h, w = t.shape
    bits = mask.view(-1, 1).bitwise_right_shift(torch.arange(8, device=mask.device, dtype=torch.uint8)) & 1
    keep = bits.view(n, h, w, c).permute(0, 3, 1, 2).to(t.dtype)
    return (t * keep).contiguous(memory_format=_CL)


def conv2d_with_skip(x, weight, bias=None, stride=1, padding=0, dilation=1, groups=1, wpack=None,
                     algo=_L.CNB_ALGO_AUTO, bn_stats=False, wpack_t=None, bn_holder=None):
    """conv2d(x, ...) plus an alias of x for the identity shortcut: returns (y, partials-or-None,
    x_skip).  See _Conv2dSkipFn."""
    if wpack is None:
        wpack = pack_weight(weight, x.dtype)
    y, part, xs = _Conv2dSkipFn.apply(x, weight, bias, wpack, _pair(stride), _pair(padding),
                                      _pair(dilation), int(groups), int(algo), bool(bn_stats), wpack_t,
                                      bn_holder)
    return y, (part if (bn_stats and part.shape[0] > 0) else None), xs


class _InputConvFn(torch.autograd.Function):
    """First layer of a network on the tensor-core path: explicit im2col of the fp32 NCHW batch
    (cnb_im2col_nchw) followed by a [N*P*Q, Kd] x [Kd, K] GEMM through cnb_conv2d_fwd with a 1x1
    descriptor.  The im2col matrix is kept for the weight gradient; the input gets no gradient."""

    @staticmethod
    def forward(ctx, x, weight, bias, stride, padding, dilation, dtype, want_stats, relu=False):
        if _SKIP_ZERO_GRADS:
            ctx.set_materialize_grads(False)
        _cuda(x, weight, bias)
        lib = _L.load()
        if x.dtype != torch.float32 or not x.is_contiguous():
            x = x.float().contiguous()
        n, c, h, w = x.shape
        k, cw, r, s = weight.shape
        if cw != c:
            raise RuntimeError(f"expected input with {cw} channels, got {c}")
        code = _L.CNB_BF16 if dtype == torch.bfloat16 else _L.CNB_F32
        d0 = _L.ConvDesc(n, h, w, c, k, r, s, stride[0], stride[1], padding[0], padding[1],
                         dilation[0], dilation[1], 1, code, 0, _L.CNB_ALGO_AUTO)
        P, Q = C.c_int32(), C.c_int32()
        _L.check(lib.cnb_conv2d_out_dims(C.byref(d0), C.byref(P), C.byref(Q)), "cnb_conv2d_out_dims")
        P, Q = P.value, Q.value
        rsc = r * s * c
        kd = (rsc + 7) // 8 * 8
        col = _empty_nhwc(n, kd, P, Q, dtype, x.device)
        _L.check(lib.cnb_im2col_nchw(C.byref(d0), _p(x), _p(col), kd, _stream()), "cnb_im2col_nchw")
        # weight [K, C, R, S] (any layout) -> [K, (r, s, c)] zero-padded to kd columns
        wcol = torch.zeros((k, kd), dtype=torch.float32, device=x.device)
        wcol[:, :rsc] = weight.detach().permute(0, 2, 3, 1).reshape(k, rsc)
        wcol = wcol.to(dtype).view(k, kd, 1, 1)
        flags = (_L.CNB_CONV_BIAS if bias is not None else 0) | (_L.CNB_CONV_RELU if relu else 0)
        d1 = _L.ConvDesc(n, P, Q, kd, k, 1, 1, 1, 1, 0, 0, 1, 1, 1, code, flags, _L.CNB_ALGO_AUTO)
        y = _empty_nhwc(n, k, P, Q, dtype, x.device)
        ws = _workspace(lib.cnb_conv2d_workspace_bytes(C.byref(d1), _L.CNB_FWD), x.device)
        b32 = None if bias is None else bias.detach().float().contiguous()
        rows = lib.cnb_conv2d_bn_partial_rows(C.byref(d1)) if want_stats else 0
        part = torch.empty((max(rows, 0), 2, k), dtype=torch.float32, device=x.device)
        _L.check(lib.cnb_conv2d_fwd(C.byref(d1), _p(col), _p(wcol), _p(b32), _p(y),
                                    _p(part) if rows > 0 else None, _p(ws), ws.numel(), _stream()),
                 "cnb_conv2d_fwd(input)")
        ctx.save_for_backward(col)
        ctx.desc = d1
        ctx.geom = (k, c, r, s, rsc, kd)
        ctx.has_bias = bias is not None
        ctx.relu = bool(relu)
        ctx.mark_non_differentiable(part)
        return y, part

    @staticmethod
    def backward(ctx, dy, _dpart=None):
        if ctx.relu:
            raise NotImplementedError("conv2d_input(relu=True) is the inference epilogue; no backward")
        if dy is None:
            return (None,) * 9
        lib = _L.load()
        (col,) = ctx.saved_tensors
        d = ctx.desc
        k, c, r, s, rsc, kd = ctx.geom
        dy = _nhwc(dy)
        if dy.dtype != col.dtype:
            dy = dy.to(col.dtype)
        dwc = torch.empty((k, kd), dtype=torch.float32, device=col.device)
        db = torch.empty(k, dtype=torch.float32, device=col.device) if ctx.has_bias else None
        ws = _workspace(lib.cnb_conv2d_workspace_bytes(C.byref(d), _L.CNB_WGRAD), col.device)
        _L.check(lib.cnb_conv2d_wgrad(C.byref(d), _p(col), _p(dy), _p(dwc), _p(db), _p(ws), ws.numel(),
                                      _stream()), "cnb_conv2d_wgrad(input)")
        dw = dwc[:, :rsc].view(k, r, s, c).permute(0, 3, 1, 2)
        return None, dw, db, None, None, None, None, None, None


def conv2d_input(x, weight, bias=None, stride=1, padding=0, dilation=1, dtype=torch.bfloat16,
                 bn_stats=False, relu=False):
    """The network's first convolution applied to the fp32 NCHW batch (reference
    src/basemodel.py:185-188 feeds `self(inputs)` with it); see _InputConvFn."""
    y, part = _InputConvFn.apply(x, weight, bias, _pair(stride), _pair(padding), _pair(dilation),
                                 dtype, bool(bn_stats), bool(relu))
    if bn_stats:
        return y, (part if part.shape[0] > 0 else None)
    return y


def rowim2col(x, kernel_w, stride_w, pad_w, dil_w, dtype):
    """fp32 NCHW batch -> [N, S*C rounded up, H, Q] in `dtype`, channels_last: the S horizontal taps of
    the stem filter unrolled into channels (cnb_rowim2col_nchw).  No gradient: x is the input batch."""
    _cuda(x)
    lib = _L.load()
    if x.dtype != torch.float32 or not x.is_contiguous():
        x = x.float().contiguous()
    n, c, h, w = x.shape
    v = 8 if dtype == torch.bfloat16 else 4
    cd = (kernel_w * c + v - 1) // v * v
    q = (w + 2 * pad_w - dil_w * (kernel_w - 1) - 1) // stride_w + 1
    code = _L.CNB_BF16 if dtype == torch.bfloat16 else _L.CNB_F32
    d = _L.ConvDesc(n, h, w, c, 8, 1, kernel_w, 1, stride_w, 0, pad_w, 1, dil_w, 1, code, 0,
                    _L.CNB_ALGO_AUTO)
    out = _empty_nhwc(n, cd, h, q, dtype, x.device)
    _L.check(lib.cnb_rowim2col_nchw(C.byref(d), _p(x), _p(out), cd, _stream()), "cnb_rowim2col_nchw")
    return out


def conv2d_input_rows(x, weight, bias=None, stride=1, padding=0, dilation=1, dtype=torch.bfloat16,
                      bn_stats=False, relu=False):
    """The network's first convolution as an R x 1 convolution over the row-im2col of the batch
    (see rowim2col): same result as conv2d_input, a third of the bytes.  The filter is rearranged by
    differentiable torch ops ([K, C, R, S] -> [K, S*C (+pad), R, 1]), so the weight gradient flows
    back through them.  The default stem path (layers: CNB_STEM_ROWS=0 selects conv2d_input)."""
    stride, padding, dilation = _pair(stride), _pair(padding), _pair(dilation)
    k, c, r, s_ = weight.shape
    x2 = rowim2col(x, s_, stride[1], padding[1], dilation[1], dtype)
    cd = x2.shape[1]
    w2 = weight.permute(0, 2, 3, 1).reshape(k, r, s_ * c)
    w2 = torch.nn.functional.pad(w2, (0, cd - s_ * c)).permute(0, 2, 1).unsqueeze(-1)
    w2 = w2.contiguous(memory_format=_CL)                      # [K, Cd, R, 1], KRSC storage
    return conv2d(x2, w2, bias, (stride[0], 1), (padding[0], 0), (dilation[0], 1), 1,
                  wpack=pack_weight(w2, dtype), bn_stats=bn_stats, relu=relu)


def conv2d_uses_tc(x_shape, w_shape, stride, padding, dilation, groups, dtype, direction) -> bool:
    """Which path a conv would take: True = tcgen05 implicit GEMM, False = CUDA cores."""
    code = _L.CNB_BF16 if dtype == torch.bfloat16 else _L.CNB_F32
    d = _conv_desc(x_shape, w_shape, _pair(stride), _pair(padding), _pair(dilation), groups, code, 0,
                   _L.CNB_ALGO_AUTO)
    r = _L.load().cnb_conv2d_uses_tc(C.byref(d), direction)
    if r < 0:
        _L.check(r, "cnb_conv2d_uses_tc")
    return bool(r)


# ----------------------------------------------------------------------------- batch norm
# Residual tail `output += identity; F.relu(output)` (reference src/resnet.py:191-194): the gradient of
# `identity` is the block-output gradient dy masked by the output ReLU.  When the identity is the
# alias handed out by _Conv2dSkipFn (convert._residual_forward_folded), BatchNorm backward returns dy
# ITSELF for it and registers the 1-bit mask here, keyed by dy's address; _Conv2dSkipFn.backward pops
# the entry and lets the data-grad epilogue mask the addend (cnb_conv2d_dgrad_add addend_mask).
# CNB_LAZY_SKIP_GRAD=0 turns it off (BatchNorm backward writes the masked tensor, as before).
_LAZY_DRES: dict = {}
_LAZY_SKIP = __import__("os").environ.get("CNB_LAZY_SKIP_GRAD", "1") == "1"


class _BatchNormFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, residual, gamma, beta, running_mean, running_var, training, momentum, eps,
                relu, partials, nbt=None, lazy_dres=False, holder=None):
        _cuda(x, residual, gamma, beta)
        lib = _L.load()
        x = _nhwc(x)
        n, c, h, w = x.shape
        flags = (_L.CNB_BN_RELU if relu else 0) | (_L.CNB_BN_ADD if residual is not None else 0)
        if residual is not None:
            residual = _nhwc(residual)
            if residual.shape != x.shape or residual.dtype != x.dtype:
                raise RuntimeError("batch_norm: residual must match x in shape and dtype")
        if nbt is not None and (not training or not nbt.is_cuda or nbt.dtype != torch.int64):
            raise RuntimeError("batch_norm: num_batches_tracked must be a CUDA int64 scalar (training only)")
        d = _L.BnDesc(n * h * w, c, _dt(x), flags, float(eps), float(momentum),
                      None if nbt is None else nbt.data_ptr())
        y = torch.empty_like(x, memory_format=_CL)
        ws = _workspace(lib.cnb_bn_workspace_bytes(C.byref(d)), x.device)
        g32 = None if gamma is None else gamma.detach()
        b32 = None if beta is None else beta.detach()
        if training:
            if n * h * w <= 1:
                raise ValueError(
                    f"Expected more than 1 value per channel when training, got input size {list(x.shape)}")
            mean = torch.empty(c, dtype=torch.float32, device=x.device)
            rstd = torch.empty(c, dtype=torch.float32, device=x.device)
            nparts = 0
            if partials is not None:
                if tuple(partials.shape[1:]) != (2, c) or partials.dtype != torch.float32:
                    raise RuntimeError("batch_norm: conv partials do not belong to this input")
                nparts = partials.shape[0]
            # ReLU mask: 1 bit per element instead of re-reading y in backward
            mask = None
            if relu and x.dtype == torch.bfloat16 and c % 8 == 0:
                mask = torch.empty(n * h * w * c // 8, dtype=torch.uint8, device=x.device)
            _L.check(lib.cnb_bn_fwd_train(C.byref(d), _p(x), _p(residual), _p(g32), _p(b32),
                                          _p(running_mean), _p(running_var), _p(y), _p(mean),
                                          _p(rstd), _p(partials), nparts, _p(mask), _p(ws),
                                          ws.numel(), _stream()), "cnb_bn_fwd_train")
            ctx.save_for_backward(x, (y if mask is None else None) if relu else None, mask, g32,
                                  mean, rstd)
            if holder is not None and mask is not None:
                # what a sole consumer's data-grad needs to form this layer's backward sums
                holder["x"], holder["mask"], holder["mean"] = x, mask, mean
        else:
            _L.check(lib.cnb_bn_fwd_eval(C.byref(d), _p(x), _p(residual), _p(g32), _p(b32),
                                         _p(running_mean), _p(running_var), _p(y), _p(ws), ws.numel(),
                                         _stream()), "cnb_bn_fwd_eval")
            ctx.save_for_backward()
        ctx.training = training
        ctx.desc = d
        ctx.has_res = residual is not None
        ctx.affine = gamma is not None
        # lazy residual gradient: see _LAZY_DRES
        ctx.holder = holder
        ctx.lazy_dres = bool(lazy_dres and training and residual is not None and relu
                             and x.dtype == torch.bfloat16 and c % 32 == 0)
        return y

    @staticmethod
    def backward(ctx, dy):
        if not ctx.training:
            raise NotImplementedError(
                "backward through inference-mode BatchNorm is not part of the reference hot path "
                "(valid_model runs under no_grad, src/basemodel.py:304)")
        lib = _L.load()
        x, y, mask, gamma, mean, rstd = ctx.saved_tensors
        d = ctx.desc
        dy = _nhwc(dy)
        if dy.dtype != x.dtype:
            dy = dy.to(x.dtype)
        dx = torch.empty_like(x, memory_format=_CL)
        lazy = getattr(ctx, "lazy_dres", False) and mask is not None and ctx.needs_input_grad[1]
        dres = torch.empty_like(x, memory_format=_CL) if (ctx.has_res and ctx.needs_input_grad[1]
                                                           and not lazy) else None
        c = x.shape[1]
        dgamma = torch.empty(c, dtype=torch.float32, device=x.device) if ctx.affine else None
        dbeta = torch.empty(c, dtype=torch.float32, device=x.device) if ctx.affine else None
        ws = _workspace(lib.cnb_bn_workspace_bytes(C.byref(d)), x.device)
        h = getattr(ctx, "holder", None)
        sums = h.pop("bwd", None) if h else None
        if sums is not None and mask is not None and sums[1] == dy.data_ptr() and sums[2].shape == dy.shape:
            # the data-grad that produced dy already formed sum dy', sum dy' * (x - mean) (_dgrad_with_bn_sums)
            part = sums[0]
            _L.check(lib.cnb_bn_bwd_from_partials(C.byref(d), _p(x), _p(mask), _p(dy), _p(gamma), _p(mean),
                                                  _p(rstd), _p(part), part.shape[0], _p(dx), _p(dres),
                                                  _p(dgamma), _p(dbeta), _p(ws), ws.numel(), _stream()),
                     "cnb_bn_bwd_from_partials")
        else:
            _L.check(lib.cnb_bn_bwd(C.byref(d), _p(x), _p(y), _p(mask), _p(dy), _p(gamma), _p(mean),
                                    _p(rstd), _p(dx), _p(dres), _p(dgamma), _p(dbeta), _p(ws), ws.numel(),
                                    _stream()), "cnb_bn_bwd")
        if lazy:
            # d_residual = dy where the output ReLU was active.  Its only consumer is the skip-gradient
            # epilogue of the block's first convolution (_Conv2dSkipFn), which can apply the 1-bit mask
            # itself: hand it dy and the mask instead of writing a tensor the size of the block output.
            if len(_LAZY_DRES) > 256:
                _LAZY_DRES.clear()
            _LAZY_DRES[dy.data_ptr()] = (mask, dy)
            dres = dy
        return dx, dres, dgamma, dbeta, None, None, None, None, None, None, None, None, None, None


def batch_norm(x, gamma, beta, running_mean, running_var, training, momentum=0.1, eps=1e-5,
               relu=False, residual=None, partials=None, num_batches_tracked=None,
               lazy_residual_grad=False):
    """Drop-in for F.batch_norm on NCHW feature maps (+ optional fused residual add / ReLU).
    `partials`: the per-tile sums the producing conv2d(..., bn_stats=True) emitted for x.
    `num_batches_tracked`: the module's int64 counter; the statistics epilogue adds 1 to it.
    In training with a fused ReLU the result carries `_cnb_bn_holder` (see _dgrad_with_bn_sums)."""
    holder = {} if (training and relu and _BN_BWD_SUMS and torch.is_grad_enabled()) else None
    y = _BatchNormFn.apply(x, residual, gamma, beta, running_mean, running_var, bool(training),
                           float(momentum), float(eps), bool(relu),
                           partials if training else None,
                           num_batches_tracked if training else None, bool(lazy_residual_grad), holder)
    if holder:
        y._cnb_bn_holder = holder
    return y


# ----------------------------------------------------------------------------- element-wise
class _ReluFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        _cuda(x)
        lib = _L.load()
        if not (x.is_contiguous() or (x.dim() == 4 and x.is_contiguous(memory_format=_CL))):
            x = x.contiguous()
        y = torch.empty_like(x)
        _L.check(lib.cnb_relu_fwd(_p(x), _p(y), x.numel(), _dt(x), _stream()), "cnb_relu_fwd")
        ctx.save_for_backward(y)
        return y

    @staticmethod
    def backward(ctx, dy):
        lib = _L.load()
        (y,) = ctx.saved_tensors
        dy = _like(dy, y)
        dx = torch.empty_like(y)
        _L.check(lib.cnb_relu_bwd(_p(y), _p(dy), _p(dx), y.numel(), _dt(y), _stream()), "cnb_relu_bwd")
        return dx


def _like(g: torch.Tensor, ref: torch.Tensor) -> torch.Tensor:
    """Bring an incoming gradient to the dtype and exact memory layout of `ref`."""
    if g.dtype != ref.dtype:
        g = g.to(ref.dtype)
    if g.shape != ref.shape:
        raise RuntimeError(f"gradient shape {tuple(g.shape)} != {tuple(ref.shape)}")
    if g.stride() != ref.stride():
        out = torch.empty_like(ref)
        out.copy_(g)
        return out
    return g


def relu(x):
    """Drop-in for nn.ReLU / F.relu (reference src/basemodel.py:1012; src/resnet.py:154,194)."""
    return _ReluFn.apply(x)


class _AddReluFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, a, b):
        _cuda(a, b)
        lib = _L.load()
        a = _nhwc(a)
        b = _like(b, a) if b.shape == a.shape else b
        if a.shape != b.shape or a.dtype != b.dtype:
            raise RuntimeError("add_relu: operands must match in shape and dtype")
        y = torch.empty_like(a)
        _L.check(lib.cnb_add_relu_fwd(_p(a), _p(b), _p(y), a.numel(), _dt(a), _stream()),
                 "cnb_add_relu_fwd")
        ctx.save_for_backward(y)
        return y

    @staticmethod
    def backward(ctx, dy):
        lib = _L.load()
        (y,) = ctx.saved_tensors
        dy = _like(dy, y)
        dx = torch.empty_like(y)
        _L.check(lib.cnb_add_relu_bwd(_p(y), _p(dy), _p(dx), y.numel(), _dt(y), _stream()),
                 "cnb_add_relu_bwd")
        return dx, dx


def add(a, b):
    """a + b on feature maps, no gradient (inference tail of a residual block without ReLU)."""
    _cuda(a, b)
    a = _nhwc(a)
    b = _like(b, a)
    y = torch.empty_like(a)
    _L.check(_L.load().cnb_add(_p(a), _p(b), _p(y), a.numel(), _dt(a), _stream()), "cnb_add")
    return y


def add_relu(a, b):
    """relu(a + b): the `output += identity; F.relu(output)` tail of every residual block
    (reference src/resnet.py:151-154,191-194 and the resnext/se/sk equivalents)."""
    return _AddReluFn.apply(a, b)


# ----------------------------------------------------------------------------- pooling
def _pool_desc(x, kernel_size, stride, padding):
    kh, kw = _pair(kernel_size)
    sh, sw = _pair(stride if stride is not None else kernel_size)
    ph, pw = _pair(padding)
    n, c, h, w = x.shape
    if ph * 2 > kh or pw * 2 > kw:
        raise RuntimeError("pad should be at most half of effective kernel size")
    P = (h + 2 * ph - kh) // sh + 1
    Q = (w + 2 * pw - kw) // sw + 1
    if P <= 0 or Q <= 0:
        raise RuntimeError(f"Given input size: ({c}x{h}x{w}). Calculated output size: ({c}x{P}x{Q}). "
                           "Output size is too small")
    return _L.PoolDesc(n, h, w, c, kh, kw, sh, sw, ph, pw, _dt(x)), P, Q


class _MaxPoolFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, kernel_size, stride, padding, return_indices):
        _cuda(x)
        lib = _L.load()
        x = _nhwc(x)
        d, P, Q = _pool_desc(x, kernel_size, stride, padding)
        n, c = x.shape[0], x.shape[1]
        y = _empty_nhwc(n, c, P, Q, x.dtype, x.device)
        # int32 h*W+w indices only when the caller wants them; backward uses 1-byte window codes
        idx = torch.empty((n, c, P, Q) if return_indices else (0,), dtype=torch.int32, device=x.device,
                          **({"memory_format": _CL} if return_indices else {}))
        kh, kw = _pair(kernel_size)
        code = None if kh * kw > 255 else torch.empty((n, c, P, Q), dtype=torch.uint8,
                                                      device=x.device, memory_format=_CL)
        if code is None and not return_indices:
            idx = torch.empty((n, c, P, Q), dtype=torch.int32, device=x.device, memory_format=_CL)
        _L.check(lib.cnb_maxpool_fwd(C.byref(d), _p(x), _p(y), _p(idx) if idx.numel() else None,
                                     _p(code), _stream()), "cnb_maxpool_fwd")
        ctx.save_for_backward(idx if code is None else code)
        ctx.coded = code is not None
        ctx.desc = d
        ctx.xshape = tuple(x.shape)
        ctx.mark_non_differentiable(idx)
        return y, idx

    @staticmethod
    def backward(ctx, dy, _didx):
        lib = _L.load()
        (sel,) = ctx.saved_tensors
        dy = _nhwc(dy)
        n, c, h, w = ctx.xshape
        dx = _empty_nhwc(n, c, h, w, dy.dtype, dy.device)
        _L.check(lib.cnb_maxpool_bwd(C.byref(ctx.desc), _p(dy), None if ctx.coded else _p(sel),
                                     _p(sel) if ctx.coded else None, _p(dx), _stream()),
                 "cnb_maxpool_bwd")
        return dx, None, None, None, None


def max_pool2d(x, kernel_size, stride=None, padding=0, return_indices=False):
    """Drop-in for F.max_pool2d (reference src/basemodel.py:1017-1024).  Indices are int32
    h*W+w with PyTorch's first-maximum rule."""
    y, idx = _MaxPoolFn.apply(x, kernel_size, stride, padding, bool(return_indices))
    return (y, idx) if return_indices else y


class _AvgPoolFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, kernel_size, stride, padding):
        _cuda(x)
        lib = _L.load()
        x = _nhwc(x)
        d, P, Q = _pool_desc(x, kernel_size, stride, padding)
        y = _empty_nhwc(x.shape[0], x.shape[1], P, Q, x.dtype, x.device)
        _L.check(lib.cnb_avgpool_fwd(C.byref(d), _p(x), _p(y), _stream()), "cnb_avgpool_fwd")
        ctx.desc = d
        ctx.xshape = tuple(x.shape)
        return y

    @staticmethod
    def backward(ctx, dy):
        lib = _L.load()
        dy = _nhwc(dy)
        n, c, h, w = ctx.xshape
        dx = _empty_nhwc(n, c, h, w, dy.dtype, dy.device)
        _L.check(lib.cnb_avgpool_bwd(C.byref(ctx.desc), _p(dy), _p(dx), _stream()), "cnb_avgpool_bwd")
        return dx, None, None, None


def avg_pool2d(x, kernel_size, stride=None, padding=0):
    """Drop-in for F.avg_pool2d with count_include_pad=True, ceil_mode=False
    (reference src/basemodel.py:1026-1033)."""
    return _AvgPoolFn.apply(x, kernel_size, stride, padding)


class _GapFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        _cuda(x)
        lib = _L.load()
        x = _nhwc(x)
        n, c, h, w = x.shape
        y = torch.empty((n, c, 1, 1), dtype=torch.float32, device=x.device)
        _L.check(lib.cnb_gap_fwd(_p(x), _p(y), n, h * w, c, _dt(x), _stream()), "cnb_gap_fwd")
        ctx.xshape = tuple(x.shape)
        ctx.xdtype = x.dtype
        return y

    @staticmethod
    def backward(ctx, dy):
        lib = _L.load()
        n, c, h, w = ctx.xshape
        dy = dy.float().contiguous()
        dx = _empty_nhwc(n, c, h, w, ctx.xdtype, dy.device)
        _L.check(lib.cnb_gap_bwd(_p(dy), _p(dx), n, h * w, c, _dt(dx), _stream()), "cnb_gap_bwd")
        return dx


def global_avg_pool2d(x):
    """Drop-in for nn.AdaptiveAvgPool2d(1) (reference src/basemodel.py:1035-1042); the pooled
    vector is produced in fp32 (the classifier / gate MLPs consume it in fp32)."""
    return _GapFn.apply(x)


# ----------------------------------------------------------------------------- data movement
class _ShuffleFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, groups):
        _cuda(x)
        lib = _L.load()
        x = _nhwc(x)
        n, c, h, w = x.shape
        if c % groups != 0:
            raise RuntimeError(f"channel_shuffle: {c} channels not divisible by {groups} groups")
        y = torch.empty_like(x, memory_format=_CL)
        _L.check(lib.cnb_channel_shuffle(_p(x), _p(y), n * h * w, c, groups, _dt(x), _stream()),
                 "cnb_channel_shuffle")
        ctx.groups = groups
        return y

    @staticmethod
    def backward(ctx, dy):
        lib = _L.load()
        dy = _nhwc(dy)
        n, c, h, w = dy.shape
        dx = torch.empty_like(dy, memory_format=_CL)
        # inverse permutation of a (g, C/g) transpose is the (C/g, g) transpose
        _L.check(lib.cnb_channel_shuffle(_p(dy), _p(dx), n * h * w, c, c // ctx.groups, _dt(dy),
                                         _stream()), "cnb_channel_shuffle(bwd)")
        return dx, None


def channel_shuffle(x, groups: int):
    """Drop-in for ShuffleUnit.channel_shuffle (reference src/shufflenet_v1.py:173-190)."""
    return _ShuffleFn.apply(x, int(groups))


class _CatFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, *xs):
        _cuda(*xs)
        lib = _L.load()
        xs = [_nhwc(t) for t in xs]
        n, _, h, w = xs[0].shape
        for t in xs:
            if t.shape[0] != n or t.shape[2] != h or t.shape[3] != w or t.dtype != xs[0].dtype:
                raise RuntimeError("cat: tensors must match except in the channel dimension")
        ctot = sum(t.shape[1] for t in xs)
        y = _empty_nhwc(n, ctot, h, w, xs[0].dtype, xs[0].device)
        off = 0
        for t in xs:
            c = t.shape[1]
            _L.check(lib.cnb_copy_channels(_p(t), c, 0, _p(y), ctot, off, n * h * w, c, _dt(t),
                                           _stream()), "cnb_copy_channels")
            off += c
        ctx.sizes = [t.shape[1] for t in xs]
        return y

    @staticmethod
    def backward(ctx, dy):
        lib = _L.load()
        dy = _nhwc(dy)
        n, ctot, h, w = dy.shape
        outs = []
        off = 0
        for i, c in enumerate(ctx.sizes):
            if ctx.needs_input_grad[i]:
                g = _empty_nhwc(n, c, h, w, dy.dtype, dy.device)
                _L.check(lib.cnb_copy_channels(_p(dy), ctot, off, _p(g), c, 0, n * h * w, c, _dt(dy),
                                               _stream()), "cnb_copy_channels(bwd)")
                outs.append(g)
            else:
                outs.append(None)
            off += c
        return tuple(outs)


def cat_channels(xs: Sequence[torch.Tensor]):
    """Drop-in for torch.cat(xs, dim=1) on feature maps (reference src/densenet.py:140,
    src/inceptionnet_v1.py:142, src/squeezenet.py:153, src/shufflenet_v1.py:167)."""
    return _CatFn.apply(*xs)


# ----------------------------------------------------------------------------- gating
class _SeScaleFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, s):
        _cuda(x, s)
        lib = _L.load()
        x = _nhwc(x)
        n, c, h, w = x.shape
        s2 = s.reshape(n, c).float().contiguous()
        y = torch.empty_like(x, memory_format=_CL)
        _L.check(lib.cnb_se_scale_fwd(_p(x), _p(s2), _p(y), n, h * w, c, _dt(x), _stream()),
                 "cnb_se_scale_fwd")
        ctx.save_for_backward(x, s2)
        ctx.sshape = tuple(s.shape)
        ctx.sdtype = s.dtype
        return y

    @staticmethod
    def backward(ctx, dy):
        lib = _L.load()
        x, s2 = ctx.saved_tensors
        n, c, h, w = x.shape
        dy = _nhwc(dy)
        if dy.dtype != x.dtype:
            dy = dy.to(x.dtype)
        dx = torch.empty_like(x, memory_format=_CL)
        ds = torch.empty((n, c), dtype=torch.float32, device=x.device)
        _L.check(lib.cnb_se_scale_bwd(_p(x), _p(s2), _p(dy), _p(dx), _p(ds), n, h * w, c, _dt(x),
                                      _stream()), "cnb_se_scale_bwd")
        return dx, ds.reshape(ctx.sshape).to(ctx.sdtype)


def se_scale(x, s):
    """x * s.expand_as(x) with s of shape [N, C, 1, 1] (reference src/se_resnet.py:253)."""
    return _SeScaleFn.apply(x, s)


class _SkFuseGapFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, u0, u1):
        _cuda(u0, u1)
        lib = _L.load()
        u0, u1 = _nhwc(u0), _nhwc(u1)
        n, c, h, w = u0.shape
        g = torch.empty((n, c, 1, 1), dtype=torch.float32, device=u0.device)
        _L.check(lib.cnb_sk_fuse_gap_fwd(_p(u0), _p(u1), _p(g), n, h * w, c, _dt(u0), _stream()),
                 "cnb_sk_fuse_gap_fwd")
        ctx.xshape = tuple(u0.shape)
        ctx.xdtype = u0.dtype
        return g

    @staticmethod
    def backward(ctx, dg):
        lib = _L.load()
        n, c, h, w = ctx.xshape
        dg = dg.float().contiguous()
        du = _empty_nhwc(n, c, h, w, ctx.xdtype, dg.device)
        _L.check(lib.cnb_gap_bwd(_p(dg), _p(du), n, h * w, c, _dt(du), _stream()), "cnb_gap_bwd")
        return du, du


def sk_fuse_gap(u0, u1):
    """global_avgpool(u0 + u1) (reference src/sknet.py:226-231), fp32 [N, C, 1, 1]."""
    return _SkFuseGapFn.apply(u0, u1)


class _SkSelectFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, u0, u1, logits):
        _cuda(u0, u1, logits)
        lib = _L.load()
        u0, u1 = _nhwc(u0), _nhwc(u1)
        n, c, h, w = u0.shape
        lg = logits.reshape(n, 2, c).float().contiguous()
        attn = torch.empty_like(lg)
        y = torch.empty_like(u0, memory_format=_CL)
        _L.check(lib.cnb_sk_select_fwd(_p(u0), _p(u1), _p(lg), _p(attn), _p(y), n, h * w, c, _dt(u0),
                                       _stream()), "cnb_sk_select_fwd")
        ctx.save_for_backward(u0, u1, attn)
        ctx.lshape = tuple(logits.shape)
        ctx.ldtype = logits.dtype
        return y

    @staticmethod
    def backward(ctx, dy):
        lib = _L.load()
        u0, u1, attn = ctx.saved_tensors
        n, c, h, w = u0.shape
        dy = _nhwc(dy)
        if dy.dtype != u0.dtype:
            dy = dy.to(u0.dtype)
        du0 = torch.empty_like(u0, memory_format=_CL)
        du1 = torch.empty_like(u0, memory_format=_CL)
        dl = torch.empty((n, 2, c), dtype=torch.float32, device=u0.device)
        _L.check(lib.cnb_sk_select_bwd(_p(u0), _p(u1), _p(attn), _p(dy), None, _p(du0), _p(du1),
                                       _p(dl), n, h * w, c, _dt(u0), _stream()), "cnb_sk_select_bwd")
        return du0, du1, dl.reshape(ctx.lshape).to(ctx.ldtype)


def sk_select(u0, u1, logits):
    """sum_i u_i * softmax_i(logits) over the two SK paths (reference src/sknet.py:233-240).
    `logits` is [N, 2*C, 1, 1] (path-major, as produced by torch.cat of the attention convs)."""
    return _SkSelectFn.apply(u0, u1, logits)
