"""Drop-in nn.Module layers backed by libconvnets_b200.

Each class subclasses the torch.nn layer the reference constructs in its factories
(reference src/basemodel.py:966-1067), so `isinstance` checks in `init_params`
(src/basemodel.py:92-104), attribute reads in `save_conv_outshape` (:734-763), parameter names
and state_dict keys are all unchanged; only `forward` differs — it calls the CUDA library.
"""
from __future__ import annotations

import os

import torch
import torch.nn as nn

from . import functional as F_
from . import lib as _L


# Packed operand copies of the weights are cached per layer and must be rebuilt whenever the
# parameter changes.  The tensor version counter is NOT enough: torch's fused (multi-tensor) Adam /
# SGD kernels update parameters without bumping it, so a global post-step hook on every optimizer
# advances an epoch that is part of the cache key (load_state_dict / manual in-place edits still
# move the version counter).
_WEIGHT_EPOCH = [0]


def _bump_weight_epoch(opt=None, *_args, **_kwargs) -> None:
    _WEIGHT_EPOCH[0] += 1
    # an optimizer that wrote the operand copies itself (optim.FusedAdam) re-validates them
    restamp = getattr(opt, "restamp", None)
    if restamp is not None:
        restamp()


from torch.optim.optimizer import register_optimizer_step_post_hook as _reg_post_hook  # noqa: E402

_reg_post_hook(_bump_weight_epoch)


def invalidate_packed_weights() -> None:
    """Force every layer to re-pack its weights on the next forward (call after editing
    parameters through a path neither an optimizer step nor the version counter sees)."""
    _bump_weight_epoch()


# One launch re-packs every tensor-core filter of a model (cnb_pack_filters) instead of one KB-sized
# launch per layer (bit-identical to the per-layer cnb_pack_filter:
# tests/test_gpu_fused_paths.py::test_batched_pack_equals_per_layer_pack).  CNB_BATCHED_PACK=0 = per layer.
_BATCHED_PACK = os.environ.get("CNB_BATCHED_PACK", "1") == "1"
# RGB stems through the row im2col (functional.conv2d_input_rows) instead of the full im2col matrix:
# a third of the bytes, but 7 k-steps of 64 with 24 valid channels each instead of 3 — measured
# slower on the ResNet-50 step (profiles/r02a_switches.md: 11 201 vs 11 527 img/s), so it stays an
# option (CNB_STEM_ROWS=1).  tests/test_gpu_fused_paths.py::test_rgb_stem_vs_oracle holds both
# paths to the float64 oracle.
_STEM_ROWS = os.environ.get("CNB_STEM_ROWS", "0") == "1"
# Inference: fold BatchNorm running statistics into the producing convolution (Conv2d._folded).
# CNB_FOLD_BN=0 keeps the separate cnb_bn_fwd_eval pass.
_FOLD_BN = os.environ.get("CNB_FOLD_BN", "1") == "1"


class PackGroup:
    """The dense tensor-core Conv2d layers of one converted model and their persistent bf16 operand
    buffers.  `refresh()` rebuilds every copy with one kernel; the device table of pointers is
    built on first use and again only if a parameter was re-allocated (model.to(), load)."""

    def __init__(self, layers):
        self.layers = list(layers)
        self.ptrs = None
        self.table = None
        self.bufs = None
        self.n_blocks = 0

    def _build(self):
        import ctypes as C
        lib = _L.load()
        items = (_L.PackItem * len(self.layers))()
        self.bufs = []
        first = 0
        for i, m in enumerate(self.layers):
            w = m.weight.data
            if not w.is_contiguous(memory_format=torch.channels_last):
                raise RuntimeError("PackGroup: master weights must be KRSC (convert() stores them so)")
            k, c, r, s_ = w.shape
            wb = torch.empty_like(w, dtype=torch.bfloat16)
            wt = None if m.cast_input else torch.empty((c, r * s_, k), dtype=torch.bfloat16,
                                                       device=w.device)
            self.bufs.append((wb, wt))
            items[i] = _L.PackItem(w.data_ptr(), wb.data_ptr(), None if wt is None else wt.data_ptr(),
                                   k, r * s_, c, first)
            first += int(lib.cnb_pack_filter_blocks(k, r * s_, c))
        self.n_blocks = first
        raw = torch.frombuffer(bytearray(bytes(items)), dtype=torch.uint8)
        self.table = raw.to(self.layers[0].weight.device)
        self.ptrs = tuple(m.weight.data_ptr() for m in self.layers)

    def refresh(self):
        if self.ptrs != tuple(m.weight.data_ptr() for m in self.layers):
            self._build()
        lib = _L.load()
        _L.check(lib.cnb_pack_filters(self.table.data_ptr(), len(self.layers), self.n_blocks,
                                      torch.cuda.current_stream().cuda_stream), "cnb_pack_filters")
        for m, pair in zip(self.layers, self.bufs):
            w = m.weight
            m.__dict__["_cnb_pack"] = ((_WEIGHT_EPOCH[0], w._version, w.data_ptr(), torch.bfloat16), pair)


class Conv2d(nn.Conv2d):
    """nn.Conv2d whose forward / backward run in libconvnets_b200 (tcgen05 implicit GEMM for
    dense bf16 layers, CUDA cores for fp32 / grouped / depthwise)."""

    algo = _L.CNB_ALGO_AUTO
    # set by convert(): arithmetic dtype of the feature maps, and whether this layer is the
    # network entry (its fp32 NCHW input batch is converted to `compute_dtype` NHWC first)
    compute_dtype = None
    cast_input = False
    # set by convert() when the next module is a BatchNorm2d: the forward epilogue then also emits
    # the per-tile sums that BatchNorm's statistics pass would otherwise re-read the output for
    emit_bn_stats = False

    @staticmethod
    def _tag(out):
        if isinstance(out, tuple):
            y, part = out
            if part is not None:
                y._cnb_bn_partials = part
            return y
        return out

    def _packed(self, dtype: torch.dtype):
        """(operand copy of the weight in `dtype`, [C, R*S, K] copy for data-grad or None), cached
        until the parameter changes (optimizer.step() bumps its version counter)."""
        w = self.weight
        key = (_WEIGHT_EPOCH[0], w._version, w.data_ptr(), dtype)
        cache = self.__dict__.get("_cnb_pack")
        if cache is None or cache[0] != key:
            dense_tc = (dtype == torch.bfloat16 and self.groups == 1 and w.dtype == torch.float32
                        and w.shape[0] % 8 == 0 and w.shape[1] % 8 == 0)
            group = self.__dict__.get("_cnb_group") if _BATCHED_PACK else None
            if dense_tc and group is not None and w.is_cuda:
                group.refresh()   # re-packs every member, this layer included
                return self.__dict__["_cnb_pack"][1]
            if dense_tc:
                pair = F_.pack_filter(w, transposed=not self.cast_input)
            else:
                pair = (F_.pack_weight(w, dtype), None)
            cache = (key, pair)
            self.__dict__["_cnb_pack"] = cache
        return cache[1]

    def _folded(self, bn, dtype: torch.dtype):
        """Inference: (filter, bias, packed operand) with the following BatchNorm2d's running
        statistics folded in — w' = w * gamma / sqrt(var + eps), b' = beta + (b - mean) * gamma /
        sqrt(var + eps) — so that conv -> BN -> ReLU is ONE kernel with a bias / ReLU epilogue
        (reference valid_model / evaluate / test, src/basemodel.py:304-346,508-576,601-722).
        Rebuilt when the weights or the running statistics change (every training forward bumps
        num_batches_tracked)."""
        w = self.weight
        nbt = bn.num_batches_tracked
        key = (_WEIGHT_EPOCH[0], w._version, w.data_ptr(), dtype, bn.running_mean._version,
               bn.running_var._version, -1 if nbt is None else nbt._version,
               -1 if bn.weight is None else bn.weight._version,
               -1 if bn.bias is None else bn.bias._version)
        cache = self.__dict__.get("_cnb_fold")
        if cache is None or cache[0] != key:
            with torch.no_grad():
                scale = torch.rsqrt(bn.running_var.float() + bn.eps)
                if bn.weight is not None:
                    scale = scale * bn.weight.float()
                shift = -bn.running_mean.float() * scale
                if bn.bias is not None:
                    shift = shift + bn.bias.float()
                if self.bias is not None:
                    shift = shift + self.bias.float() * scale
                wf = (w.detach().float() * scale.view(-1, 1, 1, 1)).contiguous(memory_format=torch.channels_last)
                dense_tc = (dtype == torch.bfloat16 and self.groups == 1 and w.shape[0] % 8 == 0
                            and w.shape[1] % 8 == 0)
                wpack = F_.pack_filter(wf, transposed=False)[0] if dense_tc else F_.pack_weight(wf, dtype)
            cache = (key, (wf, shift.contiguous(), wpack))
            self.__dict__["_cnb_fold"] = cache
        return cache[1]

    def forward(self, x: torch.Tensor, with_skip: bool = False):
        """`with_skip`: also return an alias of x for the identity shortcut of a residual block —
        (conv(x), x_skip) — so that the skip gradient is folded into this layer's data-grad
        (functional._Conv2dSkipFn)."""
        if self.padding_mode != "zeros" or isinstance(self.padding, str):
            raise NotImplementedError("convnets_b200.Conv2d supports integer zero padding only")
        bn = self.__dict__.get("_cnb_bn")
        if (bn is not None and not self.training and not bn.training and bn.track_running_stats
                and bn.running_mean is not None and not torch.is_grad_enabled() and not with_skip
                and _FOLD_BN):
            return self._forward_folded(x, bn)
        stats = self.emit_bn_stats and self.training
        # x is the output of a BatchNorm + ReLU and this conv its only reader (marked by the block
        # forward): the data-grad epilogue will also form that BatchNorm's backward sums
        holder = getattr(x, "_cnb_bn_holder", None) if getattr(x, "_cnb_sole", False) else None
        if with_skip:
            wpack, wpack_t = self._packed(x.dtype)
            y, part, xs = F_.conv2d_with_skip(x, self.weight, self.bias, self.stride, self.padding,
                                              self.dilation, self.groups, wpack=wpack, algo=self.algo,
                                              bn_stats=stats, wpack_t=wpack_t, bn_holder=holder)
            return self._tag((y, part)), xs
        if self.cast_input and self.compute_dtype is not None and not x.requires_grad:
            # network entry: fp32 NCHW batch -> compute dtype NHWC.  A stem whose channel count
            # cannot feed the tensor-core path (RGB: 16-byte TMA rows need C % 8 == 0) goes through
            # an explicit im2col + GEMM instead of the CUDA-core kernel.
            if self.compute_dtype == torch.bfloat16 and self.groups == 1 and x.shape[1] % 8:
                entry = F_.conv2d_input_rows if _STEM_ROWS else F_.conv2d_input
                return self._tag(entry(x, self.weight, self.bias, self.stride, self.padding,
                                       self.dilation, self.compute_dtype, bn_stats=stats))
            x = F_.to_nhwc(x, self.compute_dtype)
        wpack, wpack_t = self._packed(x.dtype)
        return self._tag(F_.conv2d(x, self.weight, self.bias, self.stride, self.padding,
                                   self.dilation, self.groups, wpack=wpack, algo=self.algo,
                                   bn_stats=stats, wpack_t=wpack_t, bn_holder=holder))


    def _forward_folded(self, x: torch.Tensor, bn) -> torch.Tensor:
        relu = bool(bn.fuse_relu)
        if self.cast_input and self.compute_dtype is not None:
            if self.compute_dtype == torch.bfloat16 and self.groups == 1 and x.shape[1] % 8:
                wf, bf, _ = self._folded(bn, self.compute_dtype)
                entry = F_.conv2d_input_rows if _STEM_ROWS else F_.conv2d_input
                y = entry(x, wf, bf, self.stride, self.padding, self.dilation, self.compute_dtype,
                          relu=relu)
                y._cnb_bn_folded = bn
                return y
            x = F_.to_nhwc(x, self.compute_dtype)
        wf, bf, wpack = self._folded(bn, x.dtype)
        y = F_.conv2d(x, wf, bf, self.stride, self.padding, self.dilation, self.groups, wpack=wpack,
                      algo=self.algo, relu=relu)
        y._cnb_bn_folded = bn
        return y


class BatchNorm2d(nn.BatchNorm2d):
    """nn.BatchNorm2d (per-replica batch statistics, as in the reference) with an optional fused
    ReLU: `fuse_relu` is set by convert() when the next module is nn.ReLU."""

    fuse_relu = False

    def forward(self, x: torch.Tensor, residual: torch.Tensor | None = None,
                relu: bool | None = None, lazy_residual_grad: bool = False) -> torch.Tensor:
        self._check_input_dim(x)
        training = self.training or (self.running_mean is None and self.running_var is None)
        if not training and getattr(x, "_cnb_bn_folded", None) is self:
            # inference: the producing convolution already applied this layer (and its fused ReLU)
            # through its folded filter / bias epilogue (Conv2d._folded); what is left is the
            # residual tail of a block, `output += identity; F.relu(output)` (src/resnet.py:191-194)
            r = self.fuse_relu if relu is None else relu
            if residual is not None:
                return F_.add_relu(x, residual) if r else F_.add(x, residual)
            return F_.relu(x) if (r and not self.fuse_relu) else x
        momentum = 0.0 if self.momentum is None else self.momentum
        nbt = None
        if self.training and self.track_running_stats and self.num_batches_tracked is not None:
            if self.momentum is None or not self.num_batches_tracked.is_cuda:
                self.num_batches_tracked.add_(1)
                if self.momentum is None:   # cumulative average: the factor is read on the host
                    momentum = 1.0 / float(self.num_batches_tracked)
            else:
                nbt = self.num_batches_tracked   # +1 inside the statistics epilogue (no extra launch)
        rm = self.running_mean if (not training or self.track_running_stats) else None
        rv = self.running_var if (not training or self.track_running_stats) else None
        partials = getattr(x, "_cnb_bn_partials", None) if training else None
        return F_.batch_norm(x, self.weight, self.bias, rm, rv, training, momentum, self.eps,
                             relu=self.fuse_relu if relu is None else relu, residual=residual,
                             partials=partials, num_batches_tracked=nbt,
                             lazy_residual_grad=lazy_residual_grad and F_._LAZY_SKIP)


class ReLU(nn.ReLU):
    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return F_.relu(x)


class MaxPool2d(nn.MaxPool2d):
    def forward(self, x: torch.Tensor):
        if self.dilation not in (1, (1, 1)) or self.ceil_mode:
            raise NotImplementedError("convnets_b200.MaxPool2d: dilation=1, ceil_mode=False only")
        return F_.max_pool2d(x, self.kernel_size, self.stride, self.padding, self.return_indices)


class AvgPool2d(nn.AvgPool2d):
    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if self.ceil_mode or not self.count_include_pad or self.divisor_override is not None:
            raise NotImplementedError(
                "convnets_b200.AvgPool2d: count_include_pad=True, ceil_mode=False only (the "
                "reference's configuration, src/basemodel.py:1026-1033)")
        return F_.avg_pool2d(x, self.kernel_size, self.stride, self.padding)


class AdaptiveAvgPool2d(nn.AdaptiveAvgPool2d):
    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if self.output_size not in (1, (1, 1)):
            raise NotImplementedError("convnets_b200.AdaptiveAvgPool2d: output_size=1 only")
        return F_.global_avg_pool2d(x)


class Identity(nn.Identity):
    """Placeholder left where convert() fused an nn.ReLU into the preceding BatchNorm2d."""
